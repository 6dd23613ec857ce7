"""Summarises a .ncu-rep (raw page CSV) into a markdown table under profiles/.
    python tools/ncu_summary.py report.ncu-rep out.md "title" [traffic.json key]
With the last two arguments the DRAM bytes (read + write) of the first launch are stored under
`key` in traffic.json — the figure bench.py republishes as roofline.traffic."""
import csv, json, os, subprocess, sys
rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
pick = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput', 'gpu__dram_throughput',
        'sm__warps_active', 'launch__', 'sm__throughput', 'lts__t_bytes.sum', 'lts__t_sector_hit_rate', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct', 'sm__inst_executed_pipe_fp64', 'sm__pipe_fp64', 'smsp__average_warp', 'l1tex__data_bank_conflicts',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.max']
keys = [h for h in hdr if any(p in h for p in pick)]
with open(out, "w") as f:
    f.write(f"# {title}\n\n")
    for i, r in enumerate(rows[2:]):
        f.write(f"## launch {i}\n\n| metric | value | unit |\n|---|---|---|\n")
        for k in keys:
            v, u = r[hdr.index(k)], units[hdr.index(k)]
            if 'stalled' in k:
                try:
                    if float(v) < 0.05: continue
                except ValueError:
                    pass
            f.write(f"| {k} | {v} | {u} |\n")
        f.write("\n")
r = rows[2]
g = lambda k: r[hdr.index(k)]
print(g('Kernel Name')[:70], g('gpu__time_duration.sum'), units[hdr.index('gpu__time_duration.sum')], g('dram__bytes_read.sum'), g('dram__bytes_write.sum'), units[hdr.index('dram__bytes_write.sum')])

if len(sys.argv) > 5:
    path, key = sys.argv[4], sys.argv[5]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tot = sum(float(g(k)) * scale[units[hdr.index(k)]] for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
    d = json.load(open(path)) if os.path.exists(path) else {}
    d[key] = int(tot)
    d.setdefault("_kernels", {})[key] = g('Kernel Name')
    d.setdefault("_time_us", {})[key] = float(g('gpu__time_duration.sum')) * {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}[units[hdr.index('gpu__time_duration.sum')]]
    json.dump(d, open(path, "w"), indent=1)
