#!/bin/bash
# ncu --set full capture of every kernel family on the shipped build, one launch each, plus the
# launch list of the bench command.  Run under gpurun from the repo root:
#   gpurun --timeout 1500 -- 'REV=$(cat gpurun_rev.txt) bash tools/ncu_capture.sh'
# Writes gpurun_out/r02_ncu_*.ncu-rep, gpurun_out/r02_ncu_*.md, gpurun_out/r02_traffic.json, r02_launches_bench.csv.
set -u
O=gpurun_out
REV=${REV:-unknown}
SO=$(sha256sum llm-mcp_b200/libembed_truncate_norm.so | cut -c1-16)
rm -f $O/r02_traffic.json
cap() {   # name  kernel-regex  skip  title  command...
  local name=$1 kre=$2 skip=$3 title=$4; shift 4
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$kre -s $skip -c 1 -f -o $O/r02_ncu_$name "$@" > $O/r02_ncu_$name.log 2>&1
  python tools/ncu_summary.py $O/r02_ncu_$name.ncu-rep $O/r02_ncu_$name.md "ncu --set full --clock-control none: $title (source $REV, library sha256 $SO)" $O/r02_traffic.json $name
  [ "$name" = c3 ] || rm -f $O/r02_ncu_$name.ncu-rep      # gpurun brings back at most 64 MiB: the summaries travel, one report stays
}
cap c3      rows_reg    1 "C3 1,048,576 x 4096 -> 1024 float32"            python tools/one_shape.py 1048576 4096 1024 f32
cap c2      rows_reg    1 "C2 65,536 x 768 -> 256 float32"                 python tools/one_shape.py 65536 768 256 f32
cap c4      rows_reg    1 "C4 shard 1,048,576 x 3072 -> 256 float32"       python tools/one_shape.py 1048576 3072 256 f32
cap c1      rows_bulk   1 "C1 1,024 x 4096 -> 1024 float64"                python tools/one_shape.py 1024 4096 1024 f64
cap c3f64   rows_bulk   1 "C3 1,048,576 x 4096 -> 1024 float64"            python tools/one_shape.py 1048576 4096 1024 f64
cap c5      rows_ring   1 "C5 ragged 1,048,576 x 4096 float64, k in {64..1024}"  python tools/ragged_once.py
cap c1x64   rows_ring   1 "64 merged C1 requests (65,536 rows of 8 KiB float64) through the gather launch" python tools/ragged_once.py c1x64
cap wide    rows_wide   1 "wide rows 131,072 x 8192 -> 4096 float64 (32 KiB prefixes)" python tools/one_shape.py 131072 8192 4096 f64
cap general rows_general 1 "off-grid float32 1,000,000 x 771 -> 257 (general kernel)" python tools/one_shape.py 1000000 771 257 f32
cap ringodd rows_ring   1 "off-grid float64 1,000,000 x 771 -> 257 (page ring, aligned span staged)" python tools/one_shape.py 1000000 771 257 f64
cap jsonparse  parse_vectors  0 "JSON packer parse_vectors, C1 response (1,024 vectors x 1024 kept numbers)"  python tools/json_renorm_bench.py
cap jsonformat format_vectors 0 "JSON packer format_vectors, C1 response" python tools/json_renorm_bench.py
# launch list of the bench command (kernel share of the step), after the same command ran clean without ncu
python bench.py --steps 20 --warmup 3 --no-others --no-cpu > $O/r02_bench_for_launchlist.json 2> $O/r02_bench_for_launchlist.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_bench.csv python bench.py --steps 20 --warmup 3 --no-others --no-cpu > $O/r02_launches_bench.out 2>&1
cat $O/r02_traffic.json
