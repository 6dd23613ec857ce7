"""-m gpu parity tests: the CUDA path, called through the C-ABI, against the golden vectors
and the CPU oracle (oracle/etn_oracle.c).  Bar (BASELINE.json north_star): output count and
order exact; TRUNCATE bit-exact; TRUNCATE_RENORM within 4 ulp per element of the exactly
rounded result.  The library documents 2 ulp, which is asserted as well.

Reference semantics under test: core/internal/api/handlers.go:2062-2078."""
import threading

import numpy as np
import pytest

from tests.golden_util import kat_cases, synthetic

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from llm_mcp_b200 import embedgpu as eg  # noqa: E402
from tests import gpu_util as gu  # noqa: E402

ULP_BAR = 4       # north_star
ULP_DOC = 2       # include/embed_truncate_norm.h
KATS = kat_cases()


def assert_ulp(oracle, got: np.ndarray, want: np.ndarray, bar=ULP_DOC, what=""):
    assert got.shape == want.shape, (got.shape, want.shape)
    d, i = oracle.max_ulp(got, want)
    assert 0 <= d <= bar, f"{what}: {d} ulp at flat index {i}: got {got.ravel()[i]!r} want {want.ravel()[i]!r}"
    assert bar <= ULP_BAR


# --------------------------------------------------------------------- golden ----
@pytest.mark.parametrize("case", KATS, ids=[c["name"] for c in KATS])
def test_kat_truncate_bit_exact(case):
    x = gu.to_dev(case["in"])
    got = gu.run_dense(x, case["k"], eg.TRUNCATE).cpu().numpy()
    assert got.shape == (x.shape[0], min(case["k"], case["D"]))
    for r, want in enumerate(case["truncate"]):
        assert np.array_equal(gu.bits(got[r]), gu.bits(want))


@pytest.mark.parametrize("case", KATS, ids=[c["name"] for c in KATS])
def test_kat_renorm_within_ulp(oracle, case):
    x = gu.to_dev(case["in"])
    got = gu.run_dense(x, case["k"], eg.TRUNCATE_RENORM).cpu().numpy()
    assert np.isfinite(got).all()
    assert_ulp(oracle, got, case["renorm"], what=case["name"])
    # zeros keep their sign, zero rows are copied
    z = case["renorm"] == 0
    assert np.array_equal(gu.bits(got[z]), gu.bits(case["renorm"][z]))


def test_synthetic_golden(oracle):
    z, manifest = synthetic()
    for m in manifest:
        dt = np.float32 if m["dtype"] == "f32" else np.float64
        x = gu.dev_fill(m["seed"], m["rows"], m["D"], dt)
        assert np.array_equal(x.cpu().numpy(), oracle.fill(m["seed"], m["rows"], m["D"], dt)), "device generator"
        if m["k"] == "ragged":
            kpr = z[m["name"] + "_k"]
            off = oracle.ragged_offsets(kpr, m["D"])
            got = gu.run_ragged(x, kpr, off, eg.TRUNCATE_RENORM).cpu().numpy()
            assert_ulp(oracle, got, z[m["name"]], what=m["name"])
            continue
        got = gu.run_dense(x, m["k"], eg.TRUNCATE_RENORM).cpu().numpy()
        assert_ulp(oracle, got, z[m["name"]], what=m["name"])


# --------------------------------------------------------------- oracle parity ----
SHAPES = [
    # (rows, D, k, dtype, seed)
    (1024, 4096, 1024, np.float64, 1),     # C1 in full
    (65536, 768, 256, np.float32, 2),      # C2 in full
    (4096, 4096, 1024, np.float32, 3),     # C3 rows
    (4096, 3072, 256, np.float32, 4),      # C4 rows
    (2048, 4096, 64, np.float64, 5), (2048, 4096, 128, np.float64, 5), (2048, 4096, 512, np.float64, 5),
    (777, 768, 256, np.float64, 6), (513, 1536, 512, np.float32, 6),
    (300, 4096, 2048, np.float32, 7), (300, 4096, 2048, np.float64, 7),   # 8 KiB / 16 KiB rows
    (64, 8192, 8192, np.float64, 8), (64, 12288, 9000, np.float32, 8),     # beyond the register path
    (1, 4096, 1024, np.float32, 9), (7, 4096, 1024, np.float64, 9), (9, 768, 256, np.float32, 9),
    (1000, 770, 257, np.float32, 10), (1000, 771, 100, np.float64, 10), (500, 33, 32, np.float32, 10),
    (500, 40, 40, np.float64, 11), (500, 40, 39, np.float64, 11), (500, 4096, 1, np.float32, 11),
    (500, 4, 2, np.float64, 12), (500, 100, 96, np.float32, 12), (333, 1000, 36, np.float32, 12),
    (200, 48, 1000, np.float32, 13), (200, 48, 48, np.float64, 13),         # k >= D
    # short rows: 8 / 16 lanes per row
    (4099, 768, 8, np.float32, 14), (4099, 768, 64, np.float32, 14), (4099, 768, 128, np.float32, 14),
    (4099, 4096, 4, np.float64, 14), (4099, 4096, 32, np.float64, 14), (4099, 4096, 64, np.float64, 14),
    (4099, 4096, 24, np.float64, 14), (4099, 772, 12, np.float32, 14),        # 16-byte tiles only
    # several warps per row: 12, 16, 32, 64 KiB prefixes, then beyond
    (257, 3072, 1536, np.float64, 15), (257, 4096, 4096, np.float32, 15), (129, 4096, 4096, np.float64, 15),
    (65, 8192, 8192, np.float64, 15), (65, 16384, 16384, np.float32, 15), (33, 20000, 20000, np.float32, 15),
    (129, 4096, 3000, np.float32, 15), (129, 4096, 1100, np.float64, 15),
    # odd k on 16-byte aligned rows: staged through the ring, element-wise
    (3001, 4096, 257, np.float64, 17), (3001, 4096, 1001, np.float64, 17), (3001, 3072, 1001, np.float32, 17),
    (3001, 4096, 1, np.float64, 17), (3001, 4096, 1021, np.float64, 17), (3001, 4096, 2043, np.float32, 17),
    (3001, 4096, 1023, np.float64, 17), (3001, 4096, 2047, np.float32, 17), (77, 4096, 1025, np.float64, 17),
    # wide-row kernel edges: just over 8 KiB, 96 KiB, the 112 KiB limit, one element past it
    (67, 4096, 2052, np.float32, 16), (35, 14336, 12288, np.float64, 16), (19, 28672, 28672, np.float32, 16),
    (19, 28680, 28676, np.float32, 16), (3, 14340, 14336, np.float64, 16), (1, 4096, 4096, np.float64, 16),
]


@pytest.mark.parametrize("rows,D,k,dt,seed", SHAPES)
def test_dense_against_oracle(oracle, rows, D, k, dt, seed):
    x = gu.dev_fill(seed, rows, D, dt)
    xh = oracle.fill(seed, rows, D, dt)
    tr = gu.run_dense(x, k, eg.TRUNCATE).cpu().numpy()
    assert np.array_equal(gu.bits(tr), gu.bits(oracle.truncate(xh, k)))
    rn = gu.run_dense(x, k, eg.TRUNCATE_RENORM).cpu().numpy()
    assert_ulp(oracle, rn, oracle.renorm_exact(xh, k), what=f"{rows}x{D}->{k} {np.dtype(dt).name}")


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_dispatch_boundaries(oracle, dt):
    """Every size at which the dispatcher changes kernel or tiling (lanes per row, vectors per
    lane, bulk ring, wide rows, the 112 KiB limit), one step below, at, and one step above,
    with k < D and k == D, 16- and 32-byte aligned pitches."""
    es = np.dtype(dt).itemsize
    edges = [16, 32, 256, 512, 1024, 2048, 4096, 6144, 8192, 16384, 57344, 114688]
    rows = 37
    for kb in edges:
        for delta in (-16, 0, 16, es):
            k = (kb + delta) // es
            if k <= 0:
                continue
            for D in (k, k + 16 // es, k + 32 // es + (32 // es) * 3):
                xh = oracle.fill(kb % 97 + 1, rows, D, dt)
                x = gu.to_dev(xh)
                tr = gu.run_dense(x, k, eg.TRUNCATE).cpu().numpy()
                assert np.array_equal(gu.bits(tr), gu.bits(xh[:, :k])), (kb, delta, D)
                rn = gu.run_dense(x, k, eg.TRUNCATE_RENORM).cpu().numpy()
                assert_ulp(oracle, rn, oracle.renorm_exact(xh, k), what=f"k_bytes={kb}{delta:+d} D={D}")


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_random_mantissas_and_scales(oracle, dt):
    """Full-mantissa values over a wide range of row magnitudes, including rows whose squares
    overflow or underflow the element type."""
    rng = np.random.default_rng(21)
    rows, D, k = 4096, 512, 384
    emax = 120 if dt == np.float32 else 1000
    scale = np.exp2(rng.integers(-emax, emax, size=(rows, 1)).astype(np.float64))
    xh = (rng.standard_normal((rows, D)) * scale).astype(dt)
    xh[5, :k] = 0
    xh[6, 3:] = 0
    xh[7, :k:2] = -0.0
    assert np.isfinite(xh).all()
    got = gu.run_dense(gu.to_dev(xh), k, eg.TRUNCATE_RENORM).cpu().numpy()
    assert np.isfinite(got).all()
    assert_ulp(oracle, got, oracle.renorm_exact(xh, k), what="random scales")
    # subnormal-only rows
    tiny = np.finfo(dt).smallest_subnormal
    xs = (rng.integers(-1000, 1000, size=(64, D)) * tiny).astype(dt)
    got = gu.run_dense(gu.to_dev(xs), k, eg.TRUNCATE_RENORM).cpu().numpy()
    assert_ulp(oracle, got, oracle.renorm_exact(xs, k), what="subnormal rows")


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("k", [8, 16, 64, 2048, 6144])
def test_mixed_row_kinds_side_by_side(oracle, dt, k):
    """Neighbouring rows that need different treatment (all zeros, squares overflowing,
    squares underflowing, ordinary) must not disturb each other: short rows share a warp,
    long rows share a block."""
    rng = np.random.default_rng(k)
    rows, D = 515, 6144
    fi = np.finfo(dt)
    kinds = rng.integers(0, 4, size=rows)
    x = rng.standard_normal((rows, D))
    x[kinds == 1] *= float(fi.max) / 16
    x[kinds == 2] *= float(fi.smallest_subnormal) * 64
    xh = x.astype(dt)
    xh[kinds == 0, :k] = 0
    assert np.isfinite(xh).all()
    got = gu.run_dense(gu.to_dev(xh), k, eg.TRUNCATE_RENORM).cpu().numpy()
    assert np.isfinite(got).all()
    assert_ulp(oracle, got, oracle.renorm_exact(xh, k), what=f"mixed kinds k={k}")


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_unaligned_pointers(oracle, dt):
    """Rows that start off a 16-byte boundary must take the general kernel and still match."""
    rows, D, k = 257, 772, 256
    base = gu.dev_fill(14, rows * D + 3, 1, dt).reshape(-1)
    xh = base.cpu().numpy()
    for shift in (1, 2, 3):
        x = base[shift:shift + rows * D].view(rows, D)
        outbuf = torch.empty(rows * k + 3, dtype=x.dtype, device="cuda")
        out = outbuf[shift:shift + rows * k].view(rows, k)
        gu.run_dense(x, k, eg.TRUNCATE_RENORM, out=out)
        want = oracle.renorm_exact(xh[shift:shift + rows * D].reshape(rows, D), k)
        assert_ulp(oracle, out.cpu().numpy(), want, what=f"shift {shift}")
        gu.run_dense(x, k, eg.TRUNCATE, out=out)
        assert np.array_equal(out.cpu().numpy(), xh[shift:shift + rows * D].reshape(rows, D)[:, :k])


def test_in_place_full_rows(oracle):
    """out == in with k >= D, on every kernel family: registers, bulk ring, wide rows, page
    ring (odd k) and the general kernel (odd D)."""
    for dt in (np.float32, np.float64):
        for rows, D in ((1000, 1024), (1000, 96), (300, 4096), (300, 1001), (300, 1003 if dt == np.float32 else 1001), (64, 9000)):
            xh = oracle.fill(15, rows, D, dt)
            x = gu.to_dev(xh)
            gu.run_dense(x, D, eg.TRUNCATE_RENORM, out=x)
            assert_ulp(oracle, x.cpu().numpy(), oracle.renorm_exact(xh, D), what=f"in place {rows}x{D}")


def test_ragged_against_oracle(oracle):
    """C5 in miniature: requests of 1..64 rows, one k per request from {64,...,1024}, plus odd k."""
    rng = np.random.default_rng(5)
    D = 4096
    for dt, ks in ((np.float64, [64, 128, 256, 512, 1024]), (np.float32, [1, 3, 100, 257, 1024, 4096, 5000])):
        kpr = np.concatenate([np.full(rng.integers(1, 65), rng.choice(ks)) for _ in range(200)]).astype(np.int32)
        rows = len(kpr)
        xh = oracle.fill(5, rows, D, dt)
        x = gu.dev_fill(5, rows, D, dt)
        for mode in (eg.TRUNCATE, eg.TRUNCATE_RENORM):
            want, off = oracle.ragged(xh, kpr, mode)
            got = gu.run_ragged(x, kpr, off, mode).cpu().numpy()
            if mode == eg.TRUNCATE:
                assert np.array_equal(gu.bits(got), gu.bits(want))
            else:
                assert_ulp(oracle, got, want, what="ragged")


def test_ragged_ring_wraps_and_fallbacks(oracle):
    """Row-length sequences that make the per-warp 8 KiB page ring wrap, skip pages, run empty
    and hold a single ring-sized row; rows beyond the ring and rows that are not 16-byte
    tileable take the in-kernel fallback."""
    D = 4096
    pats = [[256, 1024] * 40, [64, 1024, 64, 512, 1024, 128, 960, 1024, 2, 1024], [1024] * 33, [2] * 70,
            [2048, 64, 4096, 62, 1024, 3, 1022, 1024], [512, 512, 1024, 256, 256, 256, 1024, 64]]
    for dt in (np.float64, np.float32):
        for pat in pats:
            kpr = np.array((pat * 64)[:1500], dtype=np.int32)
            if dt == np.float32:
                kpr = np.minimum(kpr * 2, 4096)
            rows = len(kpr)
            xh = oracle.fill(31, rows, D, dt)
            x = gu.dev_fill(31, rows, D, dt)
            for mode in (eg.TRUNCATE, eg.TRUNCATE_RENORM):
                want, off = oracle.ragged(xh, kpr, mode)
                got = gu.run_ragged(x, kpr, off, mode).cpu().numpy()
                if mode == eg.TRUNCATE:
                    assert np.array_equal(gu.bits(got), gu.bits(want))
                else:
                    assert_ulp(oracle, got, want, what=f"ragged ring {pat[:4]}")


def test_idempotent_and_unit_norm(oracle):
    x = gu.dev_fill(16, 8192, 768, np.float32)
    y = gu.run_dense(x, 256, eg.TRUNCATE_RENORM)
    n = y.double().pow(2).sum(1).sqrt()
    assert (n - 1).abs().max().item() < 2e-7
    y2 = gu.run_dense(y, 256, eg.TRUNCATE_RENORM)
    d, _ = oracle.max_ulp(y2.cpu().numpy(), y.cpu().numpy())
    assert 0 <= d <= 1
    # exact invariance under power-of-two scaling of the input
    y3 = gu.run_dense(x * 1024.0, 256, eg.TRUNCATE_RENORM)
    assert torch.equal(y3, y)


# -------------------------------------------------------------- boundary errors ----
def test_empty_and_errors():
    L = eg.lib()
    x = torch.zeros((4, 8), dtype=torch.float32, device="cuda")
    o = torch.zeros((4, 8), dtype=torch.float32, device="cuda")
    s = gu.stream_ptr()
    assert L.embed_truncate_norm(0, 8, 4, eg.F32, None, None, s) == 0
    assert L.embed_truncate(0, 8, 4, eg.F64, None, None, s) == 0
    assert L.embed_truncate_norm(4, 8, 4, eg.F32, None, o.data_ptr(), s) == eg.ERR_NULL
    assert L.embed_truncate_norm(4, 8, 4, eg.F32, x.data_ptr(), None, s) == eg.ERR_NULL
    assert L.embed_truncate_norm(-1, 8, 4, eg.F32, x.data_ptr(), o.data_ptr(), s) == eg.ERR_SHAPE
    assert L.embed_truncate_norm(4, 0, 4, eg.F32, x.data_ptr(), o.data_ptr(), s) == eg.ERR_SHAPE
    assert L.embed_truncate_norm(4, 8, 0, eg.F32, x.data_ptr(), o.data_ptr(), s) == eg.ERR_SHAPE
    assert L.embed_truncate_norm(4, 8, 4, 7, x.data_ptr(), o.data_ptr(), s) == eg.ERR_ENUM
    assert L.embed_truncate_norm(4, 8, 4, eg.F64, x.data_ptr() + 4, o.data_ptr(), s) == eg.ERR_ALIGN
    assert L.embed_truncate_norm_ragged(4, 8, None, None, eg.F32, 1, x.data_ptr(), o.data_ptr(), s) == eg.ERR_NULL
    assert L.embed_truncate_norm_ragged(4, 8, x.data_ptr(), x.data_ptr(), eg.F32, 9, x.data_ptr(), o.data_ptr(), s) == eg.ERR_ENUM
    torch.cuda.synchronize()
    assert (o == 0).all()          # nothing ran


def test_concurrent_callers(oracle):
    """Handlers call from arbitrary OS threads (SURVEY.md §8(b) threading)."""
    xh = oracle.fill(17, 2048, 1024, np.float32)
    want = oracle.renorm_exact(xh, 256)
    errs = []

    def worker():
        try:
            torch.cuda.set_device(0)
            with torch.cuda.stream(torch.cuda.Stream()):
                for _ in range(10):
                    got = gu.run_dense(gu.to_dev(xh), 256, eg.TRUNCATE_RENORM)
                    torch.cuda.current_stream().synchronize()
                    d, _ = oracle.max_ulp(got.cpu().numpy(), want)
                    assert 0 <= d <= ULP_DOC
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=worker) for _ in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


# ----------------------------------------------------------- full-size properties ----
def _sample_rows_check(oracle, y, seed, rows, D, k, dt, n_samples=4096):
    rng = np.random.default_rng(seed)
    idx = np.unique(np.concatenate([[0, rows - 1], rng.integers(0, rows, n_samples)]))
    got = y[torch.from_numpy(idx).cuda()].cpu().numpy()
    want = np.empty_like(got)
    # regenerate only the sampled rows on the CPU (counter-based generator)
    runs = np.split(idx, np.where(np.diff(idx) != 1)[0] + 1)
    pos = 0
    for run in runs:
        xh = oracle.fill(seed, len(run), D, dt, row0=int(run[0]), ncols=k)
        want[pos:pos + len(run)] = oracle.renorm_exact(xh, k)
        pos += len(run)
    assert_ulp(oracle, got, want, what=f"sampled rows of {rows}x{D}")


@pytest.mark.parametrize("name,rows,D,k,dt,seed", [
    ("C3_f32", 1048576, 4096, 1024, np.float32, 3),
    ("C3_f64", 524288, 4096, 1024, np.float64, 3),
    ("C4_shard", 2097152, 3072, 256, np.float32, 4),      # two of C4's eight 1M-row shards
])
def test_full_size_properties(oracle, name, rows, D, k, dt, seed):
    """BASELINE.json sizes: unit norms over every row, sampled rows against the oracle,
    truncation equal to a strided copy, idempotence."""
    torch.cuda.empty_cache()
    x = gu.dev_fill(seed, rows, D, dt)
    y = gu.run_dense(x, k, eg.TRUNCATE_RENORM)
    n = torch.linalg.vector_norm(y.double(), dim=1)
    tol = 2e-7 if dt == np.float32 else 4e-16
    assert (n - 1).abs().max().item() < tol
    _sample_rows_check(oracle, y, seed, rows, D, k, dt)
    t = gu.run_dense(x, k, eg.TRUNCATE)
    assert torch.equal(t, x[:, :k])
    del x, t
    y2 = gu.run_dense(y, k, eg.TRUNCATE_RENORM)
    assert (y2 - y).abs().max().item() <= np.finfo(dt).eps
    del y, y2
    torch.cuda.empty_cache()


def test_cuda_graph_capture_and_replay(oracle):
    """Request-sized launches are launch-latency-bound; the device entry points only enqueue
    kernels on the caller's stream, so a caller can capture a burst of them in a CUDA graph."""
    xs = [gu.dev_fill(50 + i, 2048, 1024, np.float32) for i in range(4)]
    ys = [torch.empty((2048, 256), dtype=torch.float32, device="cuda") for _ in range(4)]
    for x, y in zip(xs, ys):                               # warm up outside the capture
        gu.run_dense(x, 256, eg.TRUNCATE_RENORM, out=y)
    torch.cuda.synchronize()
    want = [y.clone() for y in ys]
    for y in ys:
        y.zero_()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for x, y in zip(xs, ys):
            gu.run_dense(x, 256, eg.TRUNCATE_RENORM, out=y)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for y, w in zip(ys, want):
        assert torch.equal(y, w)
    d, _ = oracle.max_ulp(ys[0].cpu().numpy(), oracle.renorm_exact(xs[0].cpu().numpy(), 256))
    assert 0 <= d <= ULP_DOC
