"""shopt_polyfit_mse_noalloc_dev (include/shopt_b200.h): the allocation-free variant of shopt_polyfit_mse_dev whose
per-tile partial sums live in a static __device__ array.  No counterpart in the reference,
so the checker is the existing entry point: same inputs through both, on two calls in a row of
different shapes (the second call finds the first call's partial sums in the shared array), then the first shape again.
Collected last on purpose (file name)."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

UNSUPPORTED = -5


def make_case(seed, N, r, c, degree, n_idx):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    K = (degree + 1) * (degree + 2) // 2
    pgf = torch.rand(N, c, r, generator=g, device="cuda", dtype=torch.float64)
    u = torch.rand(N, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    v = torch.rand(N, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    poly = torch.randn(K, c, r, generator=g, device="cuda", dtype=torch.float64) * 0.3
    idx = None if n_idx is None else torch.randperm(N, generator=g, device="cuda")[:n_idx].contiguous()
    return dict(pgf=pgf, u=u, v=v, poly=poly, idx=idx, r=r, c=c, degree=degree, n=N if idx is None else n_idx)


def run(fn, case, stream=None):
    import torch
    from shopt_b200 import _lib
    out = torch.full((case["n"],), float("nan"), device="cuda", dtype=torch.float64)
    st = C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
    idx = case["idx"].data_ptr() if case["idx"] is not None else None
    _lib.check(fn(case["pgf"].data_ptr(), case["r"], case["c"], case["u"].data_ptr(), case["v"].data_ptr(), idx, case["n"],
                  case["degree"], case["poly"].data_ptr(), out.data_ptr(), st))
    return out


def same(a, b):
    import torch
    # the two entry points launch the same kernels in the same order: the tolerance of the polyfit tests (rtol 1e-9 on
    # mse) is met with room to spare -- the results are in fact bit-identical
    return bool(torch.isfinite(a).all()) and torch.allclose(a, b, rtol=1e-12, atol=0.0) and torch.equal(a, b)


def test_noalloc_mse_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    from shopt_b200 import _lib
    L = _lib.lib()
    first = make_case(11, 700, 41, 37, 3, 500)       # several pixel tiles, an index list
    second = make_case(12, 333, 13, 13, 2, None)     # another shape and degree, identity list: stale partial sums below it
    got1 = run(L.shopt_polyfit_mse_noalloc_dev, first)
    got2 = run(L.shopt_polyfit_mse_noalloc_dev, second)
    got1_again = run(L.shopt_polyfit_mse_noalloc_dev, first)
    torch.cuda.synchronize()
    want1 = run(L.shopt_polyfit_mse_dev, first)
    want2 = run(L.shopt_polyfit_mse_dev, second)
    torch.cuda.synchronize()
    assert same(got1, want1)
    assert same(got2, want2)
    assert same(got1_again, want1)
    assert float(want1.min()) > 0.0                   # a real error, not zeros


def test_noalloc_mse_rejects_sizes_beyond_the_static_scratch():
    import torch
    from shopt_b200 import _lib
    L = _lib.lib()
    r, c, degree = 300, 300, 5
    cap = L.shopt_polyfit_mse_noalloc_max_stars(r, c, degree)
    assert 0 < cap < 10000
    assert L.shopt_polyfit_mse_noalloc_max_stars(131, 131, 3) >= 7000
    dummy = torch.zeros(8, device="cuda", dtype=torch.float64)       # never dereferenced: the size check comes first
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.shopt_polyfit_mse_noalloc_dev(dummy.data_ptr(), r, c, dummy.data_ptr(), dummy.data_ptr(), None, cap + 1, degree,
                                         dummy.data_ptr(), dummy.data_ptr(), st)
    assert rc == UNSUPPORTED
    text = L.shopt_last_error().decode()
    assert "static scratch" in text and str(cap) in text
    torch.cuda.synchronize()
    # n = 0 is a no-op; the list that exactly fits is accepted (small stamp so the test stays small)
    assert L.shopt_polyfit_mse_noalloc_dev(dummy.data_ptr(), r, c, dummy.data_ptr(), dummy.data_ptr(), None, 0, degree,
                                           dummy.data_ptr(), dummy.data_ptr(), st) == 0


def test_noalloc_mse_runs_under_stream_capture():
    """What the entry point is for: a capture in global mode fails on any cudaMalloc / cudaFree inside it."""
    import torch
    from shopt_b200 import _lib
    L = _lib.lib()
    case = make_case(13, 400, 17, 15, 3, 256)
    want = run(L.shopt_polyfit_mse_dev, case)
    warm = run(L.shopt_polyfit_mse_noalloc_dev, case)                 # the symbol address is looked up outside the capture
    torch.cuda.synchronize()
    out = torch.zeros(case["n"], device="cuda", dtype=torch.float64)
    graph = torch.cuda.CUDAGraph()
    idx = case["idx"].data_ptr()
    with torch.cuda.graph(graph, capture_error_mode="global"):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(L.shopt_polyfit_mse_noalloc_dev(case["pgf"].data_ptr(), case["r"], case["c"], case["u"].data_ptr(),
                                                   case["v"].data_ptr(), idx, case["n"], case["degree"],
                                                   case["poly"].data_ptr(), out.data_ptr(), st))
    for _ in range(2):
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert same(out, want) and same(warm, want)
