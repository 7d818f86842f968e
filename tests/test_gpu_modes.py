"""GPU parity: the element-wise pixel-grid modes -- "smooth" (shopt.jl:316-348, the shipped shopt.yml default) and
"chisq" (shopt.jl:194-224, chisq.jl:1-11) -- alone and inside the whole stage, against the CPU oracle.
Tolerance: pixel-grid fits within 1e-12 of the stamp peak (element-wise arithmetic), star selection bit-exact."""
import numpy as np
import pytest

from shopt_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def peak_close(a, b, tol):
    scale = np.abs(np.nan_to_num(b)).max(axis=(1, 2), keepdims=True)
    scale[scale == 0] = 1.0
    return np.array_equal(np.isnan(a), np.isnan(b)) and np.nanmax(np.abs(a - b) / scale) <= tol


@pytest.mark.parametrize("r,kw", [(41, dict(frac_masked=0.02)), (131, {}), (64, dict(frac_masked=0.01))])
@pytest.mark.parametrize("unity", [False, True])
def test_smooth_mode(stage, orc, r, kw, unity):
    from oracle import stage_oracle
    cat = synth.make_catalog(6, r, min(r, 32), **kw)
    st = cat["stamps"].numpy()
    got = stage.smooth_mode_batch(st, 3, unity)
    want = stage_oracle.pixel_grid_elementwise(1, st, 3, unity)
    assert peak_close(got, want, 1e-12)


def test_smooth_kernel_other_widths_and_large_stamps(stage, orc):
    rng = np.random.default_rng(11)
    img = rng.normal(size=(2, 40, 37))
    for la in (1, 2, 5):                       # only a = 3 has the separable path; the others take the tap loop
        got = stage.smooth_batch(img, la, False)
        want = np.stack([orc.smooth(img[i], la) for i in range(2)])
        assert np.allclose(got, want, rtol=1e-12, atol=1e-14)
    big = rng.normal(size=(1, 301, 301))       # does not fit shared memory: taps from global memory
    got = stage.smooth_batch(big, 3, False)
    assert np.allclose(got[0], orc.smooth(big[0], 3), rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("unity", [False, True])
@pytest.mark.parametrize("with_x0", [False, True])
def test_chisq_mode(stage, unity, with_x0):
    from oracle import stage_oracle
    rng = np.random.default_rng(4)
    N, r = 5, 33
    st = rng.normal(size=(N, r, r)) + 2.0
    w = rng.uniform(0.5, 2.0, size=(N, r, r))
    st[0, 3, 4] = np.nan
    st[1, 0, 0] = np.inf
    w[2, 5, 5] = 0.0
    w[2, 6, 6] = np.inf
    w[3, 7, 7] = np.nan
    w[4, 1, 2] = -1.0
    x0 = rng.uniform(size=(N, r, r)) if with_x0 else None
    got = stage.chisq_batch(st, w, x0, fill=0.25, unity_sum=unity)
    want = stage_oracle.pixel_grid_elementwise(2, st, 3, unity, w, x0, 0.25)
    assert peak_close(got, want, 1e-13)
    dead = [(0, 3, 4), (1, 0, 0), (2, 5, 5), (2, 6, 6), (3, 7, 7), (4, 1, 2)]
    if not unity:
        for (i, a, b) in dead:
            assert got[i, a, b] == (x0[i, a, b] if with_x0 else 0.25)
        live = np.ones_like(st, bool)
        for d in dead:
            live[d] = False
        assert np.array_equal(got[live], st[live])


@pytest.mark.parametrize("mode,unity", [(_lib.MODE_SMOOTH, False), (_lib.MODE_SMOOTH, True), (_lib.MODE_CHISQ, False)])
def test_stage_in_elementwise_modes(orc, mode, unity):
    from oracle import stage_oracle
    from shopt_b200 import pipeline
    from test_gpu_stage import check_against_oracle
    N, r, n_a = 160, 41, 32
    cat = synth.make_catalog(N, r, n_a, frac_bad_s=0.03)
    a = {k: cat[k].numpy() for k in ("stamps", "u", "v", "x0_model", "x0_learned", "perm")}
    rng = np.random.default_rng(8)
    w = rng.uniform(1e-10, 2e-10, size=a["stamps"].shape)
    w[5, 20, 20] = np.inf
    x0p = rng.uniform(size=a["stamps"].shape) * 1e-6
    p = pipeline.default_params(n_a=n_a, degree=2, iterations=3, training_ratio=0.8, mode=mode, unity_sum=int(unity),
                                chisq_fill=0.0)
    extra = dict(weights=w, x0_pixels=x0p) if mode == _lib.MODE_CHISQ else {}
    got = pipeline.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"], **extra)
    want = stage_oracle.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"],
                                  nthreads=orc.max_threads(), **extra)
    check_against_oracle(got, want, a["x0_model"], a["x0_learned"])
    assert peak_close(got["pgf"], want["pgf"], 1e-12)
    if mode == _lib.MODE_CHISQ:
        assert got["pgf"][5, 20, 20] == x0p[5, 20, 20]


def test_stage_with_median_constraint(orc):
    from oracle import stage_oracle
    from shopt_b200 import pipeline
    from test_gpu_stage import check_against_oracle
    N, r, n_a = 200, 33, 32
    cat = synth.make_catalog(N, r, n_a)
    a = {k: cat[k].numpy() for k in ("stamps", "u", "v", "x0_model", "x0_learned", "perm")}
    p = pipeline.default_params(n_a=n_a, degree=2, iterations=2, training_ratio=0.8, median_constraint=1)
    got = pipeline.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"])
    want = stage_oracle.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"],
                                  nthreads=orc.max_threads())
    check_against_oracle(got, want, a["x0_model"], a["x0_learned"], strict_poly=False)
    # coefficient 1 is the per-pixel median of pixel-grid fits that agree to ~1e-15: compare against the peak
    scale = np.abs(want["poly"]).max(axis=(1, 2), keepdims=True)
    assert np.max(np.abs(got["poly"] - want["poly"]) / scale) <= 1e-6
