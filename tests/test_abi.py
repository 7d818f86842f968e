"""CPU: the C-ABI library loads and exports every symbol include/shopt_b200.h declares; argument checks that do not
touch a device behave; without a device compute calls fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_cuda


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "shopt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(shopt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported():
    from shopt_b200 import _lib
    L = C.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/shopt_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes signature table out of sync with the header"


def test_version_and_defaults(stage):
    from shopt_b200 import _lib
    L = _lib.lib()
    assert L.shopt_b200_version() >= 100
    p = _lib.Params()
    L.shopt_default_params(C.byref(p))
    assert (p.n_a, p.pca_terms, p.degree, p.iterations, p.max_iter) == (64, 50, 3, 3, 1000)
    assert p.pratio == 0.99 and p.s_lo == 0.075 and p.s_hi == 1.0


def test_host_side_helpers(stage, orc):
    s = np.array([0.3, 0.074, 1.0, 1.0000001, 0.075, 0.0, np.nan])
    assert stage.s_out_of_bounds(s).tolist() == orc.s_out_of_bounds(s, 0.075, 1.0).tolist() == [1, 3, 5]
    rng = np.random.default_rng(0)
    for n in (1, 9, 10, 11, 257):
        err = rng.integers(0, 6, n).astype(float)            # many ties
        assert stage.worst_10_percent(err).tolist() == orc.worst_10_percent(err).tolist()
        assert stage.worst_10_percent(err, True).tolist() == orc.worst_10_percent(err, True).tolist()
    e = np.array([1.0, np.nan, 3.0, np.nan, 2.0] * 4)
    assert stage.worst_10_percent(e).tolist() == orc.worst_10_percent(e).tolist()


def test_bad_arguments(stage):
    from shopt_b200 import _lib
    L = _lib.lib()
    assert L.shopt_polyfit_acc_len(131, 131, 3) == 100 + 10 * 131 * 131
    assert L.shopt_polyfit_acc_len(131, 131, 9) == _lib.E_BADARG
    a = np.zeros(4)
    rc = L.shopt_fit_analytic_batch(a.ctypes.data_as(_lib.dp), 2, 2, 1, 3, a.ctypes.data_as(_lib.dp), 1e-6, 10,
                                    a.ctypes.data_as(_lib.dp), a.ctypes.data_as(_lib.dp), a.ctypes.data_as(_lib.dp), None, None)
    assert rc == _lib.E_BADARG                                # n_a > r


@pytest.mark.skipif(has_cuda(), reason="checks the no-device behaviour")
def test_no_device_fails_loudly(stage):
    from shopt_b200 import _lib
    st = np.zeros((2, 8, 8)); x0 = np.full((2, 3), 0.5)
    with pytest.raises(_lib.ShoptError) as e:
        stage.fit_analytic_batch(st, 8, x0)
    assert e.value.code == _lib.E_NODEVICE


def test_snr_cut_is_the_reference_quantile_rule(stage):
    """outliers_filter (outliers.jl:16-34): host code, runs without a device."""
    from oracle import stage_oracle
    from shopt_b200 import _lib
    rng = np.random.default_rng(7)
    for n, k in [(1, 0.33), (2, 0.33), (10, 0.33), (291, 0.33), (1000, 0.05), (1000, 0.0), (1000, 1.0), (57, 0.5)]:
        snr = rng.normal(30, 10, n).astype(np.float32)
        snr[: n // 5] = np.round(snr[: n // 5])                        # ties
        idx, q1 = stage.snr_cut(snr, k)
        widx, wq1 = stage_oracle.snr_cut(snr, k)
        assert q1 == wq1 and idx.tolist() == widx.tolist()
        assert np.isclose(q1, np.quantile(snr.astype(np.float64), k), rtol=1e-6)
    with pytest.raises(_lib.ShoptError):
        stage.snr_cut(np.array([1.0, np.nan, 2.0], np.float32), 0.33)
