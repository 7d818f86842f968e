"""The committed bench lines (profiles/r02_bench_*.json, written by bench.py on the GPU box) carry every key of the
driver's contract, and tools/ncu_facts.py still parses the committed ncu export they take their traffic figures from.
CPU-only: nothing here computes on a GPU or loads the CUDA library."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINES = sorted(glob.glob(os.path.join(ROOT, "profiles", "r02_bench_c*_n[0-9].json")))


def test_lines_exist():
    names = {os.path.basename(p) for p in LINES}
    for want in ("r02_bench_c3_n1.json", "r02_bench_c3_n2.json", "r02_bench_c3_n8.json", "r02_bench_c2_n1.json",
                 "r02_bench_c4_n1.json", "r02_bench_c5_n1.json", "r02_bench_c4_n8.json", "r02_bench_c5_n8.json"):
        assert want in names


@pytest.mark.parametrize("path", LINES, ids=[os.path.basename(p) for p in LINES])
def test_bench_line_contract(path):
    d = json.load(open(path))
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in d, k
    assert d["unit"] == "stars/s" and d["dtype"] == "f64" and d["scaling"] == "weak" and d["higher_is_better"] is True
    assert d["vs_baseline"] is None                          # BASELINE.md publishes no number for this metric
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["warmup"] >= 3 and d["gpu_launches"] > 0
    n = int(os.path.basename(path).split("_n")[1].split(".")[0])
    assert d["n_gpus"] == n
    # value = the units all ranks processed / the max-over-ranks step time
    assert abs(d["value"] - d["config"]["stars_total"] / (d["ms_per_step"] * 1e-3)) <= 1e-6 * d["value"]
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert e["value"] < d["value"]                           # host buffers, copies inside the timed region
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"]
    assert not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    r = d["roofline"]
    assert r["bound"] in ("hbm", "tensor") and r["unit"] in ("GB/s", "TFLOP/s")
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    if n == 1:
        cb = d["cpu_baseline"]
        assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] > 0 and cb["sample"]


def test_ncu_facts_parses_the_committed_export(tmp_path):
    src = os.path.join(ROOT, "profiles", "r02_ncu_full_c3_raw.csv")
    facts_path = str(tmp_path / "ncu_facts.json")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_facts.py"), src, "c3"], capture_output=True,
                         text=True, check=True, env=dict(os.environ, NCU_FACTS_OUT=facts_path)).stdout
    facts = json.load(open(facts_path))["c3"]
    assert facts == json.load(open(os.path.join(ROOT, "profiles", "ncu_facts.json")))["c3"]   # the committed facts are these
    assert "fit_kernel" in out
    for k in ("fit_kernel", "pca_power_kernel", "normal_ring_kernel", "mse_ring_kernel"):
        assert facts[k]["dram_bytes_per_launch"] > 0 and facts[k]["ms_under_ncu"] > 0
    # the ring kernels move what the algorithm needs and little else (8 r c bytes per training star)
    alg = 8 * 131 * 131 * 87213
    assert 0.95 * alg < facts["normal_ring_kernel"]["dram_bytes_per_launch"] < 1.1 * alg
    assert 0.95 * alg < facts["mse_ring_kernel"]["dram_bytes_per_launch"] < 1.1 * alg
