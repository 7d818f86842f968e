#!/usr/bin/env python
"""bench.py -- stars/s of the per-star PSF fitting stage (analytic LBFGS model fit + PCA pixel-grid fit + learned
analytic fit + degree-3 (u,v) polynomial interpolation with 3 rounds of worst-10 % rejection) on synthetic catalogues.

  python bench.py --gpus N --steps K --warmup W            our arm   (N > 1: launched by torch.distributed.run)
  python bench.py --impl reference ...                     the reference algorithm on the host cores (CPU oracle port)

A "step" is one pass of the whole stage over one synthetic catalogue.  Workload (BASELINE.json configs[2], the one
the metric "stars/sec fitted (LBFGS+PCA+poly interp)" is quoted on): 100 000 stars of 131 x 131 per GPU, n_a = 64,
PCAterms = 50, lanczos = 3, degree 3, training_ratio 0.9, 3 rounds.  Sharding is weak (100k stars per GPU).
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for what each key means.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "stars/sec fitted (LBFGS+PCA+poly interp)"
UNIT = "stars/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--stars", type=int, default=100000, help="stars per GPU")
    ap.add_argument("--stamp", type=int, default=131)
    ap.add_argument("--n-a", type=int, default=64)
    ap.add_argument("--degree", type=int, default=3)
    ap.add_argument("--cpu-sample", type=int, default=2048, help="stars of the CPU baseline sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def stage_params(args):
    from shopt_b200 import pipeline
    return pipeline.default_params(n_a=args.n_a, degree=args.degree, iterations=3, training_ratio=0.9, pca_terms=50,
                                   lanczos=3, unity_sum=0, g_tol_model=1e-6, g_tol_learned=1e-6)


def workload_config(args, n_gpus):
    return {"workload": "synthetic %dk Gaussian-PSF stars per GPU, %dx%d stamps, analytic(n_a=%d)+PCA(n=50)+degree-%d "
                        "(u,v) interpolation, 3 rounds" % (args.stars // 1000, args.stamp, args.stamp, args.n_a, args.degree),
            "stars_per_gpu": args.stars, "stars_total": args.stars * n_gpus, "stamp": args.stamp, "n_a": args.n_a,
            "pca_terms": 50, "degree": args.degree, "iterations": 3, "training_ratio": 0.9,
            "sharding": "stars, contiguous shards, dp%d" % n_gpus,
            "l2": "inputs (%.1f GB per GPU) far exceed the 126 MB L2; no flush needed" % (args.stars * args.stamp ** 2 * 8 / 1e9)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.t = index, [], threading.Event(), None
        self.nvml = None
        try:   # in-process NVML: a query costs microseconds; spawning nvidia-smi stalls the driver for ~100 ms per call,
            import pynvml   # which at 2+ ranks showed up as +70 % step time
            pynvml.nvmlInit()
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = (pynvml, h)
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        nv, h = self.nvml
        sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        try:
            bits = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        act = lambda m: "Active" if bits & m else "Not Active"
        return [str(sm), str(mx), act(0x8), act(0x40), act(0x20), act(0x4)]

    def _run(self):
        while not self.stop.is_set():
            try:
                if self.nvml:
                    self.rows.append(self._sample_nvml())
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                    self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.05 if self.nvml else 0.5)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------------------------
# CPU side: the reference's algorithm on the host cores (oracle port; the reference itself is Julia and cannot run
# on this box -- SURVEY.md 8c).  Used for `cpu_baseline` (rank 0, N = 1) and for the whole `--impl reference` arm.
# --------------------------------------------------------------------------------------------------------------
def cpu_stage_once(args, sample, seed_shift=0):
    """One pass of the whole stage on `sample` stars with every host thread; returns (seconds, threads)."""
    import numpy as np
    from oracle import shopt_oracle as orc
    from oracle import stage_oracle
    from shopt_b200 import synth
    orc.set_reference_shaped(1)        # value and gradient evaluated separately, like NLSolversBase's fg! does
    nthr = orc.max_threads()
    cat = synth.make_catalog(sample, args.stamp, args.n_a, seed=synth.SEED + seed_shift)
    a = {k: cat[k].numpy() for k in ("stamps", "u", "v", "x0_model", "x0_learned", "perm")}
    p = stage_params(args)
    t0 = time.perf_counter()
    stage_oracle.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], a["perm"], nthreads=nthr)
    return time.perf_counter() - t0, nthr


def cpu_baseline(args):
    sec, nthr = cpu_stage_once(args, args.cpu_sample)
    return {"value": args.cpu_sample / sec, "unit": UNIT, "cores": nthr, "kind": "port",
            "sample": "%d stars of the same workload, whole stage once, OpenMP over stars (the reference's star loops are "
                      "serial; only its polyfit is threaded), value+gradient evaluated reference-shaped" % args.cpu_sample,
            "seconds": sec}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times = []
    for i in range(args.warmup + args.steps):
        sec, nthr = cpu_stage_once(args, args.cpu_sample, seed_shift=i)
        if i >= args.warmup:
            times.append(sec)
    ms = 1e3 * sum(times) / len(times)
    value = args.cpu_sample / (ms * 1e-3)
    cfg = workload_config(args, args.gpus)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg, "gpu_launches": 0,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthr, "kind": "port",
                             "sample": "each step = the whole stage on %d stars of the workload with %d OpenMP threads over "
                                       "stars; the reference (Julia) is not runnable here, this is the CPU restatement in "
                                       "oracle/ (reference-shaped evaluation)" % (args.cpu_sample, nthr)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------------------
KERNELS_PER_STEP = 4 + 2 + 3 * 7   # fit, pca, fit, flags; field normal + reduce; 3 rounds x (normal, gram, reduce, median_fix, solve, mse, mse_reduce)


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from shopt_b200 import _lib, pipeline, synth
    if not torch.cuda.is_available() or _lib.lib().shopt_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    p = stage_params(args)
    n, r = args.stars, args.stamp
    n_global, offset = n * world, n * rank
    cat = synth.make_catalog(n, r, args.n_a, seed=synth.SEED + rank, device=dev)
    perm = np.random.default_rng(synth.SEED).permutation(n_global).astype(np.int64)
    pgf = torch.empty(n, r, r, dtype=torch.float64, device=dev)
    rc = r * r

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        return pipeline.run_stage_dev(p, cat["stamps"], cat["u"], cat["v"], cat["x0_model"], cat["x0_learned"], perm,
                                      n_global=n_global, offset=offset, pgf=pgf)

    for _ in range(args.warmup):
        out = step()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        ev0.record()
        for _ in range(args.steps):
            out = step()
        ev1.record()
        barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item()) / args.steps
    value = n_global / (ms_per_step * 1e-3)

    # ---- per-kernel device times (CUDA events on the launching stream), for the roofline of the dominant kernel ----
    stages = stage_breakdown(args, p, cat, pgf, out)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback (of fallback)"
    dom = max(stages, key=lambda k: stages[k]["ms"])
    roofline = {"bound": "hbm", "kernel": dom, "achieved": stages[dom]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": stages[dom]["gbs"] / peak, "traffic": None, "peak_source": peak_src,
                "note": "the dominant kernel is FP64-compute-bound (Jacobi PCA); the HBM-bound kernels of the step are "
                        "listed under stages with their own fractions",
                "stages": {k: dict(v, frac=v["gbs"] / peak) for k, v in stages.items()},
                "step_algorithmic_bytes_per_star": 6.878 * 8 * rc, "step_hbm_frac": value / world * 6.878 * 8 * rc / 1e9 / peak}

    # ---- end to end: host buffers in, results out, copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, p, cat, perm, world, n_global, offset, pgf, barrier)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args)
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": workload_config(args, world), "clocks": clk.summary(),
                "gpu_launches": KERNELS_PER_STEP * args.steps, "e2e": e2e, "roofline": roofline, "cpu_baseline": cpu,
                "counts": out["counts"], "n_per_round": [int(x) for x in out["n_per_round"]]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def stage_breakdown(args, p, cat, pgf, out):
    """Each kernel of the step timed alone with CUDA events on the launching (torch current) stream; GB/s from the
    ALGORITHMIC bytes of DESIGN.md (what must cross HBM once), not from counters."""
    import ctypes as C
    import torch
    from shopt_b200 import _lib, stage
    L = _lib.lib()
    n, r = args.stars, args.stamp
    rc = r * r
    dev = pgf.device
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    train = torch.nonzero(out["flags"] & _lib.FLAG_TRAINING).flatten().contiguous()
    nt = int(train.numel())
    K = (args.degree + 1) * (args.degree + 2) // 2
    acc = torch.empty(L.shopt_polyfit_acc_len(r, r, args.degree), dtype=torch.float64, device=dev)
    poly = torch.empty(K, r, r, dtype=torch.float64, device=dev)
    mse = torch.empty(max(nt, 1), dtype=torch.float64, device=dev)
    info = torch.zeros(4, dtype=torch.int32, device=dev)
    s = torch.empty(3, n, dtype=torch.float64, device=dev)
    ist = torch.empty(2, n, dtype=torch.int32, device=dev)
    u, v = cat["u"], cat["v"]

    def timed(fn, reps=2):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def fit(src, x0, tol):
        _lib.check(L.shopt_fit_analytic_batch_dev(src.data_ptr(), r, r, n, args.n_a, x0.data_ptr(), tol, p.max_iter,
                                                  s[0].data_ptr(), s[1].data_ptr(), s[2].data_ptr(), ist[0].data_ptr(),
                                                  ist[1].data_ptr(), st))
    res = {}
    fit_bytes = n * (8 * args.n_a ** 2 + 32 + 24)
    t = timed(lambda: fit(cat["stamps"], cat["x0_model"], p.g_tol_model))
    res["fit_kernel(model)"] = {"ms": t, "bytes": fit_bytes, "gbs": fit_bytes / t / 1e6}
    t = timed(lambda: _lib.check(L.shopt_pca_batch_dev(cat["stamps"].data_ptr(), r, r, n, p.pca_terms, p.pratio, p.lanczos,
                                                       p.unity_sum, pgf.data_ptr(), ist[0].data_ptr(), ist[1].data_ptr(), st)))
    res["pca_kernel"] = {"ms": t, "bytes": 2 * 8 * rc * n, "gbs": 2 * 8 * rc * n / t / 1e6}
    t = timed(lambda: fit(pgf, cat["x0_learned"], p.g_tol_learned))
    res["fit_kernel(learned)"] = {"ms": t, "bytes": fit_bytes, "gbs": fit_bytes / t / 1e6}
    if nt >= 5:
        t = timed(lambda: _lib.check(L.shopt_polyfit_normal_dev(pgf.data_ptr(), r, r, u.data_ptr(), v.data_ptr(),
                                                                train.data_ptr(), nt, args.degree, acc.data_ptr(), st)), 3)
        res["normal_kernel(round 1)"] = {"ms": t, "bytes": 8 * rc * nt, "gbs": 8 * rc * nt / t / 1e6}
        t = timed(lambda: _lib.check(L.shopt_polyfit_solve_dev(acc.data_ptr(), r, r, args.degree, poly.data_ptr(),
                                                               info.data_ptr(), st)), 3)
        res["solve_kernel"] = {"ms": t, "bytes": 16 * K * rc, "gbs": 16 * K * rc / t / 1e6}
        t = timed(lambda: _lib.check(L.shopt_polyfit_mse_dev(pgf.data_ptr(), r, r, u.data_ptr(), v.data_ptr(),
                                                             train.data_ptr(), nt, args.degree, poly.data_ptr(),
                                                             mse.data_ptr(), st)), 3)
        res["mse_kernel(round 1)"] = {"ms": t, "bytes": 8 * rc * nt, "gbs": 8 * rc * nt / t / 1e6}
    return res


def run_e2e(args, p, cat, perm, world, n_global, offset, pgf, barrier):
    """The same step through the public API with HOST buffers: pinned inputs -> H2D -> stage -> D2H of the results
    (model/learned parameters, flags, polynomial cube) through the C ABI host entry points: shopt_stage_run at N = 1,
    shopt_stage_run_comm (this rank's shard + the all-reduce callback) at N > 1; the copies happen inside the library,
    chunked and overlapped with the per-star kernels."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from shopt_b200 import pipeline
    n, r = args.stars, args.stamp
    K = (args.degree + 1) * (args.degree + 2) // 2
    names = ("stamps", "u", "v", "x0_model", "x0_learned")
    host = {k: torch.empty(cat[k].shape, dtype=torch.float64, pin_memory=True) for k in names}
    for k in names:
        host[k].copy_(cat[k])
    torch.cuda.synchronize()
    h2d = sum(host[k].numel() * 8 for k in names)
    d2h = 8 * (6 * n + K * r * r) + 8 * n
    steps = max(1, min(args.steps, 2))

    def once():
        a = {k: host[k].numpy() for k in names}
        if world == 1:
            return pipeline.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], perm, want_pgf=False)
        return pipeline.run_stage(p, a["stamps"], a["u"], a["v"], a["x0_model"], a["x0_learned"], perm, want_pgf=False,
                                  n_global=n_global, offset=offset)
    once()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        once()
    barrier()
    sec = torch.tensor([(time.perf_counter() - t0) / steps], dtype=torch.float64, device=pgf.device)
    if world > 1:
        dist.all_reduce(sec, op=dist.ReduceOp.MAX)
    return {"value": n_global / float(sec.item()), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "ms_per_step": 1e3 * float(sec.item()), "steps": steps,
            "api": "shopt_stage_run (C ABI, host buffers)" if world == 1 else
            "shopt_stage_run_comm (C ABI, host buffers of the shard, torch.distributed all-reduce callback)"}


def main():
    args = parse_args()
    # stdout carries exactly one JSON line: anything libraries print to fd 1 meanwhile (NCCL's version banner, ...) is
    # sent to stderr, and print() keeps the real stdout
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    real.flush()


if __name__ == "__main__":
    main()
