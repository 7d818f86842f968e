"""Worker of tests/test_gpu_multirank.py (launched by torch.distributed.run, one rank per GPU): the sharded stage over
NCCL must reproduce the single-GPU stage -- flags, training set and rejection bit-exact, per-star fits bit-identical
(same kernels on the same stars), polynomial coefficients equal up to the summation order of the all-reduce."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shopt_b200 import pipeline, synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N, r, n_a = 1501, 41, 32
    cat = synth.make_catalog(N, r, n_a, frac_bad_s=0.03, frac_contaminated=0.03)
    p = pipeline.default_params(n_a=n_a, degree=3, iterations=3, training_ratio=0.9)
    perm = cat["perm"].numpy()
    a, b = pipeline.shard_bounds(N, world, rank)
    names = ("stamps", "u", "v", "x0_model", "x0_learned")
    shard = {k: cat[k][a:b].contiguous().cuda() for k in names}
    got = pipeline.run_stage_dev(p, shard["stamps"], shard["u"], shard["v"], shard["x0_model"], shard["x0_learned"], perm,
                                 n_global=N, offset=a)
    torch.cuda.synchronize()
    # single-GPU run of the whole catalogue on every rank (no process group involved: world of one)
    full = {k: cat[k].cuda() for k in names}
    io_one = pipeline.run_stage_dev(p, full["stamps"], full["u"], full["v"], full["x0_model"], full["x0_learned"], perm,
                                    exchange="none")
    torch.cuda.synchronize()
    assert got["counts"] == io_one["counts"], (got["counts"], io_one["counts"])
    assert got["n_per_round"].tolist() == io_one["n_per_round"].tolist()
    assert torch.equal(got["flags"], io_one["flags"][a:b])
    assert torch.equal(got["model"], io_one["model"][:, a:b]) and torch.equal(got["pgf"], io_one["pgf"][a:b])
    poly, ref = got["poly"].cpu().numpy(), io_one["poly"].cpu().numpy()
    scale = np.abs(ref).max(axis=(1, 2), keepdims=True)
    err = np.max(np.abs(poly - ref) / scale)
    assert err < 1e-10, err
    # s(u,v), g1(u,v), g2(u,v): the normal equations add over the shards (one 130-double all-reduce)
    assert got["field_rank"] == io_one["field_rank"] == 10
    assert np.allclose(got["field"], io_one["field"], rtol=1e-9, atol=1e-12), np.abs(got["field"] - io_one["field"]).max()
    # every rank holds the same replicated result
    gathered = [torch.empty_like(got["poly"]) for _ in range(world)]
    dist.all_gather(gathered, got["poly"])
    assert all(torch.equal(g, gathered[0]) for g in gathered)
    # ---- options that need a per-pixel median over the training stars of EVERY shard (interpolate.jl:118-130) ----
    from shopt_b200 import _lib

    def sharded_vs_single(params, stamps, **kw):
        sh = {k: v[a:b].contiguous().cuda() for k, v in kw.items()}
        g = pipeline.run_stage_dev(params, stamps[a:b].contiguous().cuda(), shard["u"], shard["v"], shard["x0_model"],
                                   shard["x0_learned"], perm, n_global=N, offset=a, **sh)
        torch.cuda.synchronize()
        one = pipeline.run_stage_dev(params, stamps.cuda(), full["u"], full["v"], full["x0_model"], full["x0_learned"], perm,
                                     exchange="none", **{k: v.cuda() for k, v in kw.items()})
        torch.cuda.synchronize()
        assert g["counts"] == one["counts"], (g["counts"], one["counts"])
        assert g["n_per_round"].tolist() == one["n_per_round"].tolist()
        assert torch.equal(g["flags"], one["flags"][a:b])
        gp, rp = g["poly"].cpu().numpy(), one["poly"].cpu().numpy()
        assert np.isfinite(rp).all() and np.isfinite(gp).all()
        sc = np.abs(rp).max(axis=(1, 2), keepdims=True)
        return gp, rp, float(np.max(np.abs(gp - rp) / sc))

    # (B) median_constraint: coefficient 1 IS the median, selected by integer radix passes -> bit-identical
    pm = pipeline.default_params(n_a=n_a, degree=2, iterations=2, training_ratio=0.9, median_constraint=1)
    gp, rp, err_b = sharded_vs_single(pm, cat["stamps"])
    assert np.array_equal(gp[0], rp[0]), "per-pixel medians differ between the sharded and the single-GPU run"
    assert err_b < 1e-10, err_b
    # (C) NaN pixel-grid values among the training stars (mode chisq, masked pixels keep a NaN start value): the NaN ->
    #     median repair before the solve needs the other shards' values of that pixel
    catm = synth.make_catalog(N, r, n_a, frac_masked=0.01)
    pc = pipeline.default_params(n_a=n_a, degree=2, iterations=2, training_ratio=0.9, mode=_lib.MODE_CHISQ,
                                 chisq_fill=float("nan"))
    gp, rp, err_c = sharded_vs_single(pc, catm["stamps"], weights=torch.ones_like(catm["stamps"]))
    assert err_c < 1e-10, err_c
    if rank == 0:
        print("MULTIRANK_OK world=%d coef_err=%.2e median_constraint_err=%.2e nan_repair_err=%.2e counts=%s"
              % (world, err, err_b, err_c, got["counts"]))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
