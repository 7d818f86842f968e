"""CPU, world_size 2 over gloo: the host-side protocol of a sharded run (shopt_stage_run_dev with an all-reduce
callback).  Stars shard contiguously; the only exchanges are sum-all-reduces: zero-padded gathers of the survivor
flags and of the per-star errors, and the [V'V | V'P] partial sums.  The test drives the SAME host functions the
library uses (shopt_training_split, shopt_worst_10_percent, shard_bounds) on two ranks and checks that both ranks
derive exactly the single-process training set and rejection sequence, and that shard partial sums add up."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from shopt_b200 import pipeline, stage


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def single_process(flags, err0, perm, ratio, rounds):
    train = pipeline.training_split(flags, perm, ratio)
    cur = train.copy()
    sizes = []
    for t in range(rounds):
        sizes.append(len(cur))
        e = err0[cur] * (1.0 + 0.01 * t)
        if t < rounds - 1:
            bad = stage.worst_10_percent(e)
            keep = np.ones(len(cur), bool)
            keep[bad] = False
            cur = cur[keep]
    return train, cur, sizes


def make_inputs(n):
    rng = np.random.default_rng(123)
    flags = np.where(rng.uniform(size=n) < 0.05, 1, 0).astype(np.int32) | np.where(rng.uniform(size=n) < 0.03, 4, 0).astype(np.int32)
    err = np.round(rng.uniform(0, 1, n), 2)             # many exact ties: the stable order must survive sharding
    perm = rng.permutation(n).astype(np.int64)
    return flags, err, perm


def worker(rank, world, port, n, ratio, rounds, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        flags, err0, perm = make_inputs(n)
        a, b = pipeline.shard_bounds(n, world, rank)
        # survivor flags: zero-padded gather through a sum-all-reduce (+32 marks "reported", as in api_stage.cu)
        g = torch.zeros(n, dtype=torch.int32)
        g[a:b] = torch.from_numpy(flags[a:b]) + 32
        dist.all_reduce(g)
        assert bool((g >= 32).all()), "every star is reported by exactly one rank"
        gflags = (g.numpy() & 31).astype(np.int32)
        train = pipeline.training_split(gflags, perm, ratio)
        cur = train.copy()
        sizes = []
        acc_total = None
        for t in range(rounds):
            sizes.append(len(cur))
            mine = (cur >= a) & (cur < b)
            # normal-equation partials of this shard (a stand-in payload with the same additive structure)
            part = torch.zeros(4, dtype=torch.float64)
            part[0] = float(mine.sum())
            part[1] = float(err0[cur[mine]].sum())
            dist.all_reduce(part)
            acc_total = part
            # per-star errors: each rank fills the positions of its own stars, zeros elsewhere; x + 0 is exact
            e = torch.zeros(len(cur), dtype=torch.float64)
            e[torch.from_numpy(np.nonzero(mine)[0])] = torch.from_numpy(err0[cur[mine]] * (1.0 + 0.01 * t))
            dist.all_reduce(e)
            if t < rounds - 1:
                bad = stage.worst_10_percent(e.numpy())
                keep = np.ones(len(cur), bool)
                keep[bad] = False
                cur = cur[keep]
        q.put((rank, train.tolist(), cur.tolist(), sizes, float(acc_total[0])))
    finally:
        dist.destroy_process_group()


def test_two_rank_protocol_matches_single_process():
    n, ratio, rounds, world = 1003, 0.8, 3, 2
    flags, err0, perm = make_inputs(n)
    want_train, want_cur, want_sizes = single_process(flags, err0, perm, ratio, rounds)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, world, port, n, ratio, rounds, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, train, cur, sizes, cnt in res:
        assert train == want_train.tolist()
        assert cur == want_cur.tolist()
        assert sizes == want_sizes
        assert cnt == want_sizes[-1]                      # shard partial sums add up to the whole


def test_shard_bounds_cover_and_split_rule():
    for n in (0, 1, 7, 100, 100001):
        for w in (1, 2, 3, 8):
            b = [pipeline.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(w - 1))
    # round(Int, x) rounds half to even (shopt.jl:477): 0.5 * 5 survivors = 2.5 -> 2; 0.5 * 7 = 3.5 -> 4
    flags = np.zeros(5, np.int32)
    assert len(pipeline.training_split(flags, np.arange(5), 0.5)) == 2
    assert len(pipeline.training_split(np.zeros(7, np.int32), np.arange(7), 0.5)) == 4
    flags = np.array([0, 1, 0, 2, 4, 0, 8, 16], np.int32)                 # bits 8/16 (training/rejected) do not exclude
    assert pipeline.training_split(flags, np.arange(8)[::-1].copy(), 1.0).tolist() == [7, 6, 5, 2, 0]
