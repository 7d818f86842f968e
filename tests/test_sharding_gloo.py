"""CPU, world_size 2 and 3 over gloo: the exchange protocol of a sharded run.  Stars shard contiguously; the ranks
exchange only sum-all-reduces of small buffers: the survivor bits along the permutation (+ 3 counters), and one 4096-bin
histogram per pass of the worst-10 % radix select.  The test drives the library's HOST restatement of that protocol
(shopt_training_split_sharded_host / shopt_select_worst_sharded_host: the same order relations and messages as the
device kernels of csrc/select.cu) on several ranks with torch.distributed(gloo) as the all-reduce, and checks that every
rank derives exactly the single-process training set (shopt.jl:472-478) and rejection sequence (shopt.jl:513-520,
557-575) -- including the stable order under many exact ties and NaN errors."""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from shopt_b200 import _lib, pipeline, stage


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def single_process(flags, err0, perm, ratio, rounds):
    """The reference's own order of operations on one process (host functions of the C ABI)."""
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
    err[rng.choice(n, 5, replace=False)] = np.nan       # NaN errors sort first (isless)
    err[rng.choice(n, 3, replace=False)] = -0.0
    perm = rng.permutation(n).astype(np.int64)
    return flags, err, perm


def host_allreduce():
    def _cb(buf, count, dtype, user):
        try:
            ct = C.c_double if dtype == 0 else C.c_int32
            arr = np.ctypeslib.as_array(C.cast(buf, C.POINTER(ct)), shape=(count,))
            t = torch.from_numpy(arr)
            dist.all_reduce(t)                            # in place on the library's buffer
            return 0
        except Exception as e:                            # never let an exception cross the C boundary
            print("all-reduce failed:", e)
            return 1
    return _lib.HOST_ALLREDUCE_FN(_cb)


def sharded_protocol(flags, err0, perm, ratio, rounds, a, b, cb):
    """One rank's view: returns (global ids of its training stars, of its survivors after the rounds, global sizes)."""
    L = _lib.lib()
    n, nl = len(flags), b - a
    fl = np.ascontiguousarray(flags[a:b])
    idx = np.zeros(max(nl, 1), np.int64)
    pos = np.zeros(max(nl, 1), np.uint32)
    counts = np.zeros(4, np.int64)
    fn = C.cast(cb, C.c_void_p) if cb is not None else None
    m = L.shopt_training_split_sharded_host(fl.ctypes.data_as(_lib.ip), nl, a, n, perm.ctypes.data_as(_lib.lp), ratio, fn, None,
                                            idx.ctypes.data_as(_lib.lp), pos.ctypes.data_as(C.POINTER(C.c_uint32)),
                                            counts.ctypes.data_as(_lib.lp))
    assert m >= 0
    idx, pos = idx[:m].copy(), pos[:m].copy()
    train = idx + a
    nc = int(counts[3])
    sizes = []
    for t in range(rounds):
        sizes.append(nc)
        if t == rounds - 1:
            break
        e = np.ascontiguousarray(err0[idx + a] * (1.0 + 0.01 * t))
        mrej = int(np.ceil(0.10 * nc))
        rej = np.zeros(max(len(idx), 1), np.uint8)
        r = L.shopt_select_worst_sharded_host(e.ctypes.data_as(_lib.dp), pos.ctypes.data_as(C.POINTER(C.c_uint32)), len(idx),
                                              mrej, fn, None, rej.ctypes.data_as(C.POINTER(C.c_uint8)))
        assert r >= 0
        keep = rej[:len(idx)] == 0
        idx, pos = idx[keep], pos[keep]
        nc -= mrej
    return train, idx + a, sizes, counts


def worker(rank, world, port, n, ratio, rounds, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        flags, err0, perm = make_inputs(n)
        a, b = pipeline.shard_bounds(n, world, rank)
        cb = host_allreduce()
        train, cur, sizes, counts = sharded_protocol(flags, err0, perm, ratio, rounds, a, b, cb)
        q.put((rank, train.tolist(), cur.tolist(), sizes, counts.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_protocol_matches_single_process(world):
    n, ratio, rounds = 1003, 0.8, 4
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
    pos_of = {int(g): i for i, g in enumerate(perm)}
    train_all, cur_all = [], []
    for rank, train, cur, sizes, counts in sorted(res):
        a, b = pipeline.shard_bounds(n, world, rank)
        assert all(a <= g < b for g in train)
        assert train == sorted(train, key=pos_of.get)                    # training order = permutation order
        assert sizes == want_sizes
        assert counts[3] == len(want_train) and counts[2] == int(((flags & 7) == 0).sum())
        train_all += train
        cur_all += cur
    assert sorted(train_all, key=pos_of.get) == want_train.tolist()       # the union is the single-process split ...
    assert sorted(cur_all, key=pos_of.get) == want_cur.tolist()           # ... and so are the survivors of the rejections


def test_single_rank_protocol_is_the_reference_order():
    """world of one, no all-reduce: the radix select equals the stable descending sort of shopt.jl:513-520."""
    for n in (1, 9, 10, 11, 64, 257, 1003):
        flags, err0, perm = make_inputs(max(n, 12))
        flags, err0, perm = flags[:n], err0[:n], np.argsort(np.argsort(perm[:n])).astype(np.int64)
        want_train, want_cur, want_sizes = single_process(flags, err0, perm, 0.9, 3)
        train, cur, sizes, _ = sharded_protocol(flags, err0, perm, 0.9, 3, 0, n, None)
        assert train.tolist() == want_train.tolist() and cur.tolist() == want_cur.tolist() and sizes == want_sizes


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


# ---- per-pixel median across shards (median_constraint / NaN -> median repair, interpolate.jl:118-130) -----------------
def median_inputs(n, npx):
    rng = np.random.default_rng(7)
    z = np.round(rng.normal(size=(n, npx)), 1)            # many exact ties
    z[rng.uniform(size=z.shape) < 0.05] = np.nan
    z[:, 0] = np.nan                                      # a pixel without a finite value: median NaN
    z[:, 1] = 0.25                                        # a constant pixel
    z[: n // 2, 2] = -0.0                                 # signed zeros order below +0
    z[n // 2:, 2] = 0.0
    z[:, 3] = np.where(np.arange(n) % 2 == 0, -3.0, 5.0)  # even count, distinct middles -> v1/2 + v2/2
    return z


def median_worker(rank, world, port, n, npx, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z = median_inputs(n, npx)
        a, b = pipeline.shard_bounds(n, world, rank)
        cb = host_allreduce()
        zl = np.ascontiguousarray(z[a:b])
        med = np.zeros(npx)
        rc = _lib.lib().shopt_median_sharded_host(zl.ctypes.data_as(_lib.dp), b - a, npx, C.cast(cb, C.c_void_p), None,
                                                  med.ctypes.data_as(_lib.dp))
        q.put((rank, rc, med))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 200), (3, 201)])
def test_sharded_median_matches_numpy(world, n):
    npx = 40
    z = median_inputs(n, npx)
    with np.errstate(all="ignore"):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            want = np.nanmedian(z, axis=0)                # numpy: mean of the two middle values = v1/2 + v2/2 here
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=median_worker, args=(r, world, port, n, npx, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, rc, med in res:
        assert rc == 0
        assert np.isnan(med[0]) and med[1] == 0.25 and med[3] == (1.0 if n % 2 == 0 else -3.0)
        assert np.array_equal(np.isnan(med), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.array_equal(med[ok], want[ok]), (rank, np.abs(med[ok] - want[ok]).max())
    assert all(np.array_equal(res[0][2], r[2], equal_nan=True) for r in res)      # identical on every rank
    # and a world of one (no exchange) gives the same numbers
    med1 = np.zeros(npx)
    assert _lib.lib().shopt_median_sharded_host(np.ascontiguousarray(z).ctypes.data_as(_lib.dp), n, npx, None, None,
                                                med1.ctypes.data_as(_lib.dp)) == 0
    assert np.array_equal(med1, res[0][2], equal_nan=True)
