"""The whole per-star fitting stage as one call: the host-side mirror of shopt/shopt.jl's loops :130-184 (model fit +
s-blacklist), :281-314 (PCA pixel-grid fit), :357-417 (learned fit + failed list), :472-577 (training split + polyfit
rounds with worst-10 % rejection), on top of libshopt_b200.so's shopt_stage_run / shopt_stage_run_dev.

Multi-GPU: one process per GPU, each owning a contiguous shard of the catalogue.  Stars are independent up to the
polyfit; the exchanges are sum-all-reduces of small device buffers: the normal-equation partials [V'V | V'P], the
survivor bits along the permutation (once), one 4096-bin histogram per pass of the worst-10 % radix select.  The
library owns its NCCL communicator (shopt_comm_*; torch.distributed only carries the 128-byte unique id to the other
ranks) and enqueues every exchange on the stage's stream from C; a torch.distributed callback (make_allreduce) remains
as the alternative.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check


def default_params(**over):
    """shopt.yml keys on the hot path with the README's documented defaults (README.md:300-330)."""
    p = _lib.Params()
    _lib.lib().shopt_default_params(C.byref(p))
    for k, v in over.items():
        if not hasattr(p, k):
            raise KeyError(k)
        setattr(p, k, v)
    return p


def shard_bounds(n_global, world_size, rank):
    """Contiguous shards of ceil(N/G) stars (the last ones may be shorter or empty)."""
    per = -(-n_global // world_size)
    a = min(n_global, rank * per)
    return a, min(n_global, a + per)


def num_terms(degree):
    return (degree + 1) * (degree + 2) // 2


def training_split(flags, perm, training_ratio):
    """shopt.jl:472-478 on the gathered per-star flags: global ids of the training stars, in `perm` order."""
    flags = np.ascontiguousarray(flags, dtype=np.int32)
    perm = np.ascontiguousarray(perm, dtype=np.int64)
    out = np.zeros(max(len(flags), 1), np.int64)
    n = _lib.lib().shopt_training_split(flags.ctypes.data_as(_lib.ip), len(flags), perm.ctypes.data_as(_lib.lp),
                                        training_ratio, out.ctypes.data_as(_lib.lp))
    if n < 0:
        check(int(n))
    return out[:n].copy()


class _DevView:
    """Exposes a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, count, dtype):
        self.__cuda_array_interface__ = dict(shape=(count,), typestr="<f8" if dtype == 0 else "<i4",
                                             data=(ptr, False), version=2)


def make_allreduce(group=None):
    """ctypes callback: in-place sum-all-reduce of a device buffer over `group` with torch.distributed."""
    import torch
    import torch.distributed as dist

    def _cb(ptr, count, dtype, stream, user):
        try:
            t = torch.as_tensor(_DevView(ptr, count, dtype), device="cuda")
            cur = torch.cuda.current_stream()
            if stream and int(stream) != cur.cuda_stream:    # the library's own stream (host entry points)
                with torch.cuda.stream(torch.cuda.ExternalStream(int(stream))):
                    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
            else:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
            return 0
        except Exception as e:      # never let an exception cross the C boundary
            print("shopt_b200 all-reduce failed:", e)
            return 1
    return _lib.ALLREDUCE_FN(_cb)


_COMMS = {}


def library_comm(group=None):
    """The library's own NCCL communicator over the ranks of `group` (created once per group): rank 0 draws the unique
    id, torch.distributed broadcasts its 128 bytes, every rank calls shopt_comm_init on its current device."""
    import torch
    import torch.distributed as dist
    key = id(group) if group is not None else 0
    if key in _COMMS:
        return _COMMS[key]
    L = _lib.lib()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    buf = (C.c_ubyte * _lib.UNIQUE_ID_BYTES)()
    if rank == 0:
        check(L.shopt_comm_unique_id(buf))
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    t = torch.tensor(list(buf), dtype=torch.uint8, device=dev)
    src = dist.get_global_rank(group, 0) if group is not None else 0
    dist.broadcast(t, src=src, group=group)
    raw = bytes(t.cpu().tolist())
    comm = C.c_void_p()
    check(L.shopt_comm_init(C.c_char_p(raw), rank, world, C.byref(comm)))
    _COMMS[key] = comm
    return comm


def _exchange(group, exchange):
    """(callback pointer, user pointer, keep-alive) of a sharded run, or (None, None, None)."""
    import torch.distributed as dist
    sharded = (exchange != "none" and dist.is_available() and dist.is_initialized()
               and dist.get_world_size(group) > 1)
    if not sharded:
        return None, None, None
    if exchange == "library" or (exchange == "auto" and dist.get_backend(group) == "nccl"):
        comm = library_comm(group)
        return C.cast(_lib.lib().shopt_comm_allreduce, C.c_void_p), comm, comm
    cb = make_allreduce(group)
    return C.cast(cb, C.c_void_p), None, cb


def run_stage_dev(params, stamps, u, v, x0_model, x0_learned, perm, n_global=None, offset=0, group=None, pgf=None,
                  weights=None, x0_pixels=None, exchange="auto", want_status=False):
    """Device-resident shard (torch CUDA tensors, float64): stamps (n, c, r), u, v (n), x0_* (n, 3);
    perm: global permutation (host, int64).  `group`/torch.distributed initialised -> sharded run.
    Returns dict of torch tensors (model/learned (3, n), pgf, flags, pca_k, poly (K, c, r)) + host counts."""
    import torch
    import torch.distributed as dist
    n, c, r = stamps.shape
    dev = stamps.device
    n_global = n if n_global is None else n_global
    perm_dev = None
    if type(perm).__module__.startswith("torch"):      # a device copy of the permutation: nothing to upload per call
        assert perm.is_cuda and perm.dtype == torch.int64 and perm.is_contiguous()
        perm_dev, perm = perm, None
        assert perm_dev.numel() == n_global
    else:
        perm = np.ascontiguousarray(perm, dtype=np.int64)
        assert len(perm) == n_global
    f64 = torch.float64
    K = num_terms(params.degree)
    out = dict(model=torch.empty(3, n, dtype=f64, device=dev), learned=torch.empty(3, n, dtype=f64, device=dev),
               pgf=pgf if pgf is not None else torch.empty(n, c, r, dtype=f64, device=dev),
               flags=torch.empty(n, dtype=torch.int32, device=dev), pca_k=torch.empty(n, dtype=torch.int32, device=dev),
               poly=torch.empty(K, c, r, dtype=f64, device=dev))
    npr = np.zeros(max(params.iterations, 1), np.int64)
    for t in (stamps, u, v, x0_model, x0_learned):
        assert t.is_cuda and t.is_contiguous() and t.dtype == f64
    io = _lib.StageIO(stamps=stamps.data_ptr(), r=r, c=c, n_local=n, u=u.data_ptr(), v=v.data_ptr(),
                      x0_model=x0_model.data_ptr(), x0_learned=x0_learned.data_ptr(), n_global=n_global, offset=offset,
                      perm=perm.ctypes.data if perm is not None else None,
                      perm_dev=perm_dev.data_ptr() if perm_dev is not None else None,
                      model=out["model"].data_ptr(), learned=out["learned"].data_ptr(),
                      pgf=out["pgf"].data_ptr(), flags=out["flags"].data_ptr(), pca_k=out["pca_k"].data_ptr(),
                      poly=out["poly"].data_ptr(), n_per_round=npr.ctypes.data,
                      weights=weights.data_ptr() if weights is not None else None,
                      x0_pixels=x0_pixels.data_ptr() if x0_pixels is not None else None)
    if want_status:
        out["status"] = torch.empty(4, n, dtype=torch.int32, device=dev)   # model status, iters, learned status, iters
        io.model_status, io.model_iters = out["status"][0].data_ptr(), out["status"][1].data_ptr()
        io.learned_status, io.learned_iters = out["status"][2].data_ptr(), out["status"][3].data_ptr()
    fn, user, keep = _exchange(group, exchange)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(_lib.lib().shopt_stage_run_dev(C.byref(params), C.byref(io), fn, user, stream))
    del keep
    return _finish(out, io, npr, params)


def _finish(out, io, npr, params):
    out["n_per_round"] = npr[:params.iterations].copy()
    out["counts"] = dict(zip(("blacklisted", "failed", "training", "kept"), list(io.counts)))
    out["field"] = np.array(io.field).reshape(3, 10)     # s(u,v), g1(u,v), g2(u,v): shopt.jl:422-466
    out["field_rank"] = int(io.field_rank)
    out["polyfit_rank"], out["polyfit_info"] = int(io.polyfit_rank), int(io.polyfit_info)
    return out


def run_stage(params, stamps, u, v, x0_model, x0_learned, perm, want_pgf=True, weights=None, x0_pixels=None,
              n_global=None, offset=0, group=None, exchange="auto"):
    """Host arrays (H2D, stage, D2H inside the call): stamps (N, c, r), u, v (N), x0_* (N, 3), perm (n_global).
    stamps of dtype float32 are taken as the raw FITS VIGNET cube: uploaded as float32 and ingested on the device
    (nanMask, unity sum), see shopt_stage_io.stamps_f32.
    One GPU by default; with torch.distributed initialised and n_global / offset given, the arrays are this rank's
    shard of a sharded run (shopt_stage_run_comm)."""
    import ctypes as C
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    f32 = np.asarray(stamps).dtype == np.float32
    stamps = np.ascontiguousarray(stamps, dtype=np.float32) if f32 else f(stamps)
    u, v, x0_model, x0_learned = f(u), f(v), f(x0_model), f(x0_learned)
    perm = np.ascontiguousarray(perm, dtype=np.int64)
    weights = None if weights is None else f(weights)
    x0_pixels = None if x0_pixels is None else f(x0_pixels)
    N, c, r = stamps.shape
    K = num_terms(params.degree)
    out = dict(model=np.zeros((3, N)), learned=np.zeros((3, N)), pgf=np.zeros((N, c, r)) if want_pgf else None,
               flags=np.zeros(N, np.int32), pca_k=np.zeros(N, np.int32), poly=np.zeros((K, c, r)))
    npr = np.zeros(max(params.iterations, 1), np.int64)
    io = _lib.StageIO(stamps=stamps.ctypes.data, r=r, c=c, n_local=N, u=u.ctypes.data, v=v.ctypes.data,
                      x0_model=x0_model.ctypes.data, x0_learned=x0_learned.ctypes.data, n_global=N, offset=0,
                      perm=perm.ctypes.data, model=out["model"].ctypes.data, learned=out["learned"].ctypes.data,
                      pgf=out["pgf"].ctypes.data if want_pgf else None, flags=out["flags"].ctypes.data,
                      pca_k=out["pca_k"].ctypes.data, poly=out["poly"].ctypes.data, n_per_round=npr.ctypes.data,
                      weights=weights.ctypes.data if weights is not None else None,
                      x0_pixels=x0_pixels.ctypes.data if x0_pixels is not None else None, stamps_f32=int(f32))
    if n_global is not None and n_global != N:
        import torch.distributed as dist
        assert dist.is_initialized() and len(perm) == n_global
        io.n_global, io.offset = n_global, offset
        fn, user, keep = _exchange(group, exchange)
        check(_lib.lib().shopt_stage_run_comm(C.byref(params), C.byref(io), fn, user))
        del keep
    else:
        check(_lib.lib().shopt_stage_run(C.byref(params), C.byref(io)))
    return _finish(out, io, npr, params)
