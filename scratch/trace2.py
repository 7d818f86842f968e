"""2-rank trace of the sharded stage (torchrun): SHOPT_TRACE milestones on rank 0."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from shopt_b200 import pipeline, synth
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 100000
cat = synth.make_catalog(n, 131, 64, seed=synth.SEED + rank, device=torch.device("cuda", local))
perm = np.random.default_rng(1).permutation(n * world).astype(np.int64)
p = pipeline.default_params(n_a=64, degree=3, iterations=3, training_ratio=0.9)
pgf = torch.empty(n, 131, 131, dtype=torch.float64, device="cuda")
for it in range(3):
    if it == 2 and rank == 0:
        os.environ["SHOPT_TRACE"] = "1"
    dist.barrier(); torch.cuda.synchronize(); t = time.time()
    out = pipeline.run_stage_dev(p, cat["stamps"], cat["u"], cat["v"], cat["x0_model"], cat["x0_learned"], perm,
                                 n_global=n * world, offset=n * rank, pgf=pgf)
    torch.cuda.synchronize()
    print("rank", rank, "iter", it, "ms", (time.time() - t) * 1e3, flush=True)
dist.destroy_process_group()
