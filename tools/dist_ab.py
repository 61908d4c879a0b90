"""torchrun --nproc-per-node N tools/dist_ab.py : A/B of environment-selected variants of the multi-GPU schedule
inside ONE process group (the library reads these variables per block, so they can be switched between
constructions): gap counters reduce-scattered as uint32 (default) or as bytes (PSC_GAP_RS8=1).
Prints per variant the device time of the construction (max over ranks) for a few alternating repetitions."""
import os
import sys
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psascan_b200 as P  # noqa: E402
from psascan_b200 import dist as D, textgen  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, mb = 1 << 31, 1 << 28
ctx = P.Context(local)
D.init(ctx)
text = textgen.markov_torch(n, seed=3, device="cuda")
cap = D.slice_capacity(ctx, n, world)
out = torch.empty(5 * cap, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
variants = [("uint32", {}), ("bytes", {"PSC_GAP_RS8": "1"})]
res = {k: [] for k, _ in variants}
for rep in range(4):
    for name, env in variants:
        for k in ("PSC_GAP_RS8",):
            os.environ.pop(k, None)
        os.environ.update(env)
        dist.barrier(); torch.cuda.synchronize()
        D.construct_sa_slices(ctx, text.data_ptr(), n, out.data_ptr(), mb)
        torch.cuda.synchronize()
        st = ctx.stats()
        t = torch.tensor([st["ms_total"]], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rep > 0:
            res[name].append(round(float(t.item()), 2))
if rank == 0:
    for k, v in res.items():
        print(f"gap reduce-scatter as {k}: ms per construction {v}", flush=True)
D.finalize(ctx)
ctx.close()
dist.destroy_process_group()
