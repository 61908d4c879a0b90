"""torchrun --nproc-per-node N tools/run_config3.py [log2_bytes=34] [log2_block=30] [steps=1]
BASELINE configs[2]: order-5 Markov text (the bench generator, seed 3) at its stated 16 GiB over N GPUs
(csrc/dist.cu).  The text and the suffix-array slices stay in HBM; every rank verifies its slice
(psc_verify_sa_slice_d).  Prints one JSON line on rank 0."""
import json, os, sys, time
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psascan_b200 as P  # noqa: E402
from psascan_b200 import dist as D, textgen  # noqa: E402

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 34
lgb = int(sys.argv[2]) if len(sys.argv) > 2 else 30
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
n, mb = 1 << lg, 1 << lgb
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = P.Context(local)
D.init(ctx, rank, world)
t0 = time.time()
text = textgen.markov_torch(n, seed=3) if rank == 0 else torch.empty(n, dtype=torch.uint8, device="cuda")
if world > 1:
    dist.broadcast(text, src=0)
torch.cuda.synchronize()
gen_s = time.time() - t0
cap = D.slice_capacity(ctx, n, world)
out = torch.empty(5 * cap, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
res = []
for it in range(steps + 1):   # the first construction pays for the device allocations
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t1 = time.time()
    beg, ln, comm = D.construct_sa_slices(ctx, text.data_ptr(), n, out.data_ptr(), mb)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    res.append((time.time() - t1, ctx.stats()))
first = torch.zeros(8, dtype=torch.uint8, device="cuda")
first[:5] = out[:5]
firsts = [torch.zeros(8, dtype=torch.uint8, device="cuda") for _ in range(world)]
if world > 1:
    dist.all_gather(firsts, first)
nxt = int.from_bytes(bytes(firsts[rank + 1][:5].cpu().tolist()), "little") if rank + 1 < world else -1
t2 = time.time()
viol = ctx.verify_sa_slice_ptr(text.data_ptr(), n, out.data_ptr(), beg, ln, nxt)
ver_s = time.time() - t2
flag = torch.tensor([0 if any(viol) else 1], device="cuda")
ms = torch.tensor([r[1]["ms_total"] for r in res], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    st = res[-1][1]
    B = st["n_blocks"]
    steps_total = sum(max(0, n - min(n, (i + 1) * mb)) for i in range(B))
    best = float(ms[1:].min()) if steps else float(ms[0])
    print(json.dumps({"workload": "configs[2]: order-5 Markov text, seed 3", "text_bytes": n, "max_block": mb, "n_blocks": B,
                      "n_gpus": world, "ms_cold": float(ms[0]), "ms_per_construction": best, "MB_per_s": n / 1e6 / (best / 1e3),
                      "wall_s": [round(r[0], 3) for r in res], "lf_steps": steps_total,
                      "lf_steps_per_s_job": steps_total / (st["ms_walk"] / 1e3), "verified": bool(flag.item()),
                      "verify": "psc_verify_sa_slice_d on every rank", "verify_s": round(ver_s, 2), "text_gen_s": round(gen_s, 1),
                      "comm_per_rank": comm, "phases_ms_rank0": {k: round(v, 1) for k, v in st.items() if k.startswith("ms_") and v}}), flush=True)
D.finalize(ctx)
ctx.close()
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
