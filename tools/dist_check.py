"""torchrun --nproc-per-node N tools/dist_check.py : multi-GPU parity (csrc/dist.cu over NCCL)
against the single-GPU pipeline and the golden .sa5 files; prints one line per case on rank 0."""
import os
import sys
import time
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psascan_b200 as P  # noqa: E402
from psascan_b200 import dist as D, textgen  # noqa: E402
from golden_cases import golden_cases, golden_text  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = P.Context(local)
D.init(ctx)
ok = True
cases = [(name, golden_text(gen, kw, n), sa5, [1000, n // 3 + 7]) for name, gen, kw, n, sa5 in golden_cases()]
big = textgen.markov(1 << 24, seed=3)
cases.append(("unary_20000", np.zeros(20000, dtype=np.uint8), None, [5000, 20000 // 3 + 1]))   # gap counters far above one byte
cases.append(("markov_16Mi", big, None, [1 << 21]))
for name, text, sa5, mbs in cases:
    n = len(text)
    d_text = torch.from_numpy(text).cuda()
    d_out = torch.zeros(5 * n, dtype=torch.uint8, device="cuda")
    if sa5 is None:
        sa5 = ctx.construct_sa(text, max_block=mbs[0])
    for mb in mbs:
        d_out.zero_()
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        cnt = D.construct_sa_distributed(ctx, d_text.data_ptr(), n, d_out.data_ptr(), mb)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        good = bool(np.array_equal(d_out.cpu().numpy(), sa5))
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = ok and bool(flag.item())
        if rank == 0:
            st = ctx.stats()
            print(f"{name} n={n} max_block={mb} world={world} ok={bool(flag.item())} {dt*1e3:.1f} ms "
                  f"walk {st['ms_walk']:.2f} ms sort {st['ms_sort']:.2f} ms coll {cnt}", flush=True)
# the slice API: every rank keeps its share of the SA; gathered here only for the comparison
for name, text, sa5, mbs in cases[-3:]:
    n = len(text)
    d_text = torch.from_numpy(text).cuda()
    capn = D.slice_capacity(ctx, n, world)
    d_slice = torch.zeros(5 * capn, dtype=torch.uint8, device="cuda")
    if sa5 is None:
        sa5 = ctx.construct_sa(text, max_block=mbs[0])
    beg, ln, cnt = D.construct_sa_slices(ctx, d_text.data_ptr(), n, d_slice.data_ptr(), mbs[0])
    good = bool(np.array_equal(d_slice[:5 * ln].cpu().numpy(), sa5[5 * beg:5 * (beg + ln)]))
    lens = [None] * world
    dist.all_gather_object(lens, (beg, ln))
    good = good and sum(x[1] for x in lens) == n and all(lens[i][0] + lens[i][1] == lens[i + 1][0] for i in range(world - 1))
    flag = torch.tensor([1 if good else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = ok and bool(flag.item())
    if rank == 0:
        print(f"slices {name} n={n} max_block={mbs[0]} world={world} ok={bool(flag.item())} shares {lens} coll {cnt}", flush=True)
D.finalize(ctx)
ctx.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
