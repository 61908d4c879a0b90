#!/usr/bin/env python
"""bench.py -- SA construction throughput and LF steps/s of the B200 hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--size BYTES] [--impl reference]

A step = one SA construction over one synthetic text (text -> 40-bit .sa5 entries).
Default workload: the configs[2] generator (order-5 Markov, 30 symbols) at 2 GiB in 8 blocks of 256 MiB,
the largest multiple-of-8-blocks size whose text, block suffix arrays, gap arrays and output stay resident
in one B200 with room for the e2e leg's pinned buffers; it runs the whole pSAscan
pipeline (block sorts, rank structures, 7.5 G LF steps, gap-guided merge).
  value : input MB/s with the text and the output resident in HBM (device time, CUDA events recorded
          by the library on its own stream around the whole construction)
  e2e   : the same through the host-buffer entry point (pinned host text in, .sa5 bytes out to pinned
          host memory; both copies inside the timed region)
N > 1: one process per GPU (torchrun), the SAME text sharded over the ranks (csrc/dist.cu: block sorts on
separate GPUs, the tail walk of every block split over the ranks, gap counts summed over NVLink, merge
sharded by output range): strong scaling.  Every rank keeps n/N of the suffix array.
After the timed loops the last result is verified on the device (psc_verify_sa_d, O(n)); a wrong
suffix array makes the run fail.
--impl reference times the unmodified reference construct_sa (oracle/_ref) on the host cores, on the
same full text.
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (generator, kwargs, default size, max_block (0 = whole text), description)
    "markov8": ("markov", dict(seed=3), 1 << 31, 1 << 28,
                "configs[2] generator (order-5 Markov, 30 symbols, seed 3) at 2 GiB in 8 blocks of 256 MiB: "
                "block sorts + rank structures + 7.5 G LF steps + gap-guided merge"),
    "random1g": ("random", dict(seed=2), 1 << 30, 0,
                 "configs[1]: 1 GiB uniform-random bytes (sigma=255), single block in HBM"),
    "dna200m": ("dna", dict(seed=1), 200_000_000, 0, "configs[0]: 200 MB DNA-like (sigma=4), single block"),
    "markov_multi": ("markov", dict(seed=3), 1 << 30, 1 << 27,
                     "configs[2] scaled 1/16: 1 GiB order-5 Markov text, 8 blocks of 128 MiB, gap streaming"),
    "repetitive_multi": ("repetitive", dict(seed=4), 1 << 28, 1 << 26,
                         "configs[3] scaled: mutated copies of a 1 MiB seed, 4 blocks"),
}
DEFAULT_WORKLOAD = "markov8"
METRIC = "SA construction input MB/s"


def gen_text(workload, size):
    """the workload's text as a numpy array; GiB-sized Markov texts come from the GPU generator when a
    GPU is there (same model and seed on every rank and in the reference arm, so both arms sort the
    same bytes on the same box)"""
    import numpy as np
    from psascan_b200 import textgen
    gen, kw, dsize, mb, _ = WORKLOADS[workload]
    n = size or dsize
    kw = dict(kw)
    source = "numpy"
    if gen == "random" and n >= (1 << 26):
        rng = np.random.default_rng(kw["seed"])
        t = rng.integers(0, 255, size=n, dtype=np.uint8)
    elif gen == "markov" and n >= (1 << 26) and _cuda_ok():
        t = textgen.markov_torch(n, seed=kw["seed"]).cpu().numpy()
        source = "torch-cuda"
    else:
        t = textgen.GENERATORS[gen](n, **kw)
    return t, (mb if mb < n else 0), source


def _cuda_ok():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def mem_available_bytes():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 8 << 30


def reference_budget(n, threads, n_blocks=1):
    """(-m bytes, threads) for the reference: 14 MiB of its budget go to every thread and a block takes
    5.2 bytes per symbol (psascan.hpp:73-91).  n_blocks == 1: the largest budget the host allows (70 % of
    MemAvailable, the text and the 5n-byte output live in tmpfs beside it); else the budget that cuts the
    text into n_blocks blocks."""
    per_thread = 14 << 20
    if n_blocks > 1:
        block = -(-n // n_blocks)
        return int(5.2 * block) + per_thread * threads + (1 << 20), threads
    avail = mem_available_bytes()
    budget = int(0.7 * max(0, avail - 6 * n))
    threads = max(1, min(threads, (budget // 2) // per_thread))
    return max(budget, per_thread * threads + (64 << 20)), threads


def run_reference(text, threads, mem_bytes):
    """one run of oracle/_ref/construct_sa (the unmodified reference, libsais variant) on `text` in tmpfs;
    returns (wall seconds, LF steps streamed, seconds inside its streaming phases, its own MiB/s line)"""
    import oracle_lib
    d = tempfile.mkdtemp(prefix="psc_bench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    f = os.path.join(d, "t.txt")
    text.tofile(f)
    out = f + ".sa5"
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    try:
        t0 = time.perf_counter()
        p = subprocess.run([oracle_lib.REF_BIN, f, "-o", out, "-m", str(mem_bytes)], check=True, env=env,
                           stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, stdin=subprocess.DEVNULL)
        dt = time.perf_counter() - t0
    finally:
        for x in os.listdir(d):
            os.remove(os.path.join(d, x))
        os.rmdir(d)
    # compute_gap.hpp:153-156: "Stream: 100.0%. Time: %.2Lfs. Speed: %.2LfMiB/s" per streaming pass
    err = p.stderr.decode(errors="replace").replace("\r", "\n")
    steps, secs = 0.0, 0.0
    for m in re.finditer(r"Stream: 100\.0%\. Time: ([0-9.]+)s\. Speed: ([0-9.]+)MiB/s", err):
        tm, sp = float(m.group(1)), float(m.group(2))
        secs += tm
        steps += tm * sp * (1 << 20)
    own = re.search(r"speed: ([0-9.]+)MiB/s", err)
    blk = re.search(r"Max block size = (\d+)", err)
    return dt, steps, secs, (float(own.group(1)) if own else None), (int(blk.group(1)) if blk else None)


def reference_arm(a, config, n):
    """--impl reference: the unmodified reference on the SAME full text, all host cores, the largest -m the
    host allows.  Runs longer than 120 s are timed once, without warm-up (steps_timed says how many)."""
    text, _, source = gen_text(a.workload, n)
    cores = host_threads()
    mem, thr = reference_budget(n, cores)
    budget_s = 240.0
    times, info = [], None
    t_start = time.perf_counter()
    first = run_reference(text, thr, mem)
    times.append(first[0])
    info = first
    if first[0] <= 120.0:
        # the first run doubles as the warm-up when there is time for more
        want = a.steps + (1 if a.warmup > 0 else 0)
        while len(times) < want and (time.perf_counter() - t_start) + first[0] < budget_s:
            r = run_reference(text, thr, mem)
            times.append(r[0])
            info = r
        if len(times) > 1 and a.warmup > 0:
            times = times[1:]
    sec = sum(times) / len(times)
    v = n / 1e6 / sec
    sample = (f"the whole {n}-byte text ({source} generator), construct_sa -m {mem} (libsais build), "
              f"OMP_NUM_THREADS={thr} of {cores} cores, max block {info[4]}, files in tmpfs")
    cb = {"value": v, "unit": "MB/s", "cores": thr, "kind": "reference", "sample": sample,
          "text_bytes_sorted": n, "mem_budget_bytes": mem, "own_speed_mib_s": info[3]}
    if info[2] > 0:
        cb["lf_steps_per_s"] = info[1] / info[2]
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "MB/s",
                      "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "steps_timed": len(times),
                      "ms_per_step": sec * 1e3, "higher_is_better": True,
                      "scaling": "strong" if a.gpus > 1 else "weak", "vs_baseline": None, "dtype": "u8 text, u64 keys, u32 ranks",
                      "data": "synthetic", "config": dict(config, text_source=source), "cpu_baseline": cb,
                      "e2e": {"value": v, "unit": "MB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}))


def cpu_sample_baseline(text, n_blocks, sample_n):
    """cpu_baseline leg of the GPU arm: the reference on a bounded prefix of the workload, cut into the same
    NUMBER of blocks as the GPU run so that its streaming phase (LF steps/s) is exercised too"""
    cores = host_threads()
    sample_n = min(len(text), sample_n)
    mem, thr = reference_budget(sample_n, min(cores, 64), max(1, n_blocks))
    dt, steps, secs, own, blk = run_reference(text[:sample_n], thr, mem)
    cb = {"value": sample_n / 1e6 / dt, "unit": "MB/s", "cores": thr, "kind": "reference",
          "sample": (f"first {sample_n} bytes of the workload, construct_sa -m {mem} (libsais build; max block {blk}: "
                     f"{n_blocks} blocks like the GPU run), OMP_NUM_THREADS={thr}, tmpfs"),
          "text_bytes_sorted": sample_n, "seconds": dt}
    if secs > 0:
        cb["lf_steps_per_s"] = steps / secs
        cb["lf_steps"] = steps
    return cb


def walk_roofline(s, r32, r32_ws, ws_bytes, peak, peak_src):
    """roofline of the dominant kernel, lf_walk_kernel (K6) alone: every LF step reads ONE random 32-byte sector of
    the block's rank planes (plus 5.25 sequential bytes: 1 text symbol, a gt bit in and out, the 4-byte rank it
    leaves in a page).  The ceiling is the rate of independent random 32-byte sector reads over a buffer of the
    SAME size as the rank planes (so that the share the 126 MB L2 serves is the same), measured in this run;
    the rate over a 16 GiB buffer (no L2 share) is reported beside it.  The whole walk phase (kernel + grouping
    and counting of the emitted ranks) is reported as lf_steps_per_s."""
    k_ms = s.get("ms_walk_kernel") or s["ms_walk"]
    steps_s = s["lf_steps"] / (k_ms / 1e3)
    n_launch = max(1, s["n_blocks"] - 1)
    roof = {"kernel": "lf_walk_kernel<3> (LF walk, ranks leave in per-bucket pages)",
            "bound": "hbm", "unit": "GB/s", "achieved": 32.0 * steps_s / 1e9,
            "algorithmic_bytes_per_unit": 32.0, "sequential_bytes_per_unit": 5.25,
            "units": "LF steps (one random 32-byte sector each)", "units_per_step": s["lf_steps"],
            "kernel_ms_per_step": k_ms, "avg_launch_ms": k_ms / n_launch, "lf_steps_per_s_kernel": steps_s,
            "walk_phase_ms_per_step": s["ms_walk"], "working_set_bytes": ws_bytes,
            "hbm_copy_peak": peak, "hbm_copy_peak_source": peak_src,
            "frac_of_copy_peak": 37.25 * steps_s / 1e9 / peak, "traffic": None}
    ceiling = r32_ws or r32
    if ceiling:
        roof["peak"] = ceiling
        roof["peak_source"] = ("independent random 32-byte sector reads over a %d MiB buffer (the size of the block's rank planes), "
                               "full-occupancy grid, measured in this run (psc_bench_random_sectors)" % (ws_bytes >> 20)
                               if r32_ws else "R32 over a 16 GiB buffer, measured in this run (psc_bench_random_sectors)")
        roof["r32_gbs"] = r32
        roof["r32_working_set_gbs"] = r32_ws
        roof["steps_per_s_bound"] = ceiling * 1e9 / 32.0
        roof["frac"] = roof["achieved"] / ceiling
    else:
        roof["peak"] = peak
        roof["peak_source"] = peak_src
        roof["frac"] = roof["achieved"] / peak
    return roof


def radix_roofline(s, peak, peak_src):
    gen_ms, gen_b = s["ms_radix_keygen"], s["radix_keygen_bytes"]
    pair_ms, pair_b = s["ms_radix_pass"] - gen_ms, s["radix_bytes"] - gen_b
    n_gen = 1 if gen_b else 0
    launches = max(1, s["radix_passes"] - n_gen * s["n_blocks"])
    roof = {"kernel": "radix_pass_kernel", "bound": "hbm", "unit": "GB/s",
            "achieved": pair_b / (pair_ms / 1e3) / 1e9 if pair_ms else None,
            "algorithmic_bytes_per_unit": 24.0, "units": "(key64,val32) pairs per digit pass (12 B in + 12 B out)",
            "launches_per_step": launches, "avg_launch_ms": pair_ms / launches,
            "peak": peak, "peak_source": peak_src, "traffic": None}
    if gen_b:
        roof["keygen_pass"] = {"ms": gen_ms, "algorithmic_bytes": gen_b, "GBps": gen_b / (gen_ms / 1e3) / 1e9,
                               "note": "first pass of round 0: 1 B text in, 12 B out per element"}
    if roof["achieved"]:
        roof["frac"] = roof["achieved"] / peak
    return roof, pair_b, launches


def add_traffic(roof, s, pair_b=0, launches=1):
    """DRAM bytes of one launch from the committed ncu --set full capture (profiles/), scaled to the
    algorithmic bytes of an average launch of this run"""
    try:
        name = "r2_traffic.json" if os.path.exists(os.path.join(ROOT, "profiles", "r2_traffic.json")) else "r1_traffic.json"
        tr = json.load(open(os.path.join(ROOT, "profiles", name)))
        key = "lf_walk_kernel" if roof["kernel"].startswith("lf_walk") else "radix_pass_kernel"
        k = tr[key]
        units = s["lf_steps"] / max(1, s["n_blocks"] - 1) if key == "lf_walk_kernel" else pair_b / 24.0 / launches
        roof["traffic"] = k["dram_bytes"] / k["units"] * units
        roof["traffic_source"] = k["source"]
        roof["algorithmic_bytes_per_launch"] = roof["algorithmic_bytes_per_unit"] * units
    except Exception:
        roof["traffic"] = None


def single_block_record(ctx, steps, peak, peak_src):
    """configs[1] (1 GiB random bytes, one block) kept beside the headline so rounds stay comparable"""
    import torch
    text, _, _ = gen_text("random1g", 0)
    n = len(text)
    d_text = torch.from_numpy(text).cuda()
    d_out = torch.empty(5 * n, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        ctx.construct_sa_ptr(d_text.data_ptr(), n, d_out.data_ptr(), 0, device=True)
    st = []
    for _ in range(max(1, steps)):
        ctx.construct_sa_ptr(d_text.data_ptr(), n, d_out.data_ptr(), 0, device=True)
        st.append(ctx.stats())
    ms = sum(x["ms_total"] for x in st) / len(st)
    viol = ctx.verify_sa_ptr(d_text.data_ptr(), n, d_out.data_ptr())
    roof, pair_b, launches = radix_roofline(st[-1], peak, peak_src)
    add_traffic(roof, st[-1], pair_b, launches)
    del d_text, d_out
    torch.cuda.empty_cache()
    return {"workload": "random1g", "text_bytes": n, "value": n / 1e6 / (ms / 1e3), "unit": "MB/s", "ms_per_step": ms,
            "steps": len(st), "verified": not any(viol), "roofline": roof}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--size", type=int, default=0)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=64 << 20, help="bytes of the workload the cpu_baseline leg runs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-single-block", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    gen, kw, dsize, mb_default, desc = WORKLOADS[a.workload]
    n = a.size or dsize
    config = {"workload": a.workload, "description": desc, "text_bytes": n,
              "l2": "inputs larger than L2 (text + 8 B/suffix of block SA and gap arrays >> 126 MB)"}

    if a.impl == "reference":
        if rank == 0:
            reference_arm(a, config, n)
        return True

    import numpy as np
    import torch
    import psascan_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    text, max_block, source = gen_text(a.workload, n)
    config["text_source"] = source
    ctx = psascan_b200.Context(local)
    sharded = world > 1 and max_block > 0   # multi-block workload: csrc/dist.cu over NCCL (strong scaling)
    if world > 1 and not sharded:
        raise SystemExit(f"workload {a.workload} is a single block and does not shard; use --workload {DEFAULT_WORKLOAD}")
    host_text = torch.from_numpy(text).pin_memory()
    dev_text = host_text.cuda()
    if sharded:
        from psascan_b200 import dist as pdist
        dist.broadcast(dev_text, src=0)   # every rank sorts rank 0's bytes
        pdist.init(ctx)                   # NCCL communicator inside the library (csrc/comm.cu)
        capn = pdist.slice_capacity(ctx, n, world)
    else:
        capn = n
    host_out = torch.empty(5 * capn, dtype=torch.uint8).pin_memory()
    dev_out = torch.empty(5 * capn, dtype=torch.uint8, device="cuda")
    share = [0, n]

    def step_dev():
        if sharded:   # every rank ends up with its n/world slice of the SA (merge sharded by output range)
            share[0], share[1], _ = pdist.construct_sa_slices(ctx, dev_text.data_ptr(), n, dev_out.data_ptr(), max_block)
        else:
            ctx.construct_sa_ptr(dev_text.data_ptr(), n, dev_out.data_ptr(), max_block, device=True)
        return ctx.stats()

    def step_e2e():
        if sharded:   # the text crosses PCIe once in total: every rank uploads 1/N of it, NVLink all-gather completes it
            if n % world == 0:
                per = n // world
                dev_text[rank * per:(rank + 1) * per].copy_(host_text[rank * per:(rank + 1) * per], non_blocking=True)
                dist.all_gather_into_tensor(dev_text, dev_text[rank * per:(rank + 1) * per])
            else:
                dev_text.copy_(host_text, non_blocking=True)
            torch.cuda.synchronize()
            _, ln, _ = pdist.construct_sa_slices(ctx, dev_text.data_ptr(), n, dev_out.data_ptr(), max_block)
            host_out[:5 * ln].copy_(dev_out[:5 * ln], non_blocking=True)
            torch.cuda.synchronize()
        else:
            ctx.construct_sa_ptr(host_text.data_ptr(), n, host_out.data_ptr(), max_block, device=False)
        return ctx.stats()

    for _ in range(a.warmup):
        step_dev()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t0 = time.perf_counter()
    stats = [step_dev() for _ in range(a.steps)]
    barrier()
    wall = time.perf_counter() - t0
    dev_ms = sum(s["ms_total"] for s in stats)
    comm = dict(pdist.comm_stats(ctx), nranks=world, backend="NCCL inside the library (csrc/comm.cu)") if sharded else None
    # end to end through the host-buffer entry point
    step_e2e()
    barrier()
    t1 = time.perf_counter()
    for _ in range(a.steps):
        step_e2e()
    barrier()
    e2e_wall = time.perf_counter() - t1
    clocks = sampler.stop()

    tt = torch.tensor([dev_ms, wall * 1e3, e2e_wall * 1e3], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms = [float(x) for x in tt.tolist()]

    # ---- verification of what the timed loops produced (untimed): e2e result and device result
    step_dev()
    if sharded:
        # every rank checks its own slice (range + strict suffix order of adjacent entries, psc_verify_sa_slice_d);
        # the pair across a slice boundary needs the first entry of the next slice
        first = torch.zeros(8, dtype=torch.uint8, device="cuda")
        if share[1] > 0:
            first[:5] = dev_out[:5]
        firsts = [torch.zeros(8, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(firsts, first)
        lens = [pdist.share_begin(h + 1, n, world) - pdist.share_begin(h, n, world) for h in range(world)]
        nxt = -1
        for h in range(rank + 1, world):
            if lens[h] > 0:
                nxt = int.from_bytes(bytes(firsts[h][:5].cpu().tolist()), "little")
                break
        v2 = ctx.verify_sa_slice_ptr(dev_text.data_ptr(), n, dev_out.data_ptr(), share[0], share[1], nxt)
        viol = [v2[0], 0, 0, v2[1]]
        e2e_same = bool(torch.equal(host_out[:5 * share[1]], dev_out[:5 * share[1]].cpu()))
    else:
        viol = ctx.verify_sa_ptr(dev_text.data_ptr(), n, dev_out.data_ptr())
        e2e_same = bool(torch.equal(host_out, dev_out.cpu())) if n <= (1 << 32) else True
    flag = torch.tensor([0 if (any(viol) or not e2e_same) else 1], device="cuda")
    if dist:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    verified = bool(flag.item())

    if rank == 0:
        s = stats[-1]
        value = n * a.steps / 1e6 / (dev_ms / 1e3)
        e2e = n * a.steps / 1e6 / (e2e_ms / 1e3)
        peak, peak_src = measured_peaks()
        del dev_out, host_out
        torch.cuda.empty_cache()
        line = {"metric": METRIC, "value": value, "unit": "MB/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": dev_ms / a.steps, "higher_is_better": True,
                "scaling": "strong" if sharded else "weak", "vs_baseline": None,
                "dtype": "u8 text, u64 keys, u32 ranks", "data": "synthetic",
                "config": dict(config, parallelism=("block sorts per GPU + sharded tail walk + ncclReduceScatter of gap counts + ONE exchange of SA/gap slices, "
                                                    "merge by output range; every rank keeps n/N of the SA"
                                                    if sharded else "single GPU"), max_block=max_block, n_blocks=s["n_blocks"]),
                "clocks": clocks, "wall_ms_per_step": wall_ms / a.steps,
                "e2e": {"value": e2e, "unit": "MB/s", "h2d_bytes_per_step": n if (not sharded or n % world == 0) else n * world,
                        "d2h_bytes_per_step": 5 * n, "ms_per_step": e2e_ms / a.steps},
                "gpu_launches": s["kernel_launches"] * a.steps, "comm_per_step": comm,
                "verified": verified, "verify": {"violations": viol, "e2e_equals_device_result": e2e_same,
                                                 "how": ("psc_verify_sa_slice_d on every rank: entries in range + adjacent suffixes in strictly increasing order"
                                                         if sharded else "psc_verify_sa_d: permutation + first-symbol order + tie order via ISA")},
                "phases_ms": {k: s[k] for k in s if k.startswith("ms_")},
                "counters": {k: s[k] for k in s if not k.startswith("ms_")}}
        if s["lf_steps"] > 0:
            r32 = r32_ws = None
            # the walk's working set: the rank planes of one block, 32 bytes per 224 positions and symbol
            sigma = int(np.count_nonzero(np.bincount(text[:min(n, 1 << 24)], minlength=256)))
            ws_bytes = sigma * (min(n, max_block) // 224 + 1) * 32
            try:   # random 32-byte sector rate of this GPU (SURVEY 8-d "R32"), the walk's ceiling
                r32 = ctx.bench_random_sectors(16 << 30, 128)
                r32_ws = ctx.bench_random_sectors(max(ws_bytes, 1 << 20), 128)
            except Exception as e:
                line["r32_error"] = str(e)[:100]
            # LF steps of the whole job per second of its walk phase (at N > 1: rank 0's share, same wall time)
            total_steps = sum(max(0, n - min(n, (i + 1) * max_block)) for i in range(s["n_blocks"]))
            roof = walk_roofline(s, r32, r32_ws, ws_bytes, peak, peak_src)
            add_traffic(roof, s)
            line["lf_steps_per_s"] = total_steps / (s["ms_walk"] / 1e3)
            line["lf_steps_per_step"] = total_steps
            rr, pair_b, launches = radix_roofline(s, peak, peak_src)
            line["roofline_radix_pass"] = rr
        else:
            roof, pair_b, launches = radix_roofline(s, peak, peak_src)
            add_traffic(roof, s, pair_b, launches)
        line["roofline"] = roof
        if world == 1 and not a.no_single_block and a.workload == DEFAULT_WORKLOAD:
            try:
                line["single_block"] = single_block_record(ctx, min(a.steps, 3), peak, peak_src)
                verified = verified and line["single_block"]["verified"]
                line["verified"] = verified
            except Exception as e:
                line["single_block"] = {"error": str(e)[:200]}
        if world == 1 and not a.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_sample_baseline(text, s["n_blocks"], a.cpu_sample)
            except Exception as e:  # the reference binary must travel with the repo
                line["cpu_baseline"] = {"value": None, "error": str(e)[:200]}
        print(json.dumps(line), flush=True)
    if sharded:
        pdist.finalize(ctx)
    ctx.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    return verified


if __name__ == "__main__":
    sys.exit(0 if main() else 1)
