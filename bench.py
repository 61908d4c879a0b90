#!/usr/bin/env python
"""bench.py -- SA construction throughput of the B200 hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--size BYTES]
                  [--impl reference]

A step = one SA construction over one synthetic text (text -> 40-bit .sa5 entries).
  value : input MB/s with the text and the output resident in HBM (device time, CUDA events
          recorded by the library on its own stream around the whole construction)
  e2e   : the same through the host-buffer C-ABI call psc_construct_sa (pinned host text in,
          .sa5 bytes out to pinned host memory; both copies inside the timed region)
N > 1: one process per GPU (torchrun), each rank builds the SA of its own text of the same shape
(the single-block configuration does not shard: "replicas", weak scaling).
--impl reference times the unmodified reference construct_sa (oracle/_ref) on the host cores.
"""
import argparse
import json
import os
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
    "random1g": ("random", dict(seed=2), 1 << 30, 0,
                 "configs[1]: 1 GiB uniform-random bytes (sigma=255), single block in HBM"),
    "dna200m": ("dna", dict(seed=1), 200_000_000, 0, "configs[0]: 200 MB DNA-like (sigma=4), single block"),
    "markov_multi": ("markov", dict(seed=3), 1 << 30, 1 << 27,
                     "configs[2] scaled 1/16: 1 GiB order-5 Markov text, 8 blocks of 128 MiB, gap streaming"),
    "repetitive_multi": ("repetitive", dict(seed=4), 1 << 28, 1 << 26,
                         "configs[3] scaled: mutated copies of a 1 MiB seed, 4 blocks"),
}


def gen_text(workload, size, rank):
    import numpy as np
    from psascan_b200 import textgen
    gen, kw, dsize, mb, _ = WORKLOADS[workload]
    n = size or dsize
    kw = dict(kw)
    kw["seed"] = kw.get("seed", 0) + 1000 * rank
    if gen == "random" and n >= (1 << 26):
        # fast path for big uniform texts: raw generator bytes folded into 0..254
        rng = np.random.default_rng(kw["seed"])
        t = rng.integers(0, 255, size=n, dtype=np.uint8)
        return t, (mb if mb < n else 0)
    if gen == "markov" and n >= (1 << 26):
        import torch
        if torch.cuda.is_available():   # GiB-sized Markov texts are generated on the GPU (same model)
            return textgen.markov_torch(n, seed=kw["seed"]).cpu().numpy(), (mb if mb < n else 0)
    return textgen.GENERATORS[gen](n, **kw), (mb if mb < n else 0)


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


def run_reference_sample(text, threads, mem="8gi"):
    """times oracle/_ref/construct_sa (the unmodified reference, libsais variant) on `text`."""
    import oracle_lib
    d = tempfile.mkdtemp(prefix="psc_bench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    f = os.path.join(d, "t.txt")
    text.tofile(f)
    out = f + ".sa5"
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    t0 = time.perf_counter()
    subprocess.run([oracle_lib.REF_BIN, f, "-o", out, "-m", mem], check=True, env=env,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, stdin=subprocess.DEVNULL)
    dt = time.perf_counter() - t0
    for x in os.listdir(d):
        os.remove(os.path.join(d, x))
    os.rmdir(d)
    return dt


def host_threads():
    c = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    return max(1, min(c, 128))  # 14 MiB of budget per thread (psascan.hpp:73-78); 128 keeps -m 8gi valid


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="random1g", choices=sorted(WORKLOADS))
    ap.add_argument("--size", type=int, default=0)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=64 << 20, help="bytes of the workload the CPU baseline runs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    gen, kw, dsize, _, desc = WORKLOADS[a.workload]
    n = a.size or dsize
    config = {"workload": a.workload, "description": desc, "text_bytes": n,
              "l2": "inputs larger than L2 (text + 12 B/suffix sort state >> 126 MB)"}

    if a.impl == "reference":
        if rank != 0:
            return
        thr = host_threads()
        sample_n = min(n, a.cpu_sample)
        text, _ = gen_text(a.workload, sample_n, 0)
        for _ in range(min(a.warmup, 1)):
            run_reference_sample(text, thr)
        ts = [run_reference_sample(text, thr) for _ in range(a.steps)]
        sec = sum(ts) / len(ts)
        v = sample_n / 1e6 / sec
        cb = {"value": v, "unit": "MB/s", "cores": thr, "kind": "reference",
              "sample": f"first {sample_n} bytes of the workload, construct_sa -m 8gi (libsais build), tmpfs"}
        print(json.dumps({"impl": "reference", "metric": "SA construction input MB/s", "value": v, "unit": "MB/s",
                          "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": sec * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64 keys",
                          "data": "synthetic", "config": config, "cpu_baseline": cb,
                          "e2e": {"value": v, "unit": "MB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return

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

    text, max_block = gen_text(a.workload, n, rank)
    ctx = psascan_b200.Context(local)
    host_text = torch.from_numpy(text).pin_memory()
    host_out = torch.empty(5 * n, dtype=torch.uint8).pin_memory()
    dev_text = host_text.cuda()
    dev_out = torch.empty(5 * n, dtype=torch.uint8, device="cuda")

    sharded = world > 1 and max_block > 0   # multi-block workload: csrc/dist.cu over NCCL (strong scaling)
    if sharded:
        from psascan_b200 import dist as pdist
        if rank != 0:   # every rank must hold the same text
            text, _ = gen_text(a.workload, n, 0)
            host_text.copy_(torch.from_numpy(text))
            dev_text.copy_(host_text)

    def step_dev():
        if sharded:   # every rank ends up with its n/world slice of the SA (merge sharded by output range)
            pdist.construct_sa_slices(ctx, dev_text.data_ptr(), n, dev_out.data_ptr(), max_block)
        else:
            ctx.construct_sa_ptr(dev_text.data_ptr(), n, dev_out.data_ptr(), max_block, device=True)
        return ctx.stats()

    def step_e2e():
        if sharded:   # host text -> HBM on every rank, sharded construction, every rank's slice back to its host
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
    ok = True
    # cheap integrity check of the last result: a permutation checksum
    sa_sum = int(psascan_b200.lib.decode_sa5(host_out.numpy()[: 5 * min(n, 1 << 20)]).sum()) if n else 0
    if rank == 0:
        s = stats[-1]
        total_bytes = n if sharded else n * world
        value = total_bytes * a.steps / 1e6 / (dev_ms / 1e3)
        e2e = total_bytes * a.steps / 1e6 / (e2e_ms / 1e3)
        peak, peak_src = measured_peaks()
        if s["lf_steps"] > 0 and s["ms_walk"] > s["ms_radix_pass"]:
            roof = {"kernel": "lf_walk_kernel", "bound": "hbm", "unit": "GB/s",
                    "achieved": 41.25 * s["lf_steps"] / (s["ms_walk"] / 1e3) / 1e9,
                    "algorithmic_bytes_per_unit": 41.25, "units": "LF steps", "peak": peak, "peak_source": peak_src,
                    "lf_steps_per_s": s["lf_steps"] / (s["ms_walk"] / 1e3), "traffic": None}
            try:   # random 32-byte sector rate of this GPU (SURVEY 8-d "R32"), the walk's real ceiling
                r32 = ctx.bench_random_sectors(8 << 30, 128)
                roof["r32_gbs"] = r32
                roof["frac_of_r32"] = roof["achieved"] / r32
            except Exception as e:
                roof["r32_error"] = str(e)[:100]
        else:
            # dominant kernel: the digit pass over (key64, val32) pairs; the one pass per block sort that builds
            # its keys from the text (8-bit codes) moves fewer bytes and is reported beside it
            gen_ms, gen_b = s["ms_radix_keygen"], s["radix_keygen_bytes"]
            pair_ms, pair_b = s["ms_radix_pass"] - gen_ms, s["radix_bytes"] - gen_b
            n_gen = 1 if gen_b else 0
            roof = {"kernel": "radix_pass_kernel", "bound": "hbm", "unit": "GB/s",
                    "achieved": pair_b / (pair_ms / 1e3) / 1e9 if pair_ms else None,
                    "algorithmic_bytes_per_unit": 24.0, "units": "(key64,val32) pairs per digit pass (12 B in + 12 B out)",
                    "launches_per_step": s["radix_passes"] - n_gen * s["n_blocks"],
                    "avg_launch_ms": pair_ms / max(1, s["radix_passes"] - n_gen * s["n_blocks"]),
                    "peak": peak, "peak_source": peak_src, "traffic": None}
            if gen_b:
                roof["keygen_pass"] = {"ms": gen_ms, "algorithmic_bytes": gen_b, "GBps": gen_b / (gen_ms / 1e3) / 1e9,
                                       "note": "first pass of round 0: 1 B text in, 12 B out per element"}
        # DRAM bytes of one launch from the committed ncu --set full capture (profiles/), scaled to
        # the algorithmic bytes of an average launch of this run
        try:
            tr = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_traffic.json")))
            k = tr[roof["kernel"]]
            if roof["kernel"] == "lf_walk_kernel":
                units = s["lf_steps"] / max(1, s["n_blocks"] - 1)
            else:
                units = pair_b / 24.0 / max(1, roof["launches_per_step"])
            roof["traffic"] = k["dram_bytes"] / k["units"] * units
            roof["traffic_source"] = k["source"]
            roof["algorithmic_bytes_per_launch"] = roof["algorithmic_bytes_per_unit"] * units
        except Exception:
            roof["traffic"] = None
        if roof["achieved"]:
            roof["frac"] = roof["achieved"] / peak
        line = {"metric": "SA construction input MB/s", "value": value, "unit": "MB/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": dev_ms / a.steps, "higher_is_better": True,
                "scaling": "strong" if sharded else "weak", "vs_baseline": None,
                "dtype": "u8 text, u64 keys, u32 ranks", "data": "synthetic",
                "config": dict(config, parallelism=("block sorts per GPU + sharded tail walk + NCCL all-reduce of gap counts + merge sharded by "
                                                    "output range (one all-to-all per level); every rank keeps n/N of the SA"
                                                    if sharded else "replicas" if world > 1 else "single GPU"),
                               max_block=max_block),
                "clocks": clocks, "wall_ms_per_step": wall_ms / a.steps,
                "e2e": {"value": e2e, "unit": "MB/s", "h2d_bytes_per_step": n * world if sharded else n,
                        "d2h_bytes_per_step": 5 * n,
                        "ms_per_step": e2e_ms / a.steps},
                "gpu_launches": s["kernel_launches"] * a.steps, "roofline": roof,
                "phases_ms": {k: s[k] for k in s if k.startswith("ms_")},
                "counters": {k: s[k] for k in s if not k.startswith("ms_")}, "sa_checksum_1Mi": sa_sum}
        if not a.no_cpu_baseline:
            thr = host_threads()
            sample_n = min(n, a.cpu_sample)
            try:
                sec = run_reference_sample(text[:sample_n], thr)
                line["cpu_baseline"] = {"value": sample_n / 1e6 / sec, "unit": "MB/s", "cores": thr, "kind": "reference",
                                        "sample": f"first {sample_n} bytes of the workload, construct_sa -m 8gi (libsais build), tmpfs"}
            except Exception as e:  # the reference binary must travel with the repo
                line["cpu_baseline"] = {"value": None, "error": str(e)[:200]}
        print(json.dumps(line))
    ctx.close()
    if dist:
        dist.destroy_process_group()
    return ok


if __name__ == "__main__":
    main()
