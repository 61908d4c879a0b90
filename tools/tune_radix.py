"""Times one radix digit pass for every kernel variant: (key64, val32) pairs (24 B per element of
algorithmic traffic) or, with --keys, keys only (16 B)."""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = [a for a in sys.argv[1:] if a != "--keys"]
keys = "--keys" in sys.argv
VARIANTS = args[1:] or ["512x12b2k2c4", "512x12b2k3c4", "384x12b3k3c4", "384x12b3k4c4", "512x8b3k3c4", "512x16b2k2c4",
                        "256x12b4k5c4", "384x16b2k3c4", "256x16b3k4c4", "384x8b4k4c4", "256x16b3k5c4", "512x16b2k3c4",
                        "512x12b2k2c2", "512x12b2k2c3", "512x12b2k2c6", "512x12b2k2c12", "512x16b2k2c8"]
n = int(args[0]) if args else 1 << 28
code = ("import sys; sys.path.insert(0, %r); import psascan_b200 as P; c = P.Context(0); "
        "ms = c.bench_radix_pass(%d, 8); print('%%.4f' %% ms)" % (ROOT, n))
for v in VARIANTS:
    env = dict(os.environ)
    env["PSC_RADIX_VARIANT_K" if keys else "PSC_RADIX_VARIANT"] = v
    if keys:
        env["PSC_BENCH_KEYS_ONLY"] = "1"
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    try:
        ms = float(r.stdout.strip().splitlines()[-1])
        print(json.dumps({"variant": v, "keys_only": keys, "n": n, "ms_per_pass": ms,
                          "GBps": (16.0 if keys else 24.0) * n / ms / 1e6}))
    except Exception:
        print(json.dumps({"variant": v, "error": (r.stderr or r.stdout)[-300:]}))
