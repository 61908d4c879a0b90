"""Times one radix digit pass (24 B/pair of algorithmic traffic) for every kernel variant."""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VARIANTS = sys.argv[2:] or ["512x12b2m0", "512x12b2m4", "512x12b2m6", "512x12b2m8", "512x12b2m12", "384x12b3m0",
                            "384x12b3m6", "384x12b3m8", "512x8b3m0", "512x8b3m4", "512x16b2m8", "512x16b2m10",
                            "256x12b4m6", "256x16b3m8", "1024x8b1m4", "1024x12b1m8"]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 28
code = ("import sys; sys.path.insert(0, %r); import psascan_b200 as P; c = P.Context(0); "
        "ms = c.bench_radix_pass(%d, 8); print('%%.4f' %% ms)" % (ROOT, n))
for v in VARIANTS:
    env = dict(os.environ, PSC_RADIX_VARIANT=v)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    try:
        ms = float(r.stdout.strip().splitlines()[-1])
        print(json.dumps({"variant": v, "n": n, "ms_per_pass": ms, "GBps": 24.0 * n / ms / 1e6}))
    except Exception:
        print(json.dumps({"variant": v, "error": (r.stderr or r.stdout)[-300:]}))
