import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import psascan_b200 as P
c = P.Context(0)
gb = int(os.environ.get("R32_GB", "8"))
print("variant", os.environ.get("PSC_R32_VARIANT"), gb, "GiB:", round(c.bench_random_sectors(gb << 30, 256), 1), "GB/s", flush=True)
