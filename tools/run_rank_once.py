"""builds the rank structure over a random BWT (ncu / timing target): size and alphabet from argv"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import psascan_b200 as P
m = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 26
sigma = int(sys.argv[2]) if len(sys.argv) > 2 else 255
bwt = np.random.default_rng(1).integers(0, sigma, size=m, dtype=np.uint8)
bwt[12345] = 0
c = P.Context(0)
for _ in range(3):
    t0 = time.perf_counter()
    r = c.rank_build(bwt, 12345)
    dt = time.perf_counter() - t0
    r.free()
print("m", m, "sigma", sigma, "last build incl. H2D of the BWT: %.1f ms" % (dt * 1e3))
