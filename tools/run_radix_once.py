import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import psascan_b200 as P
c = P.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 27
print(c.bench_radix_pass(n, 3))
