"""one multi-block construction on a Markov text (ncu target): size, max_block from argv"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import psascan_b200 as P
from psascan_b200 import textgen
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 29
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 27
c = P.Context(0)
t = textgen.markov_torch(n, seed=3, device="cuda")
out = torch.empty(5 * n, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
c.construct_sa_ptr(t.data_ptr(), n, out.data_ptr(), max_block=mb, device=True)
print(c.stats())
