"""large-input smoke: text generated on the GPU, automatic block size, sampled order check"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import psascan_b200 as P
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3 << 30
sigma = int(sys.argv[2]) if len(sys.argv) > 2 else 4
mb = int(sys.argv[3]) if len(sys.argv) > 3 else 0
torch.manual_seed(5)
c = P.Context(0)
t = torch.randint(0, sigma, (n,), dtype=torch.uint8, device="cuda")
out = torch.empty(5 * n, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
print("auto max block:", c.lib.psc_max_block_size(c.h, n) if mb == 0 else mb, flush=True)
t0 = time.time()
c.construct_sa_ptr(t.data_ptr(), n, out.data_ptr(), max_block=mb, device=True)
torch.cuda.synchronize()
st = c.stats()
print("n", n, "sigma", sigma, "blocks", st["n_blocks"], "ms", round(st["ms_total"], 1), "MB/s", round(n / 1e3 / st["ms_total"], 1),
      {k: round(v, 1) for k, v in st.items() if k.startswith("ms_") and v > 0}, flush=True)
# sampled check: 1 M random adjacent pairs must be in suffix order (first 48 symbols decide for random text)
idx = torch.randint(0, n - 1, (1 << 20,), device="cuda")
o5 = out.view(n, 5)
def sa_at(i):
    r = o5[i].to(torch.int64)
    return r[:, 0] | (r[:, 1] << 8) | (r[:, 2] << 16) | (r[:, 3] << 24) | (r[:, 4] << 32)
a, b = sa_at(idx), sa_at(idx + 1)
L = 48
off = torch.arange(L, device="cuda")
pa = (a[:, None] + off).clamp(max=n - 1)
pb = (b[:, None] + off).clamp(max=n - 1)
ta = torch.where((a[:, None] + off) < n, t[pa].to(torch.int16), torch.full_like(pa, -1, dtype=torch.int16))
tb = torch.where((b[:, None] + off) < n, t[pb].to(torch.int16), torch.full_like(pb, -1, dtype=torch.int16))
d = (ta != tb)
first = torch.where(d.any(1), d.float().argmax(1), torch.full((len(idx),), L - 1, device="cuda"))
va = ta.gather(1, first[:, None])[:, 0]
vb = tb.gather(1, first[:, None])[:, 0]
bad = int(((va > vb) & d.any(1)).sum())
undecided = int((~d.any(1)).sum())
print("order violations", bad, "undecided", undecided, "of", len(idx), flush=True)
assert bad == 0
