"""python tools/run_config4.py [log2_bytes=31] [log2_block=28]
BASELINE configs[3]: highly repetitive text (mutated copies of a 1 MiB seed over a-z, substitution rate 1e-3,
seed 4) in several blocks on one GPU; prints stats (PSC_TRACE=1 adds the per-kernel breakdown) and verifies."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psascan_b200 as P  # noqa: E402
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 31
lgb = int(sys.argv[2]) if len(sys.argv) > 2 else 28
n, mb = 1 << lg, 1 << lgb
g = torch.Generator(device="cuda"); g.manual_seed(4)
seed = torch.randint(97, 123, (1 << 20,), dtype=torch.uint8, device="cuda", generator=g)
text = seed.repeat(n >> 20)[:n].contiguous()
nmut = n // 1000
pos = torch.randint(0, n, (nmut,), device="cuda", generator=g)
text[pos] = torch.randint(97, 123, (nmut,), dtype=torch.uint8, device="cuda", generator=g)
out = torch.empty(5 * n, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
c = P.Context(0)
res = []
for it in range(2):
    t0 = time.time()
    c.construct_sa_ptr(text.data_ptr(), n, out.data_ptr(), mb, device=True)
    res.append((time.time() - t0, c.stats()))
st = res[-1][1]
viol = c.verify_sa_ptr(text.data_ptr(), n, out.data_ptr()) if n < (1 << 32) - 1 else None
print(json.dumps({"workload": "configs[3]: mutated copies of a 1 MiB seed, rate 1e-3", "text_bytes": n, "max_block": mb,
                  "n_blocks": st["n_blocks"], "ms_cold": res[0][1]["ms_total"], "ms": st["ms_total"], "MB_per_s": n / 1e3 / st["ms_total"],
                  "verified": viol == [0, 0, 0, 0], "doubling_rounds": st["doubling_rounds"], "radix_passes": st["radix_passes"],
                  "radix_elems_over_n": st["radix_elems"] / n,
                  "phases_ms": {k: round(v, 1) for k, v in st.items() if k.startswith("ms_") and v}}))
