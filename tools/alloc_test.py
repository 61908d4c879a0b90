"""how long does the driver's stream-ordered pool take per allocation as the footprint grows?"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, ctypes as C
import psascan_b200 as P
c = P.Context(0)
torch.cuda.init(); torch.cuda.synchronize()
held = []
for k in range(9):
    gb = 16
    p = C.c_void_p()
    t0 = time.time(); rc = c.lib.psc_device_alloc(c.h, C.c_size_t(gb << 30), C.byref(p)); t1 = time.time()
    torch.cuda.synchronize(); t2 = time.time()
    print(f"alloc #{k} 16 GB: call {1e3*(t1-t0):.1f} ms, +sync {1e3*(t2-t1):.1f} ms, rc {rc}", flush=True)
    if rc != 0: break
    held.append(p)
for p in held[:4]:
    c.lib.psc_device_free(c.h, p)
torch.cuda.synchronize()
for sz in (8, 24, 12, 17):
    p = C.c_void_p()
    t0 = time.time(); rc = c.lib.psc_device_alloc(c.h, C.c_size_t(sz << 30), C.byref(p)); t1 = time.time()
    torch.cuda.synchronize(); t2 = time.time()
    print(f"after freeing 64 GB: alloc {sz} GB: call {1e3*(t1-t0):.1f} ms, +sync {1e3*(t2-t1):.1f} ms, rc {rc}", flush=True)
    c.lib.psc_device_free(c.h, p)
print("--- with the memory touched between allocations")
from psascan_b200.dist import tensor_from_pointer
held2 = []
for k in range(7):
    p = C.c_void_p()
    t0 = time.time(); rc = c.lib.psc_device_alloc(c.h, C.c_size_t(16 << 30), C.byref(p)); t1 = time.time()
    if rc != 0: print("rc", rc); break
    t = tensor_from_pointer(p.value, 16 << 30, "u1", True)
    t.zero_(); torch.cuda.synchronize(); t2 = time.time()
    print(f"alloc #{k} 16 GB: call {1e3*(t1-t0):.1f} ms, zero {1e3*(t2-t1):.1f} ms", flush=True)
    held2.append(p)
