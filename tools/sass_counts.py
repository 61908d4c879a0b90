"""SASS instruction mix of the shipped hot kernels (cuobjdump -sass of the built objects): which memory, vote and
bulk-copy instructions the compiler emitted.  python tools/sass_counts.py > profiles/<round>_sass_counts.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "psascan_b200", "csrc", "build")
KERNELS = [
    ("radix.o", "radix_pass_kernelILi512ELi12ELi2ELb1ELb0ELi1E", "radix_pass_kernel<512,12,2,pairs,no keygen,1> (the digit pass of the bench)"),
    ("radix.o", "radix_pass_kernelILi512ELi12ELi2ELb1ELb1ELi1E", "radix_pass_kernel<512,12,2,pairs,keygen,1> (first pass of round 0 for 8-bit codes)"),
    ("walk.o", "lf_walk_kernelILi3E", "lf_walk_kernel<3> (LF walk, paged rank output)"),
    ("walk.o", "page_accumulate_kernel", "page_accumulate_kernel"),
    ("finalmerge.o", "descent_kernel", "descent_kernel (final merge)"),
    ("finalmerge.o", "level_bits_kernel", "level_bits_kernel (final merge)"),
]
WATCH = ["LDG", "STG", "LDS", "STS", "LDSM", "ATOMS", "ATOMG", "ATOM", "RED", "REDG", "REDUX", "VOTE", "VOTEU", "MATCH", "SHFL", "POPC", "BAR",
         "UBLKCP", "UBLKPF", "SYNCS", "LDGSTS", "LDGDEPBAR", "CCTL", "BRA", "LOP3", "IADD3", "IMAD", "ISETP", "SHF", "PRMT", "FLO", "BREV"]

def sass(obj):
    return subprocess.run(["cuobjdump", "-sass", os.path.join(BUILD, obj)], check=True, capture_output=True, text=True).stdout

cache = {}
print("# SASS instruction mix of the shipped kernels (sm_100a, `cuobjdump -sass psascan_b200/csrc/build/*.o`)\n")
print("Static counts per kernel body (not executed counts). `UBLKPF` = `cp.async.bulk.prefetch.L2`, `UBLKCP` = `cp.async.bulk`")
print("global->shared, `SYNCS` = mbarrier operations, `LDG.E.ENL2.256` = the 32-byte sector load of the rank query.\n")
for obj, key, title in KERNELS:
    text = cache.setdefault(obj, sass(obj))
    cur, body = None, []
    for line in text.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        if cur and key in cur:
            body.append(line)
    ops = collections.Counter()
    wide = collections.Counter()
    for line in body:
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Za-z0-9_]+)*)", line)
        if not m:
            continue
        ops[m.group(1)] += 1
        if m.group(1) in ("LDG", "STG", "LDS", "STS", "RED", "REDG", "ATOMG", "ATOMS", "MATCH", "UBLKPF", "UBLKCP"):
            wide[m.group(1) + m.group(2)] += 1
    total = sum(ops.values())
    print(f"## {title}\n")
    print(f"{total} instructions. " + ", ".join(f"{k} {ops[k]}" for k in WATCH if ops.get(k)) + ".\n")
    print("Memory instructions by form: " + ", ".join(f"`{k}` {v}" for k, v in sorted(wide.items(), key=lambda x: -x[1])[:14]) + ".\n")
