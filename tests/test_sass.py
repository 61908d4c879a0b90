"""The shipped library holds the instructions the design relies on (static check of the SASS, no GPU needed):
the rank query is ONE 256-bit load with a 64-byte L2 fetch, the paged walk finds bucket peers with MATCH.ANY,
the digit pass prefetches its next tile into L2 with a bulk-async prefetch, the merge descent is branch + popcount.
A compiler or source change that silently splits the sector load or drops the prefetch shows up here, not only in
the next profile."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "psascan_b200", "libpsascan_cuda.so")


@pytest.fixture(scope="module")
def sass():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe) or not os.path.exists(LIB):
        pytest.skip("cuobjdump or the built library is missing")
    out = subprocess.run([exe, "-sass", LIB], check=True, capture_output=True, text=True).stdout
    funcs, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur:
            funcs[cur].append(line)
    return {k: "\n".join(v) for k, v in funcs.items()}


def body(sass, key):
    hits = [v for k, v in sass.items() if key in k]
    assert hits, key
    return hits[0]


def test_built_for_sm_100a():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe) or not os.path.exists(LIB):
        pytest.skip("cuobjdump or the built library is missing")
    out = subprocess.run([exe, "-lelf", LIB], check=True, capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_walk_kernel_sector_load_and_match(sass):
    b = body(sass, "lf_walk_kernelILi3E")
    assert len(re.findall(r"LDG\.E\S*\.256", b)) == 1          # one sector per LF step, one instruction
    assert "LTC64B" in b                                         # 64-byte L2 fetch, not 128
    assert "MATCH.ANY" in b                                      # lanes of one bucket
    assert "ATOMG.E.ADD.64" in b                                 # pages handed out in batches
    assert "LDL" not in b and "STL" not in b                     # nothing spilled
    for mode in ("lf_walk_kernelILi0E", "lf_walk_kernelILi2E"):
        assert len(re.findall(r"LDG\.E\S*\.256", body(sass, mode))) == 1


def test_radix_pass_prefetch_and_votes(sass):
    b = body(sass, "radix_pass_kernelILi512ELi12ELi2ELb1ELb0ELi1E")
    assert "UBLKPF" in b                                         # cp.async.bulk.prefetch.L2 of the next tile
    assert b.count("VOTE") >= 64                                 # ballot-matched ranks, 8 digit bits per item
    assert "LDL" not in b and "STL" not in b


def test_merge_descent_is_register_only(sass):
    b = body(sass, "descent_kernel")
    assert "POPC" in b and "LDS" not in b and "BAR" not in b
    assert "LDL" not in b and "STL" not in b
