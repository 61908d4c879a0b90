"""Generates tests/golden/*.npz from the UNMODIFIED reference binary oracle/_ref/construct_sa
(built by oracle/Makefile from /root/reference).  Run here (needs /root/reference built once):
    python tests/golden/make_golden.py
Each case stores the seeded generator call, the reference's -m / thread setting and the .sa5
bytes it wrote; the text itself is regenerated from the seed by the tests."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib  # noqa: E402
from psascan_b200 import textgen  # noqa: E402

def BLK(b, threads=2):
    """-m value that makes the reference use blocks of about b symbols:
    max_block_size = (ram - 14 MiB * threads) / 5.2   (include/psascan.hpp:73-91)"""
    return str(14 * (1 << 20) * threads + int(5.2 * b) + 8)


# name, generator, kwargs, n, reference -m (None = default single block), threads
CASES = [
    ("dna_5000_1blk", "dna", dict(seed=1), 5000, None, 2),
    ("dna_20000_multi", "dna", dict(seed=11), 20000, BLK(6000), 2),      # forces several blocks
    ("random_8000_1blk", "random", dict(seed=2), 8000, None, 2),
    ("random_8000_multi", "random", dict(seed=2), 8000, BLK(3000), 2),
    ("markov_12000_multi", "markov", dict(seed=3), 12000, BLK(3500), 2),
    ("repetitive_16000_multi", "repetitive", dict(seed=4, seed_len=1000, rate=2e-3), 16000, BLK(5000), 2),
    ("periodic_9000_multi", "periodic", dict(period=97, sigma=3, seed=7), 9000, BLK(2500), 2),
    ("binary_3000_1thr", "random", dict(seed=9, sigma=2), 3000, None, 1),
]


def main():
    for name, gen, kw, n, mem, thr in CASES:
        text = textgen.GENERATORS[gen](n, **kw)
        sa5 = oracle_lib.run_reference(text, mem=mem, threads=thr)
        assert len(sa5) == 5 * n, (name, len(sa5))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), sa5=sa5, gen=gen, n=n,
                            kwargs=repr(kw), mem=str(mem), threads=thr)
        print(name, "ok", len(sa5))


if __name__ == "__main__":
    main()
