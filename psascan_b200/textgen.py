"""Deterministic synthetic texts for the five BASELINE.json configs.

numpy only; every generator is seeded and never emits byte 255 (the reference rejects it,
/root/reference/include/inmem_psascan_src/initial_partial_sufsort.hpp:70-75).
"""
import numpy as np


def dna(n: int, seed: int = 1) -> np.ndarray:
    """config 1: i.i.d. uniform over ACGT."""
    rng = np.random.default_rng(seed)
    return np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=n, dtype=np.uint8)]


def random_bytes(n: int, seed: int = 2, sigma: int = 255) -> np.ndarray:
    """config 2 / 5: i.i.d. uniform over 0..sigma-1 (sigma <= 255)."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, sigma, size=n, dtype=np.uint8)


_MARKOV_ALPHA = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz .,\n", dtype=np.uint8)


def markov(n: int, seed: int = 3, order: int = 5, rows_log: int = 20) -> np.ndarray:
    """config 3: order-5 Markov over 30 symbols.  Context (last `order` symbols) hashed to
    2^rows_log rows; each row has a Zipf(1.2) distribution over 8 symbols picked by a hash
    of the row.  Generated as 4096 independent lanes advanced in lock-step so numpy can
    vectorise it; lane l owns text[l*L:(l+1)*L]."""
    lanes = (1 << 16) if n >= (1 << 24) else (4096 if n >= (1 << 16) else 1)
    per = -(-n // lanes)
    rng = np.random.default_rng(seed)
    w = 1.0 / np.arange(1, 9) ** 1.2
    cdf = np.cumsum(w / w.sum())
    cdf[-1] = 1.0
    out = np.empty((per, lanes), dtype=np.uint8)   # row t = symbol t of every lane
    ctx = rng.integers(0, 1 << 62, size=lanes, dtype=np.uint64)
    mask = np.uint64((1 << (5 * order)) - 1)
    rmask = np.uint64((1 << rows_log) - 1)
    mul = np.uint64(0x9E3779B97F4A7C15)
    sd = np.uint64(seed)
    for t in range(per):
        row = ((ctx & mask) * mul) >> np.uint64(40)
        row &= rmask
        u = rng.random(lanes)
        k = np.searchsorted(cdf, u).astype(np.uint64)          # 0..7, Zipf rank
        h = (row ^ sd) * mul + k * np.uint64(0xBF58476D1CE4E5B9)
        h ^= h >> np.uint64(31)
        sym = (h % np.uint64(30)).astype(np.uint8)
        out[t] = _MARKOV_ALPHA[sym]
        ctx = (ctx << np.uint64(5)) | sym.astype(np.uint64)
    return np.ascontiguousarray(out.T).reshape(-1)[:n].copy()


def repetitive(n: int, seed: int = 4, seed_len: int = 1 << 20, rate: float = 1e-3) -> np.ndarray:
    """config 4: copies of a `seed_len` seed string over a-z, each copy with independent
    point substitutions at `rate`."""
    rng = np.random.default_rng(seed)
    base = rng.integers(97, 123, size=min(seed_len, n), dtype=np.uint8)
    reps = -(-n // len(base))
    out = np.tile(base, reps)[:n].copy()
    nmut = int(n * rate)
    if nmut:
        pos = rng.integers(0, n, size=nmut)
        out[pos] = rng.integers(97, 123, size=nmut, dtype=np.uint8)
    return out


def periodic(n: int, period: int = 997, sigma: int = 3, seed: int = 7) -> np.ndarray:
    """long-LCP stress text over bytes {0..sigma-1} (byte 0 included on purpose)."""
    rng = np.random.default_rng(seed)
    base = rng.integers(0, sigma, size=period, dtype=np.uint8)
    return np.tile(base, -(-n // period))[:n].copy()


GENERATORS = {"dna": dna, "random": random_bytes, "markov": markov,
              "repetitive": repetitive, "periodic": periodic}


def markov_torch(n: int, seed: int = 3, order: int = 5, rows_log: int = 20, device="cuda"):
    """same model as markov(), generated with torch on `device` (bench inputs of GiB size; the
    stream differs from the numpy generator's, the statistics do not)."""
    import torch
    lanes = 1 << 20 if n >= (1 << 26) else 1 << 14
    per = -(-n // lanes)
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    w = 1.0 / torch.arange(1, 9, dtype=torch.float64) ** 1.2
    cdf = torch.cumsum(w / w.sum(), 0).to(device=device, dtype=torch.float32)
    cdf[-1] = 2.0
    alpha = torch.from_numpy(_MARKOV_ALPHA.copy()).to(device)
    out = torch.empty((per, lanes), dtype=torch.uint8, device=device)
    ctx = torch.randint(0, 1 << 62, (lanes,), generator=g, device=device, dtype=torch.int64)
    mask = (1 << (5 * order)) - 1
    rmask = (1 << rows_log) - 1
    mul = -7046029254386353131          # 0x9E3779B97F4A7C15 as int64
    mul2 = -4658895280553007687         # 0xBF58476D1CE4E5B9 as int64
    for t in range(per):
        row = (((ctx & mask) * mul) >> 40) & rmask
        u = torch.rand(lanes, generator=g, device=device)
        k = torch.searchsorted(cdf, u).to(torch.int64)
        h = (row ^ seed) * mul + k * mul2
        h = h ^ ((h >> 31) & 0x1FFFFFFFF)
        sym = torch.remainder(h, 30)
        out[t] = alpha[sym]
        ctx = (ctx << 5) | sym
    return out.t().contiguous().reshape(-1)[:n]
