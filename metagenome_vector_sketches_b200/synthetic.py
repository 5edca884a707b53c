"""Closed-form synthetic hash batches, numpy side.

`mvs-synth-v1`: hash(s, k) is a BIJECTION of the 54-bit key (s << 24 | k), so every
hash of a batch is distinct by construction (set semantics hold without a dedupe
pass) and lies in [0, 2^54) ~ the FracMinHash scaled=1000 range of the toy
signatures (max_hash 18446744073709552 ~ 2^54.03).  The same arithmetic is
implemented on the device by bench.py (torch int64 ops wrap like uint64).
"""
import numpy as np

MASK54 = (1 << 54) - 1
_M1 = 0x9E3779B97F4A7C15
_M2 = 0xBF58476D1CE4E5B9
_M3 = 0x94D049BB133111EB


def _u64(x):
    return np.asarray(x, dtype=np.uint64)


def synth_hashes(sample_ids, ks, seed=0):
    """Vectorised hash(s, k).  sample_ids < 2^30, ks < 2^24."""
    with np.errstate(over="ignore"):
        m = np.uint64(MASK54)
        x = (_u64(sample_ids) << np.uint64(24)) | _u64(ks)
        x = x ^ (np.uint64((seed * _M1) & MASK54))
        x = (x * np.uint64(_M1)) & m
        x ^= x >> np.uint64(29)
        x = (x * np.uint64(_M2)) & m
        x ^= x >> np.uint64(27)
        x = (x * np.uint64(_M3)) & m
        x ^= x >> np.uint64(31)
        return x


def uniform_batch(n_samples, hashes_per_sample, seed=0, first_sample=0):
    """CSR batch with n_samples x hashes_per_sample hashes (BASELINE config 2/3 shape)."""
    s = np.repeat(np.arange(first_sample, first_sample + n_samples, dtype=np.uint64), hashes_per_sample)
    k = np.tile(np.arange(hashes_per_sample, dtype=np.uint64), n_samples)
    offsets = np.arange(n_samples + 1, dtype=np.uint64) * np.uint64(hashes_per_sample)
    return synth_hashes(s, k, seed), offsets


def _splitmix64(x):
    with np.errstate(over="ignore"):
        x = _u64(x) + np.uint64(_M1)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(_M2)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(_M3)
        return x ^ (x >> np.uint64(31))


def lognormal_sizes(n_samples, seed=0, median=2e4, sigma=1.5, lo=1e3, hi=1e7, first_sample=0):
    """BASELINE config 4 sizes: clip(round(exp(N(ln median, sigma^2))), lo, hi), counter-based."""
    s = np.arange(first_sample, first_sample + n_samples, dtype=np.uint64)
    with np.errstate(over="ignore"):
        base = np.uint64((seed << 40) & 0xFFFFFFFFFFFFFFFF)
        a = _splitmix64(base ^ (s * np.uint64(2)))
        b = _splitmix64(base ^ (s * np.uint64(2) + np.uint64(1)))
    u1 = ((a >> np.uint64(11)).astype(np.float64) + 0.5) / 2.0 ** 53
    u2 = ((b >> np.uint64(11)).astype(np.float64) + 0.5) / 2.0 ** 53
    z = np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    n = np.rint(np.exp(np.log(median) + sigma * z))
    return np.clip(n, lo, hi).astype(np.int64)


def ragged_batch(sizes, seed=0, first_sample=0):
    sizes = np.asarray(sizes, dtype=np.int64)
    offsets = np.zeros(len(sizes) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum(sizes).astype(np.uint64)
    total = int(offsets[-1])
    s = np.repeat(np.arange(first_sample, first_sample + len(sizes), dtype=np.uint64), sizes)
    k = np.arange(total, dtype=np.uint64) - np.repeat(offsets[:-1], sizes)
    return synth_hashes(s, k, seed), offsets


def _s64(v):
    """uint64 constant as the int64 with the same bits (torch has no uint64 arithmetic)."""
    v &= 0xFFFFFFFFFFFFFFFF
    return v - (1 << 64) if v >= (1 << 63) else v


def uniform_batch_torch(n_samples, hashes_per_sample, device, seed=0, first_sample=0, chunk_samples=512):
    """Same batch as uniform_batch(), generated on `device` with torch int64 ops (bit-identical:
    products wrap modulo 2^64 and every intermediate is masked to 54 bits, so it stays >= 0).
    Returns (int64 tensor holding the uint64 bit patterns, numpy uint64 offsets)."""
    import torch
    out = torch.empty(n_samples * hashes_per_sample, dtype=torch.int64, device=device)
    k = torch.arange(hashes_per_sample, dtype=torch.int64, device=device)
    sx = (seed * _M1) & MASK54
    for s0 in range(0, n_samples, chunk_samples):
        s1 = min(n_samples, s0 + chunk_samples)
        s = torch.arange(first_sample + s0, first_sample + s1, dtype=torch.int64, device=device)
        x = ((s << 24)[:, None] | k[None, :]).reshape(-1)
        x = x ^ sx
        x = (x * _s64(_M1)) & MASK54
        x = x ^ (x >> 29)
        x = (x * _s64(_M2)) & MASK54
        x = x ^ (x >> 27)
        x = (x * _s64(_M3)) & MASK54
        x = x ^ (x >> 31)
        out[s0 * hashes_per_sample:s1 * hashes_per_sample] = x
    offsets = np.arange(n_samples + 1, dtype=np.uint64) * np.uint64(hashes_per_sample)
    return out, offsets


def ragged_batch_torch(sizes, device, seed=0, first_sample=0, chunk_hashes=1 << 27):
    """Same batch as ragged_batch(), generated on `device` in chunks of whole samples (a sample
    larger than chunk_hashes forms its own chunk).  Returns (int64 tensor, numpy uint64 offsets)."""
    import torch
    sizes = np.asarray(sizes, dtype=np.int64)
    offsets = np.zeros(len(sizes) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum(sizes).astype(np.uint64)
    total = int(offsets[-1])
    out = torch.empty(total, dtype=torch.int64, device=device)
    sx = (seed * _M1) & MASK54
    s0 = 0
    n = len(sizes)
    while s0 < n:
        s1 = s0 + 1
        while s1 < n and int(offsets[s1 + 1] - offsets[s0]) <= chunk_hashes:
            s1 += 1
        h0, h1 = int(offsets[s0]), int(offsets[s1])
        if h1 > h0:
            cnt = torch.from_numpy(sizes[s0:s1]).to(device)
            sid = torch.arange(first_sample + s0, first_sample + s1, dtype=torch.int64, device=device)
            start = torch.from_numpy(offsets[s0:s1].astype(np.int64) - h0).to(device)
            s = torch.repeat_interleave(sid, cnt)
            k = torch.arange(h1 - h0, dtype=torch.int64, device=device) - torch.repeat_interleave(start, cnt)
            x = (s << 24) | k
            del s, k
            x = x ^ sx
            x = (x * _s64(_M1)) & MASK54
            x = x ^ (x >> 29)
            x = (x * _s64(_M2)) & MASK54
            x = x ^ (x >> 27)
            x = (x * _s64(_M3)) & MASK54
            x = x ^ (x >> 31)
            out[h0:h1] = x
            del x
        s0 = s1
    return out, offsets
