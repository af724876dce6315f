"""Deterministic synthetic inputs shared by make_golden.py and the tests.

Values come from an integer hash evaluated in numpy (identical on every machine and torch
version), so large inputs (KV caches) need not be stored in the fixtures.
"""
import numpy as np
import torch


def hash_uniform(n: int, seed: int, lo: float = -1.0, hi: float = 1.0) -> np.ndarray:
    """n float64 values in [lo, hi], a pure function of (index, seed)."""
    off = np.uint64((int(seed) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        return _mix(np.arange(n, dtype=np.uint64) + off, lo, hi)


def _mix(x, lo, hi):
    x ^= x >> np.uint64(30)
    x = x * np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27)
    x = x * np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    u = (x >> np.uint64(11)).astype(np.float64) / float(1 << 53)
    return lo + (hi - lo) * u


def tensor(shape, seed: int, dtype: torch.dtype, scale: float = 1.0) -> torch.Tensor:
    n = int(np.prod(shape))
    return torch.from_numpy(hash_uniform(n, seed, -scale, scale)).reshape(shape).to(dtype)


def permutation(n: int, seed: int) -> np.ndarray:
    return np.argsort(hash_uniform(n, seed), kind="stable")
