"""Deterministic synthetic tables: a counter-based generator so
any slice of any column is reproducible on the host (NumPy, here) and on the
device (csrc/synth.cu) bit for bit.

bits(seed, stream, row) = mix64(key + (row + 1) * GOLDEN), key derived from
(seed, stream): the splitmix64 sequence seeded with `key`.  Each column uses two
streams: 2*s for values, 2*s + 1 for null / NaN placement."""
from __future__ import annotations

import numpy as np

GOLDEN = np.uint64(0x9E3779B97F4A7C15)
C2 = np.uint64(0xBF58476D1CE4E5B9)
C3 = np.uint64(0x94D049BB133111EB)

KIND_EXACT, KIND_REAL, KIND_KEY32, KIND_KEY64 = 0, 1, 2, 3


def _mix64(z: np.ndarray) -> np.ndarray:
    z = (z ^ (z >> np.uint64(30))) * C2
    z = (z ^ (z >> np.uint64(27))) * C3
    return z ^ (z >> np.uint64(31))


def _key(seed: int, stream: int) -> np.uint64:
    base = (seed * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    return np.uint64(base ^ ((stream * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF))


def bits(seed: int, stream: int, row0: int, n: int) -> np.ndarray:
    rows = np.arange(row0 + 1, row0 + n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return _mix64(_key(seed, stream) + rows * GOLDEN)


def unit(b: np.ndarray) -> np.ndarray:
    return (b >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def host_column(seed: int, stream_id: int, kind: int, n: int, row0: int = 0, modulus: int = 0,
                null_rate: float = 0.0, nan_rate: float = 0.0):
    """-> (values, valid bool array).  Same definition as etl_synth_column."""
    b = bits(seed, 2 * stream_id, row0, n)
    un = unit(bits(seed, 2 * stream_id + 1, row0, n))
    valid = un >= null_rate
    is_nan = valid & (un < null_rate + nan_rate)
    if kind == KIND_EXACT:
        v = np.floor(unit(b) * 100001.0)
        v[is_nan] = np.nan
    elif kind == KIND_REAL:
        v = unit(b) * 100000.0
        v[is_nan] = np.nan
    elif kind == KIND_KEY32:
        v = (b % np.uint64(modulus)).astype(np.int32)
    elif kind == KIND_KEY64:
        v = (b % np.uint64(modulus)).astype(np.int64)
    else:
        raise ValueError(f"kind={kind}")
    return v, valid


# The named shapes of BASELINE.json `configs`
CONFIG2 = dict(rows=10_000_000, seed=1, categories=30, key_null=0.01, val_null=(0.071, 0.074, 0.074))
CONFIG3 = dict(rows=100_000_000, seed=2, groups=1000, val_null=(0.071, 0.074, 0.074))
CONFIG4 = dict(rows=1_000_000_000, seed=3, groups=100_000_000)
