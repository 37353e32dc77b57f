"""ctypes loader of oracle/liboracle.so (the compiled CPU oracle).  TEST
INFRASTRUCTURE ONLY - see the header of etl_oracle.c."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
VP3 = C.c_void_p * 3
I32x3 = C.c_int32 * 3
F64x3 = C.c_double * 3
_NP_CODE = {np.dtype(np.float64): 0, np.dtype(np.float32): 1, np.dtype(np.int32): 2}
_lib = None


def build():
    subprocess.run(["make", "-C", HERE, "liboracle.so"], check=True, stdout=subprocess.DEVNULL)
    return LIB


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        lib = C.CDLL(LIB)
        lib.oracle_threads.restype = C.c_int
        lib.oracle_set_threads.argtypes = [C.c_int]
        lib.oracle_null_count.restype = C.c_int64
        lib.oracle_null_count.argtypes = [C.c_void_p, C.c_int64]
        lib.oracle_sum_count.restype = None
        lib.oracle_sum_count.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64,
                                         C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        lib.oracle_project_filter_impute.restype = C.c_int64
        lib.oracle_project_filter_impute.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.POINTER(VP3),
                                                     C.POINTER(I32x3), C.POINTER(F64x3), C.c_int64, C.c_void_p,
                                                     C.POINTER(VP3)]
        lib.oracle_synth_column.restype = None
        lib.oracle_synth_column.argtypes = [C.c_uint64, C.c_uint32, C.c_int, C.c_uint64, C.c_int64, C.c_int64, C.c_void_p]
        lib.oracle_groupby_sum.restype = C.c_int64
        lib.oracle_groupby_sum.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = lib
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data


def threads() -> int:
    return load().oracle_threads()


def null_count(valid_words, n: int) -> int:
    return load().oracle_null_count(_p(valid_words), n)


def sum_count(values: np.ndarray, valid_words, row_valid_words=None):
    s, c = C.c_double(), C.c_int64()
    load().oracle_sum_count(_p(values), _p(valid_words), _NP_CODE[values.dtype], _p(row_valid_words),
                            values.shape[0], C.byref(s), C.byref(c))
    return s.value, c.value


def project_filter_impute(key_code, key_valid_words, vals, vals_valid_words, fill, out_code=None, out_vals=None):
    n = key_code.shape[0]
    out_code = np.empty(n, dtype=np.int32) if out_code is None else out_code
    out_vals = [np.empty(n, dtype=np.float64) for _ in range(3)] if out_vals is None else out_vals
    val = VP3(*[_p(v) for v in vals])
    vv = VP3(*[_p(v) for v in vals_valid_words])
    dt = I32x3(*[_NP_CODE[v.dtype] for v in vals])
    fl = F64x3(*[float(f) for f in fill])
    ov = VP3(*[_p(v) for v in out_vals])
    rows = load().oracle_project_filter_impute(_p(key_code), _p(key_valid_words), C.byref(val), C.byref(vv),
                                               C.byref(dt), C.byref(fl), n, _p(out_code), C.byref(ov))
    return rows, out_code, out_vals


def synth_column(seed: int, stream_id: int, kind: int, n: int, row0: int = 0, modulus: int = 1) -> np.ndarray:
    """pyspark_tf_gke_b200.synth.host_column without nulls (kinds EXACT = 0 and KEY64 = 3), on every host core."""
    assert kind in (0, 3)
    out = np.empty(n, dtype=np.float64 if kind == 0 else np.int64)
    load().oracle_synth_column(seed, stream_id, kind, modulus, row0, n, _p(out))
    return out


def groupby_sum(keys: np.ndarray, vals: np.ndarray, scratch=None):
    """groupBy(key).agg(count, sum(value)) -> (keys, count, sum), unordered.  `scratch` = three
    arrays of len(keys) entries to reuse between calls (timing loops)."""
    n = keys.shape[0]
    assert keys.dtype == np.int64 and vals.dtype == np.float64 and vals.shape[0] == n
    ok, oc, os_ = scratch or (np.empty(n, np.int64), np.empty(n, np.int64), np.empty(n, np.float64))
    g = load().oracle_groupby_sum(_p(keys), _p(vals), n, _p(ok), _p(oc), _p(os_))
    if g < 0:
        raise MemoryError("oracle_groupby_sum: out of host memory")
    return ok[:g], oc[:g], os_[:g]
