"""The allocation-free projection entry (etl_project_filter_impute_noalloc).

It exists for callers that cannot let this library allocate on the device - a
process that handed the whole GPU to a pool allocator, or a loop that cannot
take the device-wide sync hidden in cudaMalloc - so its scratch is a fixed
__device__ array shared by every call.  The checks here go through the C ABI
and compare it, bit for bit (the tolerance every projection test uses: none),
with the handle-based entry on two calls in a row, because the second call sees
the first call's leftovers in that array.  The file name sorts last on purpose:
a run that stops at the first failure has already run everything else.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import pyspark_tf_gke_b200 as pkg
import pyspark_tf_gke_b200.synth as S
from pyspark_tf_gke_b200.engine import project_filter_impute_noalloc

pytestmark = pytest.mark.gpu


def _device_table(eng, n, seed):
    key = eng.synth_column(seed, 0, S.KIND_KEY32, n, modulus=30, null_rate=0.03)
    cols = [eng.synth_column(seed, 1 + c, S.KIND_REAL, n, null_rate=0.07, nan_rate=0.01) for c in range(3)]
    return key, cols


def _run_both(eng, n, seed, fill):
    key, cols = _device_table(eng, n, seed)
    ref_code, ref_vals, ref_rows = eng.project_filter_impute(key, cols, fill)
    # poison the outputs so stale data cannot pass for a result
    out_code = torch.full((n,), -7, dtype=torch.int32, device=eng.device)
    out_vals = [torch.full((n,), -7.0, dtype=torch.float64, device=eng.device) for _ in range(3)]
    out_rows = torch.full((1,), -7, dtype=torch.int64, device=eng.device)
    project_filter_impute_noalloc(key, cols, fill, out_code, out_vals, out_rows)
    torch.cuda.synchronize()
    r = int(ref_rows.item())
    assert int(out_rows.item()) == r
    assert torch.equal(out_code[:r], ref_code[:r])
    for a, b in zip(out_vals, ref_vals):
        assert a[:r].cpu().numpy().tobytes() == b[:r].cpu().numpy().tobytes()
    return r


def test_noalloc_matches_handle_entry_on_two_calls_in_a_row():
    eng = pkg.Engine(0)
    try:
        # different sizes and seeds back to back: the second call finds the first
        # call's chunk words and ticket in the shared array and must not care
        r1 = _run_both(eng, 3_000_003, seed=31, fill=[1.5, 2.5, 3.5])
        r2 = _run_both(eng, 700_001, seed=32, fill=[-1.0, 0.0, 1e10])
        r3 = _run_both(eng, 3_000_003, seed=31, fill=[1.5, 2.5, 3.5])
        assert r1 == r3 and r2 < r1
    finally:
        eng.close()


def test_noalloc_small_and_empty_inputs():
    eng = pkg.Engine(0)
    try:
        for n in (1, 33, 1025):
            _run_both(eng, n, seed=40 + n, fill=[0.0, 0.0, 0.0])
        rows = torch.full((1,), 5, dtype=torch.int64, device=eng.device)
        empty = pkg.Column(torch.empty(0, dtype=torch.int32, device=eng.device))
        evals = [pkg.Column(torch.empty(0, dtype=torch.float64, device=eng.device)) for _ in range(3)]
        project_filter_impute_noalloc(empty, evals, [0, 0, 0], empty.data, [c.data for c in evals], rows)
        torch.cuda.synchronize()
        assert int(rows.item()) == 0
    finally:
        eng.close()


def test_noalloc_rejects_tables_beyond_its_fixed_scratch():
    lib = pkg._lib.load()
    torch.cuda.init()
    buf = torch.zeros(64, dtype=torch.float64, device="cuda:0")
    VP3, I3, F3 = C.c_void_p * 3, C.c_int32 * 3, C.c_double * 3
    p = buf.data_ptr()
    rc = lib.etl_project_filter_impute_noalloc(p, None, C.byref(VP3(p, p, p)), C.byref(VP3(None, None, None)),
                                               C.byref(I3(0, 0, 0)), C.byref(F3(0, 0, 0)), (1 << 31) + 1, p,
                                               C.byref(VP3(p, p, p)), p, None)
    assert rc == -5  # ETL_ECAPACITY
    msg = lib.etl_last_error().decode()
    assert "exceeds the fixed scratch" in msg and str(1 << 31) in msg
