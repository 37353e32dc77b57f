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


def test_noalloc_leaves_device_memory_untouched():
    # the property the entry exists for: once the first call on the device has done its one-time set-up,
    # a call changes neither the driver's free-memory figure (no cudaMalloc / cudaFree anywhere below it)
    # nor torch's own allocator statistics (the binding allocates nothing either)
    eng = pkg.Engine(0)
    try:
        n, fill = 2_000_003, [1.0, 2.0, 3.0]
        key, cols = _device_table(eng, n, seed=51)
        out_code = torch.empty(n, dtype=torch.int32, device=eng.device)
        out_vals = [torch.empty(n, dtype=torch.float64, device=eng.device) for _ in range(3)]
        out_rows = torch.zeros(1, dtype=torch.int64, device=eng.device)
        project_filter_impute_noalloc(key, cols, fill, out_code, out_vals, out_rows)      # warm-up (set-up call)
        torch.cuda.synchronize()
        for _ in range(2):                                                               # the 2nd and 3rd calls
            free0, _ = torch.cuda.mem_get_info(eng.device)
            reserved0, allocated0 = torch.cuda.memory_reserved(eng.device), torch.cuda.memory_allocated(eng.device)
            project_filter_impute_noalloc(key, cols, fill, out_code, out_vals, out_rows)
            torch.cuda.synchronize()
            free1, _ = torch.cuda.mem_get_info(eng.device)
            assert free1 == free0
            assert torch.cuda.memory_reserved(eng.device) == reserved0
            assert torch.cuda.memory_allocated(eng.device) == allocated0
        ref_code, ref_vals, ref_rows = eng.project_filter_impute(key, cols, fill)
        r = int(ref_rows.item())
        assert int(out_rows.item()) == r and torch.equal(out_code[:r], ref_code[:r])
    finally:
        eng.close()


def test_noalloc_large_table_uses_more_tiles_per_chunk():
    # above ~67 M rows the fixed 16,384 chunk words force more than 4 tiles per chunk - the reason the
    # chunk size is not a constant; 80 M rows, compared with the handle entry bit for bit
    eng = pkg.Engine(0)
    try:
        r = _run_both(eng, 80_000_001, seed=61, fill=[0.5, -0.5, 7.0])
        assert 0 < r < 80_000_001
        _run_both(eng, 4_097, seed=62, fill=[0.5, -0.5, 7.0])       # and a small call right after the large one
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
