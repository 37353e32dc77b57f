"""ctypes binding of libetl_b200.so (include/etl_b200.h).

There is no fallback: if the library is missing or a call fails, an exception is
raised.  Nothing here, or anywhere in this package, imports oracle/."""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libetl_b200.so")

ETL_F64, ETL_F32, ETL_I32, ETL_I64 = 0, 1, 2, 3


class EtlError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libetl_b200 error {code}: {message}")
        self.code = code


_lib = None
_lock = threading.Lock()

VP3 = C.c_void_p * 3
I32x3 = C.c_int32 * 3
F64x3 = C.c_double * 3

_PROTOS = {
    "etl_abi_version": (C.c_int, []),
    "etl_last_error": (C.c_char_p, []),
    "etl_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "etl_destroy": (C.c_int, [C.c_void_p]),
    "etl_reserve": (C.c_int, [C.c_void_p, C.c_size_t]),
    "etl_project_filter_impute": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(VP3), C.POINTER(VP3),
                                            C.POINTER(I32x3), C.POINTER(F64x3), C.c_int64, C.c_void_p,
                                            C.POINTER(VP3), C.c_void_p, C.c_void_p]),
    "etl_project_filter_impute_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(VP3), C.POINTER(VP3),
                                                 C.POINTER(I32x3), C.POINTER(F64x3), C.c_int64, C.c_void_p,
                                                 C.POINTER(VP3), C.POINTER(C.c_int64)]),
    "etl_project_filter_impute_host_launches": (C.c_int64, [C.c_int64]),
    "etl_project_filter_impute_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.POINTER(VP3),
                                                    C.POINTER(I32x3), C.POINTER(F64x3), C.c_int64, C.c_void_p,
                                                    C.POINTER(VP3), C.c_void_p, C.c_void_p]),
    "etl_null_count": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "etl_column_sum_count": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(VP3), C.POINTER(VP3), C.POINTER(I32x3),
                                       C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_column_moments": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(VP3), C.POINTER(VP3), C.POINTER(I32x3),
                                     C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_csv_index_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "etl_csv_row_offsets": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "etl_csv_parse_columns": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_remap_codes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "etl_code_histogram": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_void_p]),
    "etl_assemble_features_dense": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.c_void_p, C.c_void_p, C.c_int64,
                                              C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "etl_assemble_features_sparse": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.c_void_p, C.c_void_p,
                                               C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_pack_tf_features": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(VP3), C.POINTER(VP3), C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p]),
    "etl_nan_to_null": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "etl_bucketize": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p,
                                C.c_void_p, C.c_void_p]),
    "etl_groupby_table_words": (C.c_int, [C.c_int32, C.c_int32, C.c_int64, C.POINTER(C.c_int64)]),
    "etl_groupby_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "etl_groupby_accumulate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                         C.c_int32, C.c_void_p, C.c_void_p]),
    "etl_groupby_compact": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_hash_partition": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                     C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_hash_partition_count": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                           C.c_void_p]),
    "etl_hash_partition_scatter_peers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                                                   C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_hash_subpartition": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_groupby_merge_partitioned": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32,
                                                C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "etl_slab_partition_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                          C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                          C.c_int32, C.c_void_p]),
    "etl_slab_partition_records": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                             C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p,
                                             C.c_void_p, C.c_int32, C.c_void_p]),
    "etl_peer_copy_slabs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                      C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]),
    "etl_peer_copy_slabs_bulk": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                           C.c_int64, C.c_void_p, C.c_int32, C.c_void_p]),
    "etl_copy_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "etl_groupby_merge_slabs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64,
                                          C.c_int32, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_int64, C.c_void_p, C.c_void_p]),
    "etl_peer_publish_counts": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "etl_peer_write_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "etl_parquet_expand_streams": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]),
    "etl_parquet_inflate_pages": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "etl_parquet_plain_strings": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32,
                                            C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_parquet_gather_pages": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p]),
    "etl_kmeans_prepare": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "etl_kmeans_assign": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "etl_kmeans_row_norms": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                       C.c_void_p]),
    "etl_kmeans_silhouette": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                        C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_kmeans_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                    C.c_void_p, C.c_void_p]),
    "etl_kmeans_accumulate": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VP3), C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_mlp_param_count": (C.c_int, [C.c_int32]),
    "etl_mlp_train_steps": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                                      C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_mlp_train_steps_ragged": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32,
                                             C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float,
                                             C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "etl_mlp_evaluate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p,
                                   C.c_void_p, C.c_void_p]),
    "etl_synth_column": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32, C.c_uint64, C.c_double,
                                   C.c_double, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
}


def load():
    """dlopen the in-tree library and attach prototypes.  Raises if absent."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  This package has no CPU fallback."
            )
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)  # AttributeError if the .so is stale
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib


def check(rc: int):
    if rc != 0:
        raise EtlError(rc, load().etl_last_error().decode("utf-8", "replace"))


def exported_names():
    return sorted(_PROTOS)
