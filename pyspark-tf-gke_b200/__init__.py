"""B200-native implementation of the raw-spark ETL hot path of
greg-ogs/PySpark-TF-GKE: hand-written sm_100a CUDA kernels behind
a C ABI (include/etl_b200.h, libetl_b200.so), called through ctypes; torch is
used only for device buffers, streams and torch.distributed.

Importing the package does not need a GPU; using it does."""
from . import _lib, synth  # noqa: F401
from ._lib import EtlError  # noqa: F401
from .engine import Column, Engine  # noqa: F401
from .k_means import EtlPipelineModel, KMeansWorkload  # noqa: F401
from .kmeans import ClusteringEvaluator, KMeans, KMeansModel  # noqa: F401
from .table import DeviceTable, read_csv  # noqa: F401
from .parquet import ParquetUnsupported, read_parquet  # noqa: F401
from .train_input import load_csv, load_csv_device, write_label_map  # noqa: F401
from .trainer import DeepModel, run_deep_training  # noqa: F401
from . import mlformat  # noqa: F401

__all__ = ["EtlError", "Engine", "Column"]
