"""Spark ML's on-disk model layout - SURVEY.md §8f rank 3: what `model.save(path)` /
`pipeline_model.save(path)` write (spark_workload_to_cloud_k8s.py:149-154; the
commented block of k_means.py:94-104) and what `PipelineModel.load` /
`KMeansModel.load` read back (k_means.py:118-135).

Spark is a third-party dependency the reference does not pin (`FROM spark:latest`,
unpinned `pyspark`), so this follows the layout Spark 3.0 - 4.0 share
(org.apache.spark.ml.util.DefaultParamsWriter and the per-model writers):

    <path>/metadata/part-00000       one JSON line: class, timestamp, sparkVersion, uid,
                                     paramMap, defaultParamMap     (+ _SUCCESS marker)
    <path>/data/part-00000-<uuid>-c000.snappy.parquet              (models with data)
    <path>/stages/<i>_<uid>/...      PipelineModel: one directory per stage, same layout

      KMeansModel          rows ClusterData(clusterIdx int, clusterCenter vector)
      StringIndexerModel   one row Data(labelsArray array<array<string>>)
      OneHotEncoderModel   one row Data(categorySizes array<int>)
      VectorAssembler      metadata only

A `vector` is Spark's VectorUDT: struct<type tinyint, size int, indices array<int>,
values array<double>> (dense: type 1, size / indices null); the Parquet footer carries
Spark's schema JSON under `org.apache.spark.sql.parquet.row.metadata`, which is how its
reader recovers the UDT.  The files are a few KB of host-side metadata: they go through
pyarrow.parquet (the column DATA path is csrc/parquet.cu).  Parity: no JVM on this image,
so a real Spark has not loaded these files - parity unpinned; tests check the layout, the
JSON, the Parquet schemas and the round trip through load_*.

Not part of Spark's PipelineModel in the reference either: the three imputation means
(the reference imputes with withColumn outside the pipeline, k_means.py:45-51).  They
are kept in an extra `etl_b200.json` beside `metadata/`, which Spark ignores."""
from __future__ import annotations

import json
import os
import time
import uuid
from typing import List, Optional

import numpy as np

SPARK_VERSION = "3.5.1"
ROW_METADATA_KEY = b"org.apache.spark.sql.parquet.row.metadata"
KEY_COL, INDEX_COL, VEC_COL, FEATURES_COL = "measure_name", "measure_name_index", "measure_name_vec", "features"
NUMERIC_COLS = ["value", "lower_ci", "upper_ci"]

_VECTOR_SQL_TYPE = {"type": "struct", "fields": [
    {"name": "type", "type": "byte", "nullable": False, "metadata": {}},
    {"name": "size", "type": "integer", "nullable": True, "metadata": {}},
    {"name": "indices", "type": {"type": "array", "elementType": "integer", "containsNull": False}, "nullable": True,
     "metadata": {}},
    {"name": "values", "type": {"type": "array", "elementType": "double", "containsNull": False}, "nullable": True,
     "metadata": {}}]}
_VECTOR_UDT = {"type": "udt", "class": "org.apache.spark.ml.linalg.VectorUDT", "pyClass": "pyspark.ml.linalg.VectorUDT",
               "sqlType": _VECTOR_SQL_TYPE}


def _uid(prefix: str) -> str:
    return f"{prefix}_{uuid.uuid4().hex[:12]}"       # Identifiable.randomUID


def _write_metadata(path: str, cls: str, uid: str, param_map: dict, default_map: dict) -> None:
    d = os.path.join(path, "metadata")
    os.makedirs(d, exist_ok=True)
    meta = {"class": cls, "timestamp": int(time.time() * 1000), "sparkVersion": SPARK_VERSION, "uid": uid,
            "paramMap": param_map, "defaultParamMap": default_map}
    with open(os.path.join(d, "part-00000"), "w") as fh:
        fh.write(json.dumps(meta, separators=(",", ":")) + "\n")
    open(os.path.join(d, "_SUCCESS"), "w").close()


def _read_metadata(path: str, expect_class: Optional[str] = None) -> dict:
    d = os.path.join(path, "metadata")
    parts = sorted(f for f in os.listdir(d) if f.startswith("part-"))
    if not parts:
        raise FileNotFoundError(f"no metadata part file under {d}")
    with open(os.path.join(d, parts[0])) as fh:
        meta = json.loads(fh.readline())
    if expect_class is not None and meta.get("class") != expect_class:
        raise ValueError(f"{path}: metadata class {meta.get('class')!r}, expected {expect_class!r}")
    return meta


def _write_data(path: str, table, spark_schema: dict) -> str:
    import pyarrow.parquet as pq

    d = os.path.join(path, "data")
    os.makedirs(d, exist_ok=True)
    table = table.replace_schema_metadata({ROW_METADATA_KEY: json.dumps(spark_schema, separators=(",", ":")).encode()})
    name = f"part-00000-{uuid.uuid4()}-c000.snappy.parquet"
    pq.write_table(table, os.path.join(d, name), compression="snappy", use_compliant_nested_type=True)
    open(os.path.join(d, "_SUCCESS"), "w").close()
    return name


def _read_data(path: str) -> dict:
    """{column: [value per row]} of the part files under <path>/data, through this package's own Parquet
    reader (parquet.read_nested_small: Snappy, nested lists and structs) - what Spark does on the driver."""
    from .parquet import read_nested_small

    d = os.path.join(path, "data")
    files = sorted(f for f in os.listdir(d) if f.endswith(".parquet"))
    if not files:
        raise FileNotFoundError(f"no parquet part file under {d}")
    out: dict = {}
    for f in files:
        for name, values in read_nested_small(os.path.join(d, f)).items():
            out.setdefault(name, []).extend(values)
    return out


# ---------------------------------------------------------------- KMeansModel
KMEANS_CLASS = "org.apache.spark.ml.clustering.KMeansModel"
_KMEANS_DEFAULTS = {"distanceMeasure": "euclidean", "featuresCol": "features", "initMode": "k-means||", "initSteps": 2, "k": 2,
                    "maxBlockSizeInMB": 0.0, "maxIter": 20, "predictionCol": "prediction", "solver": "auto", "tol": 1.0e-4}


def save_kmeans_model(model, path: str, params: Optional[dict] = None) -> None:
    """KMeansModel.save(path) (spark_workload_to_cloud_k8s.py:152).  `params` = the
    estimator's explicitly set params ({"k": 25, "seed": 1, "maxIter": 1000} for
    k_means.py:82); k defaults to the number of centres."""
    import pyarrow as pa

    centers = np.asarray(model.centers, dtype=np.float64)
    k = centers.shape[0]
    vec = pa.struct([pa.field("type", pa.int8(), nullable=False), pa.field("size", pa.int32()),
                     pa.field("indices", pa.list_(pa.field("element", pa.int32(), nullable=False))),
                     pa.field("values", pa.list_(pa.field("element", pa.float64(), nullable=False)))])
    rows = [{"type": 1, "size": None, "indices": None, "values": centers[i].tolist()} for i in range(k)]
    table = pa.table({"clusterIdx": pa.array(list(range(k)), pa.int32()), "clusterCenter": pa.array(rows, vec)},
                     schema=pa.schema([pa.field("clusterIdx", pa.int32(), nullable=False), pa.field("clusterCenter", vec)]))
    spark_schema = {"type": "struct", "fields": [
        {"name": "clusterIdx", "type": "integer", "nullable": False, "metadata": {}},
        {"name": "clusterCenter", "type": _VECTOR_UDT, "nullable": True, "metadata": {}}]}
    pm = {"k": k}
    pm.update(params or {})
    _write_metadata(path, KMEANS_CLASS, _uid("KMeans"), pm, dict(_KMEANS_DEFAULTS))
    _write_data(path, table, spark_schema)
    with open(os.path.join(path, "etl_b200.json"), "w") as fh:    # the structured-feature layout (not Spark's)
        json.dump({"num_labels": int(model.num_labels), "repeats": int(model.repeats)}, fh)


def load_kmeans_model(engine, path: str):
    """KMeansModel.load(path) (k_means.py:131): centres in clusterIdx order.  Reads
    what save_kmeans_model or Spark itself wrote (dense or sparse centre vectors)."""
    from .kmeans import KMeansModel

    meta = _read_metadata(path, KMEANS_CLASS)
    t = _read_data(path)
    order = np.argsort(np.asarray(t["clusterIdx"]))
    rows = []
    for i in order:
        v = t["clusterCenter"][i]
        if v["type"] == 1:
            rows.append(np.asarray(v["values"], dtype=np.float64))
        else:
            dense = np.zeros(int(v["size"]), dtype=np.float64)
            dense[np.asarray(v["indices"], dtype=np.int64)] = np.asarray(v["values"], dtype=np.float64)
            rows.append(dense)
    centers = np.stack(rows)
    k = meta["paramMap"].get("k", meta["defaultParamMap"].get("k"))
    if k is not None and int(k) < centers.shape[0]:
        raise ValueError(f"{path}: {centers.shape[0]} centres for k={k}")
    side = os.path.join(path, "etl_b200.json")
    if os.path.exists(side):
        with open(side) as fh:
            s = json.load(fh)
        num_labels, repeats = int(s["num_labels"]), int(s["repeats"])
    else:   # written by Spark: the caller's pipeline model knows the layout; assume the reference's default weight
        repeats = 5
        num_labels = (centers.shape[1] - 3) // repeats
    return KMeansModel(engine, centers, num_labels, repeats)


# -------------------------------------------------------------- PipelineModel
PIPELINE_CLASS = "org.apache.spark.ml.PipelineModel"
INDEXER_CLASS = "org.apache.spark.ml.feature.StringIndexerModel"
ENCODER_CLASS = "org.apache.spark.ml.feature.OneHotEncoderModel"
ASSEMBLER_CLASS = "org.apache.spark.ml.feature.VectorAssembler"


def _stage_dir(path: str, idx: int, n_stages: int, uid: str) -> str:
    return os.path.join(path, "stages", f"{idx:0{len(str(n_stages))}d}_{uid}")       # Pipeline.SharedReadWrite.getStagePath


def save_pipeline_model(pm, path: str) -> None:
    """PipelineModel.save(path) (spark_workload_to_cloud_k8s.py:154) for the stages of
    k_means.py:34-66: StringIndexer(handleInvalid=keep) -> OneHotEncoder ->
    VectorAssembler([measure_name_vec] * repeats + numerics, handleInvalid=keep)."""
    import pyarrow as pa

    uids = [_uid("StringIndexer"), _uid("OneHotEncoder"), _uid("VectorAssembler")]
    _write_metadata(path, PIPELINE_CLASS, _uid("PipelineModel"), {"stageUids": uids}, {})
    # stage 0: the fitted label order
    d0 = _stage_dir(path, 0, 3, uids[0])
    _write_metadata(d0, INDEXER_CLASS, uids[0], {"handleInvalid": "keep", "inputCol": KEY_COL, "outputCol": INDEX_COL},
                    {"handleInvalid": "error", "outputCol": uids[0] + "__output", "stringOrderType": "frequencyDesc"})
    inner = pa.list_(pa.field("element", pa.string()))
    t0 = pa.table({"labelsArray": pa.array([[list(pm.labels)]], pa.list_(pa.field("element", inner)))})
    _write_data(d0, t0, {"type": "struct", "fields": [{"name": "labelsArray", "type": {
        "type": "array", "elementType": {"type": "array", "elementType": "string", "containsNull": True},
        "containsNull": True}, "nullable": True, "metadata": {}}]})
    # stage 1: the index column has numLabels + 1 categories (handleInvalid=keep adds one); dropLast keeps numLabels
    d1 = _stage_dir(path, 1, 3, uids[1])
    _write_metadata(d1, ENCODER_CLASS, uids[1], {"inputCol": INDEX_COL, "outputCol": VEC_COL},
                    {"dropLast": True, "handleInvalid": "error", "outputCol": uids[1] + "__output"})
    t1 = pa.table({"categorySizes": pa.array([[len(pm.labels) + 1]], pa.list_(pa.field("element", pa.int32(), nullable=False)))})
    _write_data(d1, t1, {"type": "struct", "fields": [{"name": "categorySizes", "type": {
        "type": "array", "elementType": "integer", "containsNull": False}, "nullable": True, "metadata": {}}]})
    # stage 2: a transformer, no data
    d2 = _stage_dir(path, 2, 3, uids[2])
    _write_metadata(d2, ASSEMBLER_CLASS, uids[2], {"handleInvalid": "keep", "inputCols": [VEC_COL] * int(pm.repeats) + NUMERIC_COLS,
                                                  "outputCol": FEATURES_COL},
                    {"handleInvalid": "error", "outputCol": uids[2] + "__output"})
    with open(os.path.join(path, "etl_b200.json"), "w") as fh:
        json.dump({"means": list(pm.means), "label_counts": [int(c) for c in pm.label_counts],
                   "null_key_count": int(pm.null_key_count), "rows_after_filter": int(pm.rows_after_filter)}, fh)


def load_pipeline_model(engine, path: str):
    """PipelineModel.load(path) (k_means.py:124) -> EtlPipelineModel.  Accepts exactly
    the reference's stage sequence; anything else raises."""
    from .k_means import EtlPipelineModel

    meta = _read_metadata(path, PIPELINE_CLASS)
    uids: List[str] = meta["paramMap"]["stageUids"]
    dirs = [_stage_dir(path, i, len(uids), u) for i, u in enumerate(uids)]
    metas = [_read_metadata(d) for d in dirs]
    classes = [m["class"] for m in metas]
    if classes != [INDEXER_CLASS, ENCODER_CLASS, ASSEMBLER_CLASS]:
        raise ValueError(f"{path}: stages {classes} are not StringIndexerModel -> OneHotEncoderModel -> VectorAssembler")
    labels_array = _read_data(dirs[0])["labelsArray"][0]
    if len(labels_array) != 1:
        raise ValueError(f"{path}: StringIndexerModel with {len(labels_array)} input columns")
    labels = list(labels_array[0])
    sizes = _read_data(dirs[1])["categorySizes"][0]
    if list(sizes) != [len(labels) + 1]:
        raise ValueError(f"{path}: OneHotEncoder categorySizes {sizes} do not match {len(labels)} labels + 1")
    cols = metas[2]["paramMap"]["inputCols"]
    repeats = sum(1 for c in cols if c == metas[1]["paramMap"]["outputCol"])
    if repeats < 1 or cols[repeats:] != NUMERIC_COLS:
        raise ValueError(f"{path}: VectorAssembler inputCols {cols}")
    side, extra = os.path.join(path, "etl_b200.json"), {}
    if os.path.exists(side):
        with open(side) as fh:
            extra = json.load(fh)
    return EtlPipelineModel(engine, labels, extra.get("label_counts", [0] * len(labels)), extra.get("means", [None] * 3), repeats,
                            extra.get("null_key_count", 0), extra.get("rows_after_filter", 0),
                            features_dim=repeats * len(labels) + 3)
