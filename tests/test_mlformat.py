"""Spark ML's on-disk model layout (mlformat.py): directory
structure, metadata JSON, Parquet schemas and the round trip.  No GPU involved.
Parity unpinned: there is no JVM / pyspark on this image to load the files with."""
import json
import os
import re

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

import pyspark_tf_gke_b200 as pkg
from pyspark_tf_gke_b200 import mlformat as M


def _tree(root):
    out = []
    for d, _, files in os.walk(root):
        out += [os.path.relpath(os.path.join(d, f), root) for f in files]
    return sorted(out)


def _meta(path):
    with open(os.path.join(path, "metadata", "part-00000")) as fh:
        lines = fh.read().splitlines()
    assert len(lines) == 1                                  # one JSON object on one line
    return json.loads(lines[0])


def test_kmeans_model_layout_and_round_trip(tmp_path):
    rng = np.random.default_rng(1)
    model = pkg.KMeansModel(None, rng.normal(size=(25, 153)), 30, 5)
    path = str(tmp_path / "health_kmeans_model")
    model.save(path, params={"k": 25, "seed": 1, "maxIter": 1000})          # k_means.py:82
    files = _tree(path)
    assert "metadata/part-00000" in files and "metadata/_SUCCESS" in files and "data/_SUCCESS" in files
    data = [f for f in files if f.startswith("data/part-")]
    assert len(data) == 1 and re.fullmatch(r"data/part-00000-[0-9a-f-]{36}-c000\.snappy\.parquet", data[0])
    meta = _meta(path)
    assert meta["class"] == "org.apache.spark.ml.clustering.KMeansModel"
    assert set(meta) == {"class", "timestamp", "sparkVersion", "uid", "paramMap", "defaultParamMap"}
    assert meta["paramMap"] == {"k": 25, "seed": 1, "maxIter": 1000} and meta["uid"].startswith("KMeans_")
    assert meta["defaultParamMap"]["initMode"] == "k-means||" and meta["defaultParamMap"]["tol"] == 1e-4
    f = os.path.join(path, data[0])
    schema = pq.read_schema(f)
    assert schema.field("clusterIdx").type == pa.int32()
    vec = schema.field("clusterCenter").type
    assert [vec.field(i).name for i in range(vec.num_fields)] == ["type", "size", "indices", "values"]
    assert vec.field("type").type == pa.int8() and vec.field("values").type.value_type == pa.float64()
    spark_schema = json.loads(schema.metadata[M.ROW_METADATA_KEY])
    assert spark_schema["fields"][1]["type"]["class"] == "org.apache.spark.ml.linalg.VectorUDT"
    assert pq.ParquetFile(f).metadata.row_group(0).column(0).compression == "SNAPPY"
    back = pkg.KMeansModel.load(path)
    assert back.centers.tobytes() == model.centers.tobytes() and (back.num_labels, back.repeats) == (30, 5)


def test_kmeans_model_written_by_spark_with_sparse_centres_and_shuffled_rows(tmp_path):
    # what a Spark job could have left: no side file, rows in any order, a sparse centre vector
    path = str(tmp_path / "m")
    os.makedirs(os.path.join(path, "metadata"))
    os.makedirs(os.path.join(path, "data"))
    with open(os.path.join(path, "metadata", "part-00000"), "w") as fh:
        fh.write(json.dumps({"class": M.KMEANS_CLASS, "timestamp": 1, "sparkVersion": "3.5.1", "uid": "KMeans_x",
                             "paramMap": {"k": 2}, "defaultParamMap": {}}) + "\n")
    vec = pa.struct([("type", pa.int8()), ("size", pa.int32()), ("indices", pa.list_(pa.int32())), ("values", pa.list_(pa.float64()))])
    rows = [{"type": 0, "size": 8, "indices": [1, 7], "values": [2.5, -1.0]},
            {"type": 1, "size": None, "indices": None, "values": [float(i) for i in range(8)]}]
    pq.write_table(pa.table({"clusterIdx": pa.array([1, 0], pa.int32()), "clusterCenter": pa.array(rows, vec)}),
                   os.path.join(path, "data", "part-00000-x.snappy.parquet"), compression="snappy")
    m = M.load_kmeans_model(None, path)
    assert m.centers.tolist() == [[float(i) for i in range(8)], [0, 2.5, 0, 0, 0, 0, 0, -1.0]]
    assert (m.num_labels, m.repeats) == (1, 5)
    with open(os.path.join(path, "metadata", "part-00000"), "w") as fh:
        fh.write(json.dumps({"class": "org.apache.spark.ml.clustering.BisectingKMeansModel", "paramMap": {}, "defaultParamMap": {}}) + "\n")
    with pytest.raises(ValueError, match="expected"):
        M.load_kmeans_model(None, path)


def test_pipeline_model_layout_and_round_trip(tmp_path):
    labels = ["Asthma", "Able-Bodied", "Café, \"quoted\"", ""]
    pm = pkg.EtlPipelineModel(None, labels, [9, 7, 7, 1], [1.5, None, -3.25], 5, null_key_count=2, rows_after_filter=24)
    path = str(tmp_path / "health_kmeans_pipeline")
    pm.save(path)
    meta = _meta(path)
    assert meta["class"] == "org.apache.spark.ml.PipelineModel"
    uids = meta["paramMap"]["stageUids"]
    assert [u.split("_")[0] for u in uids] == ["StringIndexer", "OneHotEncoder", "VectorAssembler"]
    stage_dirs = sorted(os.listdir(os.path.join(path, "stages")))
    assert stage_dirs == [f"{i}_{u}" for i, u in enumerate(uids)]
    metas = [_meta(os.path.join(path, "stages", d)) for d in stage_dirs]
    assert [m["uid"] for m in metas] == uids
    assert metas[0]["paramMap"] == {"handleInvalid": "keep", "inputCol": "measure_name", "outputCol": "measure_name_index"}
    assert metas[1]["paramMap"] == {"inputCol": "measure_name_index", "outputCol": "measure_name_vec"}
    assert metas[1]["defaultParamMap"]["dropLast"] is True
    assert metas[2]["paramMap"]["inputCols"] == ["measure_name_vec"] * 5 + ["value", "lower_ci", "upper_ci"]   # k_means.py:63
    assert metas[2]["paramMap"]["handleInvalid"] == "keep" and metas[2]["paramMap"]["outputCol"] == "features"
    assert not os.path.exists(os.path.join(path, "stages", stage_dirs[2], "data"))       # a transformer: metadata only
    t0 = pq.read_table(os.path.join(path, "stages", stage_dirs[0], "data"))
    assert t0.column_names == ["labelsArray"] and t0.to_pydict()["labelsArray"] == [[labels]]
    t1 = pq.read_table(os.path.join(path, "stages", stage_dirs[1], "data"))
    assert t1.to_pydict() == {"categorySizes": [[len(labels) + 1]]}
    back = pkg.EtlPipelineModel.load(path)
    assert back.labels == labels and back.means == [1.5, None, -3.25] and back.repeats == 5
    assert back.label_counts == [9, 7, 7, 1] and back.null_key_count == 2 and back.rows_after_filter == 24
    assert back.features_dim == 5 * 4 + 3


def test_pipeline_model_rejects_other_stage_sequences(tmp_path):
    pm = pkg.EtlPipelineModel(None, ["a"], [1], [0.0, 0.0, 0.0], 1)
    path = str(tmp_path / "p")
    pm.save(path)
    meta = _meta(path)
    meta["paramMap"]["stageUids"] = meta["paramMap"]["stageUids"][:2]
    with open(os.path.join(path, "metadata", "part-00000"), "w") as fh:
        fh.write(json.dumps(meta) + "\n")
    with pytest.raises((ValueError, FileNotFoundError)):
        M.load_pipeline_model(None, path)


def test_own_reader_of_nested_model_files_matches_pyarrow():
    # the data/*.snappy.parquet files above are read with this package's own Parquet code (footer, Snappy,
    # repetition / definition levels): the same Python values pyarrow's reader gives, for every page flavour
    import io

    from pyspark_tf_gke_b200.parquet import read_nested_small

    vec = pa.struct([pa.field("type", pa.int8(), nullable=False), pa.field("size", pa.int32()),
                     pa.field("indices", pa.list_(pa.field("element", pa.int32(), nullable=False))),
                     pa.field("values", pa.list_(pa.field("element", pa.float64(), nullable=False)))])
    rows = [{"type": 1, "size": None, "indices": None, "values": [1.5, 2.5, 3.0]},
            {"type": 0, "size": 7, "indices": [1, 4], "values": [9.0, -1.0]},
            None,
            {"type": 1, "size": None, "indices": None, "values": []}]
    inner = pa.list_(pa.field("element", pa.string()))
    t = pa.table({"clusterIdx": pa.array([0, 1, 2, 3], pa.int32()), "clusterCenter": pa.array(rows, vec),
                  "labelsArray": pa.array([[["a", "b", None], []], None, [], [["é", "zz"]]], pa.list_(pa.field("element", inner))),
                  "sizes": pa.array([[31], [1, 2, 3], [], None], pa.list_(pa.field("element", pa.int32(), nullable=False))),
                  "name": pa.array(["x", None, "y", "z"])})
    for kw in (dict(compression="snappy"), dict(compression="NONE"), dict(compression="snappy", data_page_version="2.0"),
               dict(compression="snappy", use_dictionary=False), dict(compression="snappy", row_group_size=2)):
        buf = io.BytesIO()
        pq.write_table(t, buf, use_compliant_nested_type=True, **kw)
        assert read_nested_small(buf.getvalue()) == t.to_pydict(), kw
