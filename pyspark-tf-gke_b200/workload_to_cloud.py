"""The CSV-reading twin of the ETL stage, start to finish:
workloads/raw-spark/spark_checks/python_checks/spark_workload_to_cloud_k8s.py:48-154.

    read CSV (:48) -> row count (:59) -> null keys (:67-69) -> filter (:72-73) -> means + impute
    (:93-96) -> StringIndexer / OneHotEncoder / VectorAssembler with the one-hot block taken ONCE
    (:98-101, no MEASURE_NAME_WEIGHT here) -> KMeans().setK(5).setSeed(1) (:116-120) -> predictions
    (:123) -> cluster centres (:130-132) -> ClusteringEvaluator silhouette (:139-142) -> model.save /
    pipeline_model.save (:149-154)

Same steps and the same printed lines; the GCS read is a path / URL / bytes argument (the bucket
plumbing is outside SURVEY.md §8), `printSchema` / `show(5)` print the inferred column types and the
first decoded rows of the columns the stage uses."""
from __future__ import annotations

import os
from typing import Optional

from .engine import Engine
from .k_means import KEY_COL, NUMERIC_COLS, KMeansWorkload
from .kmeans import ClusteringEvaluator, KMeans
from .table import read_csv


def run(source, model_dir: Optional[str] = None, engine: Optional[Engine] = None, k: int = 5, seed: int = 1):
    """-> (pipeline_model, model, silhouette).  model_dir: where `health_kmeans_model` and
    `health_kmeans_pipeline` are written (None = the working directory, as the script does)."""
    eng = engine or Engine()
    health_df = read_csv(eng, source, NUMERIC_COLS, [KEY_COL])
    print("Dataset Schema:")
    print("root")
    for name in [KEY_COL] + NUMERIC_COLS:
        kind = "string" if name == KEY_COL else health_df.inferred_types.get(name, "double")
        print(f" |-- {name}: {kind} (nullable = true)")
    print(f"Total number of rows: {health_df.rows}")
    # the stage itself is k_means.py's, with the one-hot block used once (feature_cols = [vec] + numerics, :98)
    print("Checking for missing values in 'measure_name'...")
    wl = KMeansWorkload(eng, repeats=1, fit_kmeans=False)
    pipeline_model, _ = wl.k_means(health_df)
    print(f"Column 'measure_name' has {pipeline_model.null_key_count} missing values")
    print(f"Rows after filtering out missing 'measure_name' values: {pipeline_model.rows_after_filter}")
    print("Applying feature engineering pipeline...")
    kmeans = KMeans().setK(k).setSeed(seed)
    print("Training K-Means model...")
    model = kmeans.fit(pipeline_model)
    predictions = model.transform(pipeline_model)
    print("Sample Predictions (Cluster Assignments):")
    print(predictions[:5].cpu().tolist())
    print("Cluster Centers:")
    for center in model.clusterCenters():
        print(center)
    silhouette = ClusteringEvaluator().evaluate(pipeline_model, predictions, k=model.centers.shape[0])
    print(f"Silhouette with squared Euclidean distance = {silhouette}")
    base = model_dir or "."
    model_path, pipeline_path = os.path.join(base, "health_kmeans_model"), os.path.join(base, "health_kmeans_pipeline")
    print(f"Saving K-Means model to {model_path}")
    model.save(model_path, params={"k": kmeans.k, "seed": kmeans.seed})
    print(f"Saving K-Means pipeline to {pipeline_path}")
    pipeline_model.save(pipeline_path)
    return pipeline_model, model, silhouette
