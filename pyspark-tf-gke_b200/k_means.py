"""Drop-in for the ETL stage of workloads/raw-spark/k_means.py.

`KMeansWorkload.k_means(input_df)` keeps the reference's signature, log lines,
MEASURE_NAME_WEIGHT handling and error behaviour (log, then re-raise,
k_means.py:108-110); `input_df` is a DeviceTable instead of a Spark DataFrame
and every operation of lines 22-77 runs as a CUDA kernel through the C ABI.
The stage ends where the reference's does: `dataset = select("features")`
(k_means.py:77); then, as in the reference, KMeans(k=25, seed=1, maxIter=1000) is
fitted on it (k_means.py:83-87; kmeans.py here - same algorithm, its random
streams are not Spark's) and (pipeline_model, model) is returned.  Set
ETL_SKIP_KMEANS=1 to stop after the ETL stage (model = None)."""
from __future__ import annotations

import logging
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import dist as D
from .engine import Column, Engine
from .table import DeviceTable, read_csv

NUMERIC_COLS = ["value", "lower_ci", "upper_ci"]
KEY_COL = "measure_name"


def string_indexer_order(labels: Sequence[str], counts: Sequence[int]) -> List[int]:
    """StringIndexer's default stringOrderType="frequencyDesc": by count
    descending, ties by label ascending (Spark >= 3.0).  Labels compare as Java
    strings; for BMP text that equals Python's code-point order."""
    return sorted(range(len(labels)), key=lambda i: (-int(counts[i]), labels[i]))


@dataclass
class EtlPipelineModel:
    """What Pipeline.fit(...) produced (k_means.py:71-73): the StringIndexer
    labels, the imputation means and the assembler layout; transform() applies
    them to a table (k_means.py:74,77)."""

    engine: Engine
    labels: List[str]           # StringIndexerModel.labels, index order
    label_counts: List[int]
    means: List[Optional[float]]
    repeats: int
    null_key_count: int = 0
    rows_after_filter: int = 0
    # results of transforming the training table
    index: Optional[torch.Tensor] = None        # int32[n] label index per kept row
    numerics: Optional[List[torch.Tensor]] = None  # 3 x f64[n], imputed
    features_dim: int = 0

    def _index_table(self, table_labels: Sequence[str]) -> torch.Tensor:
        pos = {s: i for i, s in enumerate(self.labels)}
        L = len(self.labels)
        return torch.tensor([pos.get(s, L) for s in table_labels] or [L], dtype=torch.int32, device=self.engine.device)

    def transform(self, table: DeviceTable):
        """filter(isNotNull) + impute + index (unseen label -> numLabels) on a new
        table -> (index int32[n], numerics)."""
        eng = self.engine
        key = table.strings[KEY_COL]
        vals = [table.numeric[c] for c in NUMERIC_COLS]
        fill = [m if m is not None else float("nan") for m in self.means]
        code, nums, rows = eng.project_filter_impute(key.column, vals, fill)
        r = int(rows.item())
        code, nums = code[:r], [v[:r] for v in nums]
        eng.remap_codes(code, self._index_table(key.labels), len(self.labels))
        return code, nums

    def features(self, dtype=torch.float64, index=None, numerics=None) -> torch.Tensor:
        """Dense `features` matrix [n, repeats * L + 3] (DenseVector rows)."""
        index = self.index if index is None else index
        numerics = self.numerics if numerics is None else numerics
        return self.engine.assemble_dense(index, numerics, len(self.labels), self.repeats, dtype=dtype)

    def save(self, path: str) -> None:
        """PipelineModel.save(path): Spark ML's directory layout (mlformat.py)."""
        from . import mlformat
        mlformat.save_pipeline_model(self, path)

    @staticmethod
    def load(path: str, engine: Optional[Engine] = None) -> "EtlPipelineModel":
        from . import mlformat
        return mlformat.load_pipeline_model(engine, path)

    def features_sparse(self, index=None, numerics=None):
        """ELL form of the SparseVector rows: (indices int32[n, repeats+3], values f64[n, repeats+3])."""
        index = self.index if index is None else index
        numerics = self.numerics if numerics is None else numerics
        return self.engine.assemble_sparse(index, numerics, len(self.labels), self.repeats)


class KMeansWorkload:
    DB_CONFIG = None
    pipeline_model = None
    kmeans_model = None

    def __init__(self, engine: Optional[Engine] = None, repeats: Optional[int] = None, fit_kmeans: Optional[bool] = None):
        """repeats / fit_kmeans: explicit settings for callers inside this package (the CSV twin takes the
        one-hot block once and fits its own KMeans); None = what the reference reads from the environment
        (MEASURE_NAME_WEIGHT, k_means.py:56-61) / ETL_SKIP_KMEANS."""
        self.logger = logging.getLogger("KMeansWorkload")
        self.engine = engine
        self.repeats = repeats
        self.fit_kmeans = fit_kmeans

    def k_means(self, input_df: DeviceTable):
        try:
            eng = self.engine or input_df.engine
            key = input_df.strings[KEY_COL]
            self.logger.info("Checking for missing values in 'measure_name'...")
            # with torch.distributed initialised, input_df is this rank's shard of the rows: counts,
            # (sum, count) pairs and the label table are merged across ranks
            numeric_cols = [c for c in NUMERIC_COLS if c in input_df.numeric]
            if len(numeric_cols) != 3:
                raise KeyError(f"input needs the numeric columns {NUMERIC_COLS}")
            vals = [input_df.numeric[c] for c in numeric_cols]
            # k_means.py:23 (null keys) and k_means.py:45-48 (three avg jobs -> one pass; the isNotNull filter of
            # :27 is still in the plan) run back to back; their eight numbers cross the ranks in ONE all-reduce
            # and reach the host in ONE copy (exact as doubles: counts stay far below 2^53)
            nulls = eng.null_count(key.column)
            sums, counts = eng.sum_count(vals, row_valid=key.column.valid)
            local = torch.cat([nulls.to(torch.float64), torch.tensor([float(input_df.rows)], device=eng.device),
                               sums, counts.to(torch.float64)])
            merged = local.clone()
            if D.world()[1] > 1:
                D.all_reduce_counts(merged)
            both = torch.stack([local, merged]).cpu().tolist()
            local_nulls, local_rows = int(both[0][0]), int(both[0][1])
            null_measure_name_count, total_rows = int(both[1][0]), int(both[1][1])
            self.logger.info(f"Column 'measure_name' has {null_measure_name_count} missing values")
            rows_after = total_rows - null_measure_name_count                             # k_means.py:27-28
            self.logger.info(f"Rows after filtering out missing 'measure_name' values: {rows_after}")
            sums, counts = both[1][2:5], [int(c) for c in both[1][5:8]]
            means = [s / c if c else None for s, c in zip(sums, counts)]

            try:
                repeats = int(os.environ.get("MEASURE_NAME_WEIGHT", "5")) if self.repeats is None else int(self.repeats)  # k_means.py:56-61
            except Exception:
                repeats = 5
            if repeats < 1:
                repeats = 1
            self.logger.info(f"Applying measure_name weight by repeating measure_name_vec {repeats} time(s)")
            self.logger.info("Applying feature engineering pipeline...")

            # k_means.py:27 + :49-51: filter, impute, cast - one kernel
            fill = [m if m is not None else float("nan") for m in means]
            code, nums, _ = eng.project_filter_impute(key.column, vals, fill)
            r = local_rows - local_nulls            # rows this rank keeps: known from the null count, no read-back
            code, nums = code[:r], [v[:r] for v in nums]

            # StringIndexer.fit (k_means.py:34,73): group-by-count, then order on the host
            L0 = max(len(key.labels), 1)
            hist, oor = eng.code_histogram(Column(code), L0)
            hist = hist.cpu().tolist()
            seen = [i for i in range(len(key.labels)) if hist[i] > 0]
            names, ncounts = D.merge_label_tables([key.labels[i] for i in seen], [hist[i] for i in seen])
            order = string_indexer_order(names, ncounts)
            labels = [names[i] for i in order]
            label_counts = [ncounts[i] for i in order]
            model = EtlPipelineModel(eng, labels, label_counts, means, repeats, null_measure_name_count, rows_after)
            # StringIndexerModel.transform (k_means.py:74): code -> index
            eng.remap_codes(code, model._index_table(key.labels), len(labels))
            model.index, model.numerics = code, nums
            model.features_dim = repeats * len(labels) + 3
            skip = os.environ.get("ETL_SKIP_KMEANS", "").lower() in ("1", "true", "yes", "y") if self.fit_kmeans is None \
                else not self.fit_kmeans
            if skip:
                return model, None
            # K-Means Model Training (k_means.py:83-87)
            from .kmeans import KMeans

            kmeans = KMeans().setK(25).setSeed(1).setMaxIter(1000)
            self.logger.info("Training K-Means model...")
            kmeans_model = kmeans.fit(model)
            # the reference keeps this block commented out ("using in-memory models instead",
            # k_means.py:88-104); here it is live but off unless SAVE_MODELS is set
            if kmeans_model is not None and os.environ.get("SAVE_MODELS", "false").lower() in ("1", "true", "yes", "y"):
                model_path, pipeline_path = self._get_model_paths()
                self.logger.info(f"Saving K-Means model to {model_path}")
                kmeans_model.save(model_path, params={"k": kmeans.k, "seed": kmeans.seed, "maxIter": kmeans.maxIter})
                self.logger.info(f"Saving K-Means pipeline to {pipeline_path}")
                model.save(pipeline_path)
            return model, kmeans_model
        except Exception as e:
            self.logger.error(f"Error training K-Means model: {e}")
            raise

    # Legacy method (k_means.py:111-116)
    def _get_model_paths(self):
        base_dir = os.environ.get("MODEL_OUTPUT_DIR", "/opt/spark/work-dir/models")
        return os.path.join(base_dir, "health_kmeans_model"), os.path.join(base_dir, "health_kmeans_pipeline")

    # Legacy method (k_means.py:118-135): PipelineModel.load / KMeansModel.load from Spark ML's on-disk layout
    def load_models(self):
        from . import mlformat
        pr_model_path, pr_pipeline_path = self._get_model_paths()
        self.logger.info(f"Loading K-Means model from {pr_model_path}")
        self.logger.info(f"Loading Pipeline model from {pr_pipeline_path}")
        try:
            pr_pipeline_model = mlformat.load_pipeline_model(self.engine, pr_pipeline_path)
            self.logger.info(pr_pipeline_model.__class__.__name__)
        except Exception as e:
            self.logger.error(f"Error loading Pipeline model: {e}")
            raise
        try:
            pr_kmeans_model = mlformat.load_kmeans_model(self.engine, pr_model_path)
            self.logger.info(pr_kmeans_model.__class__.__name__)
        except Exception as e:
            self.logger.error(f"Error loading KMeans model: {e}")
            raise
        return pr_pipeline_model, pr_kmeans_model

    def infer_single_row(self, spark=None, entry_str: str = "Able-Bodied", entry_num: int = 0):
        """k_means.py:138-162: one hard-coded row (entry_str, n, n + 7, n + 5) through
        the in-memory pipeline and KMeans models -> (prediction, features row)."""
        try:
            if not KMeansWorkload.pipeline_model or not KMeansWorkload.kmeans_model:
                raise RuntimeError("In-memory models not available. Ensure k_means() has been executed before inference.")
            pm, km = KMeansWorkload.pipeline_model, KMeansWorkload.kmeans_model
            eng = pm.engine
            L = len(pm.labels)
            idx = pm.labels.index(entry_str) if entry_str in pm.labels else L   # handleInvalid="keep"
            index = torch.tensor([idx], dtype=torch.int32, device=eng.device)
            nums = [torch.tensor([float(v)], dtype=torch.float64, device=eng.device)
                    for v in (entry_num, entry_num + 7, entry_num + 5)]
            prediction = int(km.predict_structured(index, nums)[0].item())
            self.logger.info(f"Inference prediction: {prediction}")
            return prediction, pm.features(index=index, numerics=nums)
        except Exception as e:
            self.logger.error(f"Error during inference: {e}")
            raise

    @classmethod
    def main(cls, source: Optional[str] = None):
        """`python k_means.py` (k_means.py:164-212): the reference reads the table
        over JDBC; here the same table is read from the CSV it was loaded from
        (infra/local/mysql-database/load_csv.py:30,76).  Errors are logged, not
        raised (k_means.py:197-198)."""
        logging.basicConfig(level=logging.INFO)
        instance = cls()
        try:
            eng = Engine()
            instance.engine = eng
            src = source or os.environ.get("HEALTH_CSV", "health.csv")
            table = read_csv(eng, src, NUMERIC_COLS, [KEY_COL])
            cls.pipeline_model, cls.kmeans_model = instance.k_means(table)
            instance.logger.info("Running inference on a single row to verify the model is working correctly...")
            if os.environ.get("RUN_INFERENCE", "true").lower() in ("1", "true", "yes", "y") and cls.kmeans_model:
                try:  # k_means.py:187-195
                    data_strings = ["Able-Bodied", "Asthma", "Avoided Care Due to Cost", "Cancer",
                                    "Cardiovascular Diseases", "Child Poverty", "Premature Death"]
                    for label, num in zip(data_strings, [0, 10, 20, 30, 40, 50, 60]):
                        instance.infer_single_row(None, entry_str=label, entry_num=num)
                except Exception as ie:
                    instance.logger.warning(f"Single-row inference skipped due to error: {ie}")
        except Exception as e:
            instance.logger.error(f"An error occurred: {e}")
        return instance
