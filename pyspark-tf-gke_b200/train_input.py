"""Drop-in for load_csv of workloads/raw-tf/train_tf_ps.py:75-149 - the TF-side
ETL whose output (X float32[N,3] C-order, y int32[N], sorted vocab) is what
tf.data.Dataset.from_tensor_slices((X, y)) consumes (train_tf_ps.py:597,664).

Same signature, same return contract, same RuntimeError on an empty result.
The CSV is decoded, filtered, label-indexed and packed on the GPU; the host
only sorts the label vocabulary (15 strings on the default input)."""
from __future__ import annotations

import json
from typing import List, Optional, Tuple

import numpy as np
import torch

from .engine import Column, Engine
from .table import read_csv

_engine: Optional[Engine] = None


def _default_engine() -> Engine:
    global _engine
    if _engine is None:
        _engine = Engine()
    return _engine


def load_csv_device(source, numeric_features: Optional[List[str]] = None, label_col: str = "subpopulation",
                    engine: Optional[Engine] = None, device_text: Optional[torch.Tensor] = None):
    """As load_csv, but X / y stay on the device (the hand-off a device-resident
    trainer would take).  -> (X f32[N,3] cuda, y i32[N] cuda, vocab).  device_text: the
    file's bytes already on the GPU (table.upload_text), to keep the copy out of a timed region."""
    eng = engine or _default_engine()
    if numeric_features is None:
        numeric_features = ["value", "lower_ci", "upper_ci"]
    if len(numeric_features) != 3:
        raise ValueError("the GPU path packs exactly three numeric features (the model's input_dim, train_tf_ps.py:579)")
    table = read_csv(eng, source, numeric_features, [label_col], trim_strings=True, nan_is_null=True, device_text=device_text)
    if table.rows == 0:
        raise RuntimeError("No valid rows were parsed from the dataset.")
    lab = table.strings[label_col]
    nums = [table.numeric[c] for c in numeric_features]
    # first pass: which labels survive the row filter (vocab = sorted(set(y_raw)), train_tf_ps.py:143)
    X, y, rows = eng.pack_tf(lab.column, nums)
    n = int(rows.item())
    if n == 0:
        raise RuntimeError("No valid rows were parsed from the dataset.")
    L = len(lab.labels)
    hist, _ = eng.code_histogram(Column(y[:n]), max(L, 1))
    hist = hist.cpu().numpy()
    kept = [i for i in range(L) if hist[i] > 0]
    vocab = [lab.labels[i] for i in kept]          # lab.labels is already sorted ascending
    remap = torch.full((max(L, 1),), -1, dtype=torch.int32)
    remap[torch.tensor(kept, dtype=torch.int64)] = torch.arange(len(kept), dtype=torch.int32)
    y = y[:n]
    eng.remap_codes(y, remap.to(eng.device), -1)
    return X[:n], y, vocab


def load_csv(source: str, numeric_features: Optional[List[str]] = None,
             label_col: str = "subpopulation") -> Tuple[np.ndarray, np.ndarray, List[str]]:
    X, y, vocab = load_csv_device(source, numeric_features, label_col)
    return X.cpu().numpy(), y.cpu().numpy(), vocab


def write_label_map(vocab: List[str], path: str) -> None:
    """label_map.json exactly as run_deep_training writes it (train_tf_ps.py:582-583)."""
    with open(path, "w", encoding="utf-8") as f:
        json.dump({int(i): s for i, s in enumerate(vocab)}, f, ensure_ascii=False, indent=2)
