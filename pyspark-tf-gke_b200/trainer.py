"""Device-resident hand-off to the trainer - SURVEY.md §8f rank 1: the single-process
branch of run_deep_training (workloads/raw-tf/train_tf_ps.py:517-679) with the ETL
output never leaving the GPU.

    load_csv (:577) -> label_map.json (:582-583) -> 80/20 split with
    np.random.default_rng(1337) (:653-661) -> from_tensor_slices((X, y))
    .shuffle(min(3000, N)).batch(B).repeat() (:664-665) -> build_deep_model (:328-343)
    -> model.fit(epochs, steps_per_epoch) with validation (:669-672) -> history.json (:679)

Same steps, same files (the model is saved as model.npz - the parameter blob and its
layout - instead of a .keras archive).  What cannot be reproduced without TensorFlow:
its Glorot-uniform draws and its shuffle-buffer draws; both use NumPy generators here
(seeded), so the loss curve differs from a TF run in the way two TF runs with
different seeds differ.  The split IS the reference's (NumPy, seed 1337).

The batching is the reference's: it batches BEFORE it repeats (shuffle().batch(B)
.repeat(), :664-665), so every pass over the N training rows ends in one short batch
of N mod B rows that model.fit consumes as an ordinary step, and an epoch is the next
N // B batches of that endless stream (epochs do not line up with passes).  The epoch
kernel takes the number of rows of each step (etl_mlp_train_steps_ragged); the
epoch's loss / accuracy are the per-batch means weighted by batch size, as Keras'
metric accumulators do."""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np
import torch

from .engine import Engine
from .train_input import load_csv_device, write_label_map

LAYER_SIZES = (3, 16, 32, 64)


def glorot_uniform_params(classes: int, rng: np.random.Generator) -> np.ndarray:
    """Keras' default Dense init: kernel ~ U(-l, l), l = sqrt(6 / (fan_in + fan_out)); bias 0."""
    sizes = list(LAYER_SIZES) + [classes]
    parts = []
    for fan_in, fan_out in zip(sizes[:-1], sizes[1:]):
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        parts.append(rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(np.float32).ravel())
        parts.append(np.zeros(fan_out, dtype=np.float32))
    return np.concatenate(parts)


def batch_stream(n: int, buffer_size: int, batch: int, steps: int, rng: np.random.Generator):
    """The first `steps` batches of Dataset.range(n).shuffle(buffer_size).batch(batch).repeat()
    -> (order int32[steps, batch] padded with row 0, rows int32[steps]): every pass over the
    n rows is reshuffled and cut into ceil(n / batch) batches, the last one short."""
    per_pass = -(-n // batch)
    passes = -(-steps // per_pass)
    order = np.zeros((steps, batch), dtype=np.int32)
    rows = np.zeros(steps, dtype=np.int32)
    s = 0
    for _ in range(passes):
        idx = shuffle_buffer_order(n, buffer_size, n, rng)
        for lo in range(0, n, batch):
            if s == steps:
                break
            part = idx[lo:lo + batch]
            order[s, :part.shape[0]] = part
            rows[s] = part.shape[0]
            s += 1
    return order, rows


def shuffle_buffer_order(n: int, buffer_size: int, count: int, rng: np.random.Generator) -> np.ndarray:
    """Indices produced by Dataset.range(n).shuffle(buffer_size).repeat() - tf.data's
    shuffle-buffer semantics (fill a buffer, emit a random slot, refill it with the
    next element; each pass drains its own buffer)."""
    out = np.empty(count, dtype=np.int32)
    k = 0
    while k < count:
        buf = list(range(min(buffer_size, n)))
        nxt = len(buf)
        while buf and k < count:
            j = int(rng.integers(len(buf)))
            out[k] = buf[j]
            k += 1
            if nxt < n:
                buf[j] = nxt
                nxt += 1
            else:
                buf[j] = buf[-1]
                buf.pop()
    return out


@dataclass
class DeepModel:
    """The trained 3-16-32-64-softmax network: one float32 parameter blob on the device."""

    engine: Engine
    classes: int
    params: torch.Tensor
    history: Dict[str, List[float]] = field(default_factory=dict)

    def predict(self, X: torch.Tensor) -> torch.Tensor:
        return self.engine.mlp_evaluate(X, None, self.params, self.classes)[1]

    def evaluate(self, X: torch.Tensor, y: torch.Tensor):
        loss, pred = self.engine.mlp_evaluate(X, y, self.params, self.classes)
        return float(loss.mean()), float((pred == y).float().mean())

    def save(self, path: str) -> None:
        np.savez(path, params=self.params.cpu().numpy(), classes=self.classes, layers=np.array(LAYER_SIZES + (self.classes,)))


def fit(eng: Engine, X: torch.Tensor, y: torch.Tensor, classes: int, epochs: int, batch_size: int,
        X_val: Optional[torch.Tensor] = None, y_val: Optional[torch.Tensor] = None, seed: int = 0,
        lr: float = 1e-3) -> DeepModel:
    """model.fit(ds_train, epochs, steps_per_epoch, validation_data) on device-resident arrays:
    one kernel launch per epoch (all of its steps), one forward pass for validation."""
    n = X.shape[0]
    steps = max(1, n // batch_size)
    rng = np.random.default_rng(seed)
    params = torch.from_numpy(glorot_uniform_params(classes, rng)).to(eng.device)
    m, v = torch.zeros_like(params), torch.zeros_like(params)
    order, step_rows = batch_stream(n, min(3000, n), batch_size, epochs * steps, np.random.default_rng(seed + 1))
    d_order = torch.from_numpy(order.reshape(-1)).to(eng.device)
    d_rows = torch.from_numpy(step_rows).to(eng.device)
    weight = d_rows.to(torch.float32)
    hist: Dict[str, List[float]] = {"loss": [], "accuracy": []}
    if X_val is not None:
        hist.update({"val_loss": [], "val_accuracy": []})
    model = DeepModel(eng, classes, params, hist)
    per = steps * batch_size
    for e in range(epochs):
        w = weight[e * steps:(e + 1) * steps]
        loss, acc = eng.mlp_train_steps(X, y, d_order[e * per:(e + 1) * per], batch_size, classes, params, m, v, step0=e * steps,
                                        lr=lr, step_rows=d_rows[e * steps:(e + 1) * steps])
        hist["loss"].append(float((loss * w).sum() / w.sum()))
        hist["accuracy"].append(float((acc * w).sum() / w.sum()))
        if X_val is not None:
            vl, va = model.evaluate(X_val, y_val)
            hist["val_loss"].append(vl)
            hist["val_accuracy"].append(va)
    return model


def run_deep_training(data_source: str, output_dir: str, epochs: int, batch_size: int, engine: Optional[Engine] = None,
                      seed: int = 0) -> DeepModel:
    """train_tf_ps.py:517 (single-process branch): writes label_map.json, model.npz, history.json."""
    os.makedirs(output_dir, exist_ok=True)
    print(f"Loading dataset from: {data_source}")
    X, y, label_vocab = load_csv_device(data_source, engine=engine)
    eng = engine or X_engine(X)
    num_classes = int(y.max().item()) + 1
    write_label_map(label_vocab, os.path.join(output_dir, "label_map.json"))
    print("Running single-process (no distributed strategy).")
    n = X.shape[0]
    idx = np.arange(n)
    np.random.default_rng(1337).shuffle(idx)                      # train_tf_ps.py:655-657
    val_size = max(1, int(n * 0.2))
    train_idx = torch.from_numpy(idx[:-val_size]).to(X.device)
    val_idx = torch.from_numpy(idx[-val_size:]).to(X.device)
    model = fit(eng, X[train_idx].contiguous(), y[train_idx].contiguous(), num_classes, epochs, batch_size,
                X[val_idx].contiguous(), y[val_idx].contiguous(), seed=seed)
    save_path = os.path.join(output_dir, "model.npz")
    model.save(save_path)
    print(f"Model saved to: {save_path}")
    with open(os.path.join(output_dir, "history.json"), "w") as fh:
        json.dump(model.history, fh)
    return model


def X_engine(X: torch.Tensor) -> Engine:
    from .train_input import _default_engine

    return _default_engine()
