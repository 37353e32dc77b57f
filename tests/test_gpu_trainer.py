"""The trainer hand-off (train_tf_ps.py:328-343,517-679): the
fused multi-step CUDA training kernel against a plain PyTorch fp32 reference of the
same network, loss and Keras-form Adam on the same batches and initial weights.
Floating point: per-step losses within 1e-4 relative and parameters within 1e-4
absolute after 60 steps (fp32 sums in a different order; observed ~1e-6)."""
import json

import numpy as np
import pytest
import torch

import pyspark_tf_gke_b200 as pkg
from pyspark_tf_gke_b200 import trainer as T
from tests.conftest import health_text

pytestmark = pytest.mark.gpu
TOL_LOSS, TOL_PARAM = 1e-4, 1e-4


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


def _split(params, C):
    sizes = list(T.LAYER_SIZES) + [C]
    out, o = [], 0
    for i, j in zip(sizes[:-1], sizes[1:]):
        out.append(params[o:o + i * j].view(i, j)); o += i * j
        out.append(params[o:o + j]); o += j
    return out


def _torch_reference(X, y, order, B, C, params, steps, lr=1e-3, b1=0.9, b2=0.999, eps=1e-7, step_rows=None):
    p = params.clone().requires_grad_(True)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    losses, accs = [], []
    for s in range(steps):
        rows = order[s * B:s * B + (B if step_rows is None else int(step_rows[s]))].long()
        W1, c1, W2, c2, W3, c3, W4, c4 = _split(p, C)
        h = torch.relu(X[rows] @ W1 + c1)
        h = torch.relu(h @ W2 + c2)
        h = torch.relu(h @ W3 + c3)
        logits = h @ W4 + c4
        prob = torch.softmax(logits, dim=1)
        pl = prob[torch.arange(rows.shape[0]), y[rows].long()]
        loss = -torch.log(torch.clamp(pl, min=1e-7)).mean()
        (g,) = torch.autograd.grad(loss, p)
        with torch.no_grad():
            t = s + 1
            m = b1 * m + (1 - b1) * g
            v = b2 * v + (1 - b2) * g * g
            lr_t = lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
            p -= lr_t * m / (torch.sqrt(v) + eps)
        losses.append(float(loss.detach())); accs.append(float((logits.argmax(1) == y[rows]).float().mean()))
    return p.detach(), losses, accs


@pytest.mark.parametrize("B,C,steps", [(64, 15, 60), (100, 5, 25), (7, 2, 10), (256, 15, 8)])
def test_train_steps_match_torch_fp32_reference(eng, B, C, steps):
    rng = np.random.default_rng(B + C)
    n = 5000
    X = torch.from_numpy((rng.random((n, 3)) * 4 - 2).astype(np.float32)).to(eng.device)
    y = torch.from_numpy(rng.integers(0, C, n).astype(np.int32)).to(eng.device)
    params0 = torch.from_numpy(T.glorot_uniform_params(C, rng)).to(eng.device)
    assert params0.numel() == eng.mlp_param_count(C)
    order = torch.from_numpy(rng.integers(0, n, steps * B).astype(np.int32)).to(eng.device)
    params = params0.clone()
    m, v = torch.zeros_like(params), torch.zeros_like(params)
    half = (steps // 2) * B   # two launches: the Adam time step carries over through step0
    l1, a1 = eng.mlp_train_steps(X, y, order[:half], B, C, params, m, v, step0=0)
    l2, a2 = eng.mlp_train_steps(X, y, order[half:], B, C, params, m, v, step0=steps // 2)
    loss, acc = torch.cat([l1, l2]).cpu().numpy(), torch.cat([a1, a2]).cpu().numpy()
    ref_p, ref_loss, ref_acc = _torch_reference(X, y, order, B, C, params0, steps)
    np.testing.assert_allclose(loss, np.array(ref_loss), rtol=TOL_LOSS)
    np.testing.assert_allclose(acc, np.array(ref_acc), atol=1.5 / B)   # an arg-max tie may flip one sample
    np.testing.assert_allclose(params.cpu().numpy(), ref_p.cpu().numpy(), atol=TOL_PARAM)
    # forward-only pass agrees too
    ev_loss, pred = eng.mlp_evaluate(X, y, params, C)
    W1, c1, W2, c2, W3, c3, W4, c4 = _split(params, C)
    logits = torch.relu(torch.relu(torch.relu(X @ W1 + c1) @ W2 + c2) @ W3 + c3) @ W4 + c4
    ref_l = torch.nn.functional.cross_entropy(logits, y.long(), reduction="none")
    np.testing.assert_allclose(ev_loss.cpu().numpy(), ref_l.cpu().numpy(), rtol=1e-4, atol=1e-5)
    assert float((pred == logits.argmax(1)).float().mean()) > 0.999


def test_ragged_steps_match_torch_fp32_reference(eng):
    # batch() before repeat(): every pass ends in a short batch that is an ordinary optimiser step (train_tf_ps.py:664-665)
    rng = np.random.default_rng(99)
    n, B, C = 300, 64, 4                                  # passes of 64, 64, 64, 64, 44 rows
    X = torch.from_numpy((rng.random((n, 3)) * 4 - 2).astype(np.float32)).to(eng.device)
    y = torch.from_numpy(rng.integers(0, C, n).astype(np.int32)).to(eng.device)
    params0 = torch.from_numpy(T.glorot_uniform_params(C, rng)).to(eng.device)
    steps = 12
    order_np, rows_np = T.batch_stream(n, min(3000, n), B, steps, np.random.default_rng(5))
    assert rows_np.tolist() == [64, 64, 64, 64, 44, 64, 64, 64, 64, 44, 64, 64]
    assert sorted(np.concatenate([order_np[i, :rows_np[i]] for i in range(5)]).tolist()) == list(range(n))   # a pass = a permutation
    order = torch.from_numpy(order_np.reshape(-1)).to(eng.device)
    params = params0.clone()
    m, v = torch.zeros_like(params), torch.zeros_like(params)
    loss, acc = eng.mlp_train_steps(X, y, order, B, C, params, m, v, step_rows=torch.from_numpy(rows_np).to(eng.device))
    ref_p, ref_loss, ref_acc = _torch_reference(X, y, order, B, C, params0, steps, step_rows=rows_np)
    np.testing.assert_allclose(loss.cpu().numpy(), np.array(ref_loss), rtol=TOL_LOSS)
    np.testing.assert_allclose(acc.cpu().numpy(), np.array(ref_acc), atol=1.5 / 44)
    np.testing.assert_allclose(params.cpu().numpy(), ref_p.cpu().numpy(), atol=TOL_PARAM)


def test_shuffle_buffer_order_semantics():
    rng = np.random.default_rng(0)
    o = T.shuffle_buffer_order(10, 3, 25, rng)
    assert sorted(o[:10].tolist()) == list(range(10)) and sorted(o[10:20].tolist()) == list(range(10))  # epochs are permutations
    assert all(o[i] <= i + 2 for i in range(10))   # an element can only appear once it has entered the 3-slot buffer
    full = T.shuffle_buffer_order(50, 50, 50, np.random.default_rng(1))
    assert sorted(full.tolist()) == list(range(50)) and full.tolist() != list(range(50))


def test_run_deep_training_on_the_default_input(eng, tmp_path):
    src = tmp_path / "health.csv"
    src.write_text(health_text("workloads"))
    model = T.run_deep_training(str(src), str(tmp_path / "out"), epochs=3, batch_size=64, engine=eng, seed=3)
    assert model.classes == 15 and model.params.numel() == eng.mlp_param_count(15)
    with open(tmp_path / "out" / "history.json") as fh:
        hist = json.load(fh)
    assert set(hist) == {"loss", "accuracy", "val_loss", "val_accuracy"} and len(hist["loss"]) == 3
    assert hist["loss"][-1] < hist["loss"][0]          # it learns something
    with open(tmp_path / "out" / "label_map.json") as fh:
        lm = json.load(fh)
    assert len(lm) == 15 and list(lm) == [str(i) for i in range(15)] and list(lm.values()) == sorted(lm.values())
    saved = np.load(tmp_path / "out" / "model.npz")
    assert saved["params"].shape == (eng.mlp_param_count(15),) and saved["layers"].tolist() == [3, 16, 32, 64, 15]
    # the 80/20 split is the reference's: NumPy default_rng(1337) over the 15355 kept rows
    idx = np.arange(15355); np.random.default_rng(1337).shuffle(idx)
    X, y, _ = pkg.load_csv_device(str(src), engine=eng)
    val_loss, val_acc = model.evaluate(X[torch.from_numpy(idx[-3071:]).to(eng.device)].contiguous(),
                                       y[torch.from_numpy(idx[-3071:]).to(eng.device)].contiguous())
    assert abs(val_loss - hist["val_loss"][-1]) < 1e-5 and abs(val_acc - hist["val_accuracy"][-1]) < 1e-6
