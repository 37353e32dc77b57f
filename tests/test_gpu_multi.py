"""Multi-GPU checks: skipped on a one-GPU box, run as torchrun children when the box has >= 2 GPUs."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

from tests.conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(script, nproc, *extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, script), *extra]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_shuffle_matches_oracle():
    p = _torchrun("tests/multi_gpu/check_shuffle.py", 2)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    line = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    assert line["ok"] and all(all(v.values()) for v in line["variants"].values())
