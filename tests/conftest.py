import gzip
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def health_text(which: str = "workloads") -> str:
    with gzip.open(os.path.join(GOLDEN, f"health_{which}.csv.gz"), "rb") as fh:
        return fh.read().decode("utf-8")


@pytest.fixture(scope="session")
def health_workloads_text():
    return health_text("workloads")


@pytest.fixture(scope="session")
def health_infra_text():
    return health_text("infra")
