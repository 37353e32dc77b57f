"""The C ABI driven by a plain C++ program (tests/native/cabi_caller.cpp): no Python, no torch in
that process - the shape of the JNI / cgo / C++ binding INTEGRATION.md describes.  The program
checks the device results against host loops itself; exit code 0 = identical."""
import os
import subprocess

import pytest

from tests.conftest import ROOT

pytestmark = pytest.mark.gpu


def test_native_program_links_and_matches_host_loops(tmp_path):
    exe = str(tmp_path / "cabi_caller")
    libdir = os.path.join(ROOT, "pyspark-tf-gke_b200")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.run(["g++", "-O2", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(cuda, "include"),
                    os.path.join(ROOT, "tests", "native", "cabi_caller.cpp"), "-L", libdir, "-letl_b200",
                    "-L", os.path.join(cuda, "lib64"), "-lcudart", "-o", exe], check=True)
    env = dict(os.environ, LD_LIBRARY_PATH=libdir + ":" + os.path.join(cuda, "lib64") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    for rows in ("1000003", "31", "4097"):
        r = subprocess.run([exe, rows], capture_output=True, text=True, env=env, timeout=120)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "mismatches 0" in r.stdout
