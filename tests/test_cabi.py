"""The C-ABI library loads and exports every symbol include/etl_b200.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

import pyspark_tf_gke_b200 as pkg
from tests.conftest import ROOT


def _declared():
    with open(os.path.join(ROOT, "include", "etl_b200.h")) as fh:
        text = fh.read()
    return sorted(set(re.findall(r"ETL_API\s+[\w\s\*]+?\b(etl_\w+)\s*\(", text)))


def test_header_declares_entry_points():
    names = _declared()
    assert "etl_create" in names and "etl_project_filter_impute" in names
    assert len(names) >= 9


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in etl_b200.h but not exported"


def test_binding_covers_every_declared_symbol():
    assert sorted(pkg._lib.exported_names()) == _declared()
    lib = pkg._lib.load()
    assert lib.etl_abi_version() >= 1
    assert lib.etl_last_error() is not None


def test_product_does_not_import_oracle():
    pkg_dir = os.path.join(ROOT, "pyspark-tf-gke_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, f)) as fh:
                    src = fh.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "liboracle" not in src, f


def test_bad_arguments_are_rejected_without_a_gpu():
    lib = pkg._lib.load()
    assert lib.etl_reserve(None, 0) == -1
    assert b"NULL" in lib.etl_last_error()


def _declared_params():
    """name -> list of parameter declarations, from the header text."""
    with open(os.path.join(ROOT, "include", "etl_b200.h")) as fh:
        text = re.sub(r"/\*.*?\*/", " ", fh.read(), flags=re.S)
    out = {}
    for m in re.finditer(r"ETL_API\s+[\w\s\*]+?\b(etl_\w+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        params = [p.strip() for p in m.group(2).split(",")]
        out[m.group(1)] = [] if params == ["void"] or params == [""] else params
    return out


def test_binding_prototypes_match_the_header():
    # every ctypes prototype has as many arguments as the declaration, pointers where the header has pointers /
    # arrays, and 64-bit integers where the header says int64_t (a drifted prototype corrupts a call silently)
    decl = _declared_params()
    assert sorted(decl) == _declared()
    for name, (_, args) in pkg._lib._PROTOS.items():
        params = decl[name]
        assert len(args) == len(params), f"{name}: {len(args)} ctypes arguments, {len(params)} in etl_b200.h"
        for a, p in zip(args, params):
            is_ptr = "*" in p or "[" in p
            if is_ptr:
                assert a in (ctypes.c_void_p, ctypes.c_char_p) or issubclass(a, (ctypes._Pointer, ctypes.Array)), (name, p, a)
            elif re.match(r"^(const\s+)?int64_t\b", p):
                assert a is ctypes.c_int64, (name, p, a)
            elif re.match(r"^(const\s+)?int32_t\b", p) or re.match(r"^(const\s+)?int\b", p):
                assert a in (ctypes.c_int32, ctypes.c_int), (name, p, a)
            elif re.match(r"^(const\s+)?double\b", p):
                assert a is ctypes.c_double, (name, p, a)


def test_header_is_plain_c_and_cxx(tmp_path):
    # the boundary is a C ABI: the header must compile as C99 and as C++17 with nothing but the standard headers,
    # and the torch-free native caller must compile against it (it is built and RUN by tests/test_gpu_native_caller.py)
    import shutil
    import subprocess

    hdr = os.path.join(ROOT, "include", "etl_b200.h")
    if shutil.which("gcc"):
        subprocess.run(["gcc", "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Werror", hdr], check=True)
    if shutil.which("g++"):
        subprocess.run(["g++", "-fsyntax-only", "-x", "c++", "-std=c++17", "-Wall", "-Werror", hdr], check=True)
        caller = os.path.join(ROOT, "tests", "native", "cabi_caller.cpp")
        cuda_inc = "/usr/local/cuda/include"
        if os.path.isdir(cuda_inc):
            subprocess.run(["g++", "-fsyntax-only", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", cuda_inc, caller],
                           check=True)
