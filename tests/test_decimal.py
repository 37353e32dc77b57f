"""Host build of the product's numeral parser (csrc/decimal.cuh, the code the CSV
kernel runs per cell) against Python's float() - i.e. against correctly rounded
decimal->binary64, which is what load_csv's float(v) (train_tf_ps.py:132) and
Java's Double.parseDouble do.  CPU only."""
import ctypes as C
import os
import random
import struct
import subprocess

import numpy as np
import pytest

from tests.conftest import ROOT

SRC = os.path.join(ROOT, "tests", "native", "decimal_harness.cpp")


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("dec") / "libdec.so")
    subprocess.run(["g++", "-O2", "-shared", "-fPIC", "-o", so, SRC], check=True)
    lib = C.CDLL(so)
    lib.dec_parse_many.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def parse_many(lib, strings):
    blob = "".join(strings).encode()
    off = np.zeros(len(strings) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(s.encode()) for s in strings])
    out = np.zeros(len(strings), dtype=np.float64)
    st = np.zeros(len(strings), dtype=np.int32)
    fl = np.zeros(len(strings), dtype=np.int32)
    lib.dec_parse_many(blob, off.ctypes.data, len(strings), out.ctypes.data, st.ctypes.data, fl.ctypes.data)
    return out, st, fl


def bits(x):
    return struct.pack("<d", x)


def test_known_numerals(lib):
    ok = ["0", "8", "8.0", "-0", "1e22", "123456789012345", "12345678901234567890", "0.1", "1e-5", "16777217", ".5", "5.",
          "  12 ", "0.000", "9007199254740993", "9007199254740993.0", "0.30000000000000004", "123456789.123456789e-5",
          "1234567890123456789e-19", "+3", "-2.5", "1.5E-2", "102684", "0057", "1e0", "inf", "-Infinity", "INF"]
    out, st, fl = parse_many(lib, ok)
    for s, v, code in zip(ok, out, st):
        assert code == 0, s
        assert bits(v) == bits(float(s)), s
    out, st, fl = parse_many(lib, ["nan", "NaN", "NAN"])
    assert (st == 0).all() and np.isnan(out).all() and ((fl & 2) == 2).all()
    out, st, fl = parse_many(lib, ["", "   ", "abc", "1.2.3", ".", "1e", "+", "1e+", "--1", "1 2", "0x10", "1_000"])
    assert st.tolist() == [1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2]
    # any exponent is exact (multi-word path), through the subnormals and up to overflow
    wide = ["3.4e38", "1e23", "1.7976931348623157e308", "1.7976931348623159e308", "4.9e-324", "2.4703282292062327e-324",
            "2.4703282292062328e-324", "2.2250738585072011e-308", "1e-400", "1e400", "1234567890123456789e-300"]
    out, st, fl = parse_many(lib, wide)
    assert (st == 0).all()
    for s, v in zip(wide, out):
        assert bits(v) == bits(float(s)), s
    # 20..60 significant digits go through the multi-word mantissa, still exact
    longer = ["99999999999999999999", "0.12345678901234567891", "3.14159265358979323846264338327950288",
              "9007199254740993.000000000000000000001", "9007199254740992.999999999999999999999",
              "2.22507385850720138309023271733240406421921598046233e-308", "1" + "0" * 80]
    out, st, fl = parse_many(lib, longer)
    assert (st == 0).all()
    for s, v in zip(longer, out):
        assert bits(v) == bits(float(s)), s
    # more than 60 significant digits with a non-zero tail: reported, never approximated
    out, st, fl = parse_many(lib, ["1" * 61, "0." + "7" * 70])
    assert (st == 3).all()


def test_integer_flag_drives_schema_inference(lib):
    out, st, fl = parse_many(lib, ["8", "-15", "8.0", "1e3", "nan"])
    assert ((fl & 1) != 0).tolist() == [False, False, True, True, True]


def test_random_numerals_are_correctly_rounded(lib):
    rnd = random.Random(20261106)
    strings = []
    while len(strings) < 300_000:
        nd = rnd.choice([rnd.randint(1, 19), rnd.randint(1, 60)])
        digits = "".join(rnd.choice("0123456789") for _ in range(nd))
        k = rnd.randint(0, nd)
        s = digits[:k] + ("." if rnd.random() < 0.7 else "") + digits[k:]
        if s in ("", ".") or (s.endswith(".") and k == 0):
            continue
        if rnd.random() < 0.5:
            s += "e%d" % rnd.choice([rnd.randint(-25, 25), rnd.randint(-345, 310)])
        if rnd.random() < 0.2:
            s = "-" + s
        strings.append(s)
    out, st, fl = parse_many(lib, strings)
    assert set(st.tolist()) <= {0}          # <= 60 significant digits: always converted exactly
    exact = np.nonzero(st == 0)[0]
    ref = np.array([float(strings[i]) for i in exact])
    assert out[exact].tobytes() == ref.tobytes()


def test_health_csv_numerals(lib, health_infra_text):
    import csv, io
    cells = []
    for row in csv.DictReader(io.StringIO(health_infra_text)):
        for c in ("value", "lower_ci", "upper_ci"):
            if row[c] != "":
                cells.append(row[c])
    out, st, fl = parse_many(lib, cells)
    assert (st == 0).all() and out.tobytes() == np.array([float(c) for c in cells]).tobytes()
