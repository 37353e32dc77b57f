#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/.

Run ONCE in the build container, where /root/reference is mounted (it does not
exist on the GPU box):

    python tests/golden/make_golden.py

What it writes (all small, all committed):

  health_workloads.csv.gz / health_infra.csv.gz
      gzip of the two copies of the reference's default input
      (workloads/raw-spark/spark_checks/python_checks/health.csv and
      infra/local/mysql-database/datasets/csvs/health.csv).  These are the
      INPUT half of the config-1 golden vectors; nothing under tests/ reads
      /root/reference at run time.
  load_csv_health.npz
      X/y/vocab returned by the REFERENCE's own load_csv
      (workloads/raw-tf/train_tf_ps.py:75-149) imported read-only with
      tensorflow/matplotlib stubbed (they are only used by other functions).
  load_csv_cases.json
      small hand-written CSV texts (quoted commas/newlines, "nan", blanks,
      exponents, signs, ragged rows ...) with the reference load_csv's output
      on each, or the exception type it raised.
  spark_known_answers.json
      SURVEY.md §2.3's known answers for the Spark side, recomputed here with
      stdlib csv + fractions (exact rational arithmetic), NOT by PySpark - no
      JVM exists in this image, so the Spark side stays "parity unpinned".
"""
import csv
import gzip
import io
import json
import os
import sys
import tempfile
from collections import Counter
from fractions import Fraction
from unittest import mock

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
CSV_WORKLOADS = f"{REF}/workloads/raw-spark/spark_checks/python_checks/health.csv"
CSV_INFRA = f"{REF}/infra/local/mysql-database/datasets/csvs/health.csv"


def import_reference_load_csv():
    sys.modules.setdefault("tensorflow", mock.MagicMock())
    sys.modules.setdefault("matplotlib", mock.MagicMock())
    sys.modules.setdefault("matplotlib.pyplot", mock.MagicMock())
    sys.path.insert(0, f"{REF}/workloads/raw-tf")
    import train_tf_ps  # noqa: E402  (the reference, unmodified)

    return train_tf_ps.load_csv


CASES = {
    "basic": "subpopulation,value,lower_ci,upper_ci\nA,1,2,3\nB,4.5,5,6\nA,7,8,9\n",
    "blank_label_and_blank_numeric": "subpopulation,value,lower_ci,upper_ci\n,1,2,3\nB,,5,6\nC,1,2,\nD,1,2,3\n",
    "nan_text": "subpopulation,value,lower_ci,upper_ci\nA,nan,2,3\nB,NaN,2,3\nC,NAN,1,1\nD,1,2,3\n",
    "whitespace": "subpopulation,value,lower_ci,upper_ci\n  A  , 1 ,2 , 3\nB,   ,2,3\n   ,1,2,3\n",
    "quoted_commas": 'x,subpopulation,value,lower_ci,upper_ci\n"a,b",A,1,2,3\n"c,""d"",e",B,4,5,6\n',
    "quoted_newline": 'x,subpopulation,value,lower_ci,upper_ci\n"line1\nline2",A,1,2,3\n"z",B,4,5,6\n',
    "exponent_sign": "subpopulation,value,lower_ci,upper_ci\nA,1e3,-2.5,+3\nB,1.5E-2,0.000,12345678\n",
    "invalid_numeric": "subpopulation,value,lower_ci,upper_ci\nA,abc,2,3\nB,1,2,3\nC,1.2.3,2,3\n",
    "ragged_short": "subpopulation,value,lower_ci,upper_ci\nA,1,2\nB,1,2,3\n",
    "ragged_long": "subpopulation,value,lower_ci,upper_ci\nA,1,2,3,extra,more\nB,4,5,6\n",
    "no_trailing_newline": "subpopulation,value,lower_ci,upper_ci\nA,1,2,3\nB,4,5,6",
    "crlf": "subpopulation,value,lower_ci,upper_ci\r\nA,1,2,3\r\nB,4,5,6\r\n",
    "all_invalid": "subpopulation,value,lower_ci,upper_ci\n,1,2,3\nB,,2,3\n",
    "header_only": "subpopulation,value,lower_ci,upper_ci\n",
    "label_order": "subpopulation,value,lower_ci,upper_ci\nzeta,1,2,3\nAlpha,1,2,3\nbeta,1,2,3\nAlpha,2,2,2\n",
    "column_order": "upper_ci,value,junk,subpopulation,lower_ci\n3,1,q,A,2\n6,4,q,B,5\n",
    "big_and_small": "subpopulation,value,lower_ci,upper_ci\nA,123456789012345,0.1,1e-5\nB,16777217,3.4e38,1e22\n",
    "inf_text": "subpopulation,value,lower_ci,upper_ci\nA,inf,1,2\nB,-Infinity,1,2\n",
    "blank_lines": "subpopulation,value,lower_ci,upper_ci\n\nA,1,2,3\n\n\nB,4,5,6\n",
    # only the bare word "nan" drops a row (v.lower() == "nan", train_tf_ps.py:129); a signed one is a NaN VALUE
    "signed_nan": "subpopulation,value,lower_ci,upper_ci\nA,-nan,2,3\nB,+NaN,2,3\nC,nan,2,3\nD,1,2,3\n",
}


def run_cases(load_csv):
    out = {}
    for name, text in CASES.items():
        with tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False, newline="") as fh:
            fh.write(text)
            path = fh.name
        try:
            X, y, vocab = load_csv(path)
            out[name] = {
                "text": text,
                "X": [[float(v) for v in row] for row in X],
                "X_hex": [[float(v).hex() for v in row] for row in X],
                "y": [int(v) for v in y],
                "vocab": list(vocab),
            }
        except Exception as e:  # the reference raises RuntimeError on empty
            out[name] = {"text": text, "raises": type(e).__name__, "message": str(e)}
        finally:
            os.unlink(path)
    return out


def spark_known_answers(path):
    """k_means.py:22-77 restated with exact rationals."""
    with open(path, newline="", encoding="utf-8") as fh:
        rows = list(csv.DictReader(fh))
    key = [r["measure_name"] if r["measure_name"] != "" else None for r in rows]
    null_keys = sum(k is None for k in key)
    kept = [r for r, k in zip(rows, key) if k is not None]
    means = {}
    for c in ("value", "lower_ci", "upper_ci"):
        vals = [Fraction(r[c]) for r in kept if r[c] != ""]
        s = sum(vals)
        means[c] = {
            "sum": str(s),
            "count": len(vals),
            "mean_hex": (float(s) / len(vals)).hex(),
            "mean": float(s) / len(vals),
        }
    counts = Counter(r["measure_name"] for r in kept)
    labels = sorted(counts, key=lambda s: (-counts[s], s))
    return {
        "rows": len(rows),
        "null_measure_name": null_keys,
        "rows_after_filter": len(kept),
        "means": means,
        "labels": labels,
        "label_counts": [counts[s] for s in labels],
        "features_dim_repeats5": 5 * len(labels) + 3,
        "features_dim_repeats1": len(labels) + 3,
    }


def main():
    load_csv = import_reference_load_csv()
    for src, name in ((CSV_WORKLOADS, "health_workloads"), (CSV_INFRA, "health_infra")):
        with open(src, "rb") as fh:
            raw = fh.read()
        buf = io.BytesIO()
        with gzip.GzipFile(fileobj=buf, mode="wb", mtime=0, compresslevel=9) as gz:
            gz.write(raw)
        with open(os.path.join(HERE, f"{name}.csv.gz"), "wb") as fh:
            fh.write(buf.getvalue())

    Xw, yw, vw = load_csv(CSV_WORKLOADS)
    Xi, yi, vi = load_csv(CSV_INFRA)
    assert Xw.tobytes() == Xi.tobytes() and yw.tobytes() == yi.tobytes() and vw == vi
    np.savez_compressed(
        os.path.join(HERE, "load_csv_health.npz"), X=Xw, y=yw, vocab=np.array(vw, dtype=object).astype(str)
    )
    with open(os.path.join(HERE, "load_csv_cases.json"), "w") as fh:
        json.dump(run_cases(load_csv), fh, indent=1, sort_keys=True)
    ka = {"workloads": spark_known_answers(CSV_WORKLOADS), "infra": spark_known_answers(CSV_INFRA)}
    with open(os.path.join(HERE, "spark_known_answers.json"), "w") as fh:
        json.dump(ka, fh, indent=1, sort_keys=True)
    print("X", Xw.shape, Xw.dtype, "y", yw.shape, yw.dtype, "classes", len(vw))
    print("spark:", {k: ka["workloads"][k] for k in ("rows", "null_measure_name", "rows_after_filter")})
    print("means:", {c: v["mean"] for c, v in ka["workloads"]["means"].items()})


if __name__ == "__main__":
    main()
