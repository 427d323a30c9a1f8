"""Repo hygiene the task demands: the product never touches oracle/, and nothing GPU-side reads /root/reference."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-inference-system-based-storm_b200")


def _files(top, exts):
    for d, _, fs in os.walk(top):
        if os.sep + "build" in d or "__pycache__" in d:
            continue
        for f in fs:
            if f.endswith(exts):
                yield os.path.join(d, f)


def test_product_never_uses_the_oracle():
    pat = re.compile(r"(import\s+oracle|from\s+oracle|siboracle|sib_oracle|oracle/)")
    for f in _files(PKG, (".py", ".cpp", ".cu", ".hpp", ".h", ".java")):
        for i, line in enumerate(open(f, errors="replace"), 1):
            if pat.search(line) and "oracle has its" not in line and "oracle/sib_oracle.c)" not in line:
                raise AssertionError(f"{f}:{i} references the oracle: {line.strip()}")


def test_nothing_runtime_reads_the_reference_tree():
    for f in list(_files(PKG, (".py", ".cpp", ".cu", ".hpp"))) + [os.path.join(ROOT, "bench.py"), os.path.join(ROOT, "__graft_entry__.py")]:
        if not os.path.exists(f):
            continue
        src = open(f, errors="replace").read()
        assert "/root/reference" not in src, f
