"""pytest configuration: registers the `gpu` marker and puts the package + oracle on sys.path.

The importable Python host package lives in `implicitglobalgrid.jl_b200/igg_b200` (the outer directory name
contains a dot and cannot be a Python package name itself).  `oracle/` is test infrastructure only.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "implicitglobalgrid.jl_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


# Direct-to-halo is taken from 32 MiB of staged data per call by default; the tile-table tests use small arrays, and
# the library reads the variable once per process.
os.environ.setdefault("IGG_DIRECT_THRESHOLD", "0")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run with -m gpu on a B200)")
