"""pytest configuration: registers the `gpu` marker and puts the package on sys.path.

The package directory is `monodepth2.jl_b200/` (not an importable name because of the
dot), so the importable Python package `mdb200` lives inside it.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")
