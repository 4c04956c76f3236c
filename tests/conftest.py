"""pytest configuration: registers the `gpu` marker and puts the oracle binding (test infrastructure)
and the product's Python package on sys.path."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "gr-dtl_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")
ALIST_300 = os.path.join(GOLDEN, "n_0300_k_0152_gap_03.alist")
ALIST_100_23 = os.path.join(GOLDEN, "n_0100_k_0023_gap_10.alist")
ALIST_100_27 = os.path.join(GOLDEN, "n_0100_k_0027_gap_04.alist")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
