"""pytest configuration: registers the `gpu` marker and puts the repo root on sys.path.

`-m "not gpu"` : oracle vs the reference's property / known-answer tests, host logic, C-ABI
                 symbol export, gloo world_size-2 decomposition tests.  Runs on CPU.
`-m gpu`       : parity tests proper -- the CUDA path called through the C-ABI against the oracle.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")
