import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


class Golden:
    """Fixtures written by oracle/ref_driver.cpp (the unmodified reference, OMP backend)."""

    def __init__(self):
        self.scalars = {}
        for f in sorted(os.listdir(GOLDEN)):
            if f.startswith("manifest") and f.endswith(".json"):
                with open(os.path.join(GOLDEN, f)) as fh:
                    self.scalars.update(json.load(fh))

    def __getitem__(self, key):
        return self.scalars[key]

    def arr(self, name):
        return np.load(os.path.join(GOLDEN, name + ".npy"))

    def has(self, name):
        return os.path.exists(os.path.join(GOLDEN, name + ".npy"))


@pytest.fixture(scope="session")
def golden():
    return Golden()
