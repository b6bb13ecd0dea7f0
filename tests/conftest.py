import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# Engines draw their weight-init / dropout seed from the OS unless told otherwise; the tests pin it so that a run
# is reproducible to the bit (a different dropout mask per run made one borderline tolerance fail once in ~20 runs).
os.environ.setdefault("FNET_SEED", "20261106")

import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # keep the allocation-free entry point's test at the very end of a run
    items.sort(key=lambda it: 1 if "test_zz_" in it.nodeid else 0)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
