import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# every hj_accel_build of the test suite also runs the structural BVH check (k_verify, option verify_accel): each leaf
# triangle inside its slot's decoded box, child masks disjoint, every primitive in exactly one leaf
os.environ.setdefault("HJ_VERIFY_ACCEL", "1")
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def cornell(golden_dir):
    import henjou_renderer_b200 as hj
    return hj.load_npz(os.path.join(golden_dir, "cornelbox.npz"))


@pytest.fixture(scope="session")
def cornell_tex(golden_dir):
    import henjou_renderer_b200 as hj
    return hj.load_npz(os.path.join(golden_dir, "cornelbox_texture_test.npz"))
