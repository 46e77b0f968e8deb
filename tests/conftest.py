import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def small_run():
    """8 events x 4 cams x 40 frames at 264x420: one event of every class."""
    from autobub3hs_b200 import synth
    return synth.make_run(8, ["normal", "multi", "notrigger", "late", "short4", "short5", "flicker", "highnoise"],
                          rows=264, cols=420, seed=3)


@pytest.fixture(scope="session")
def small_trained(small_run):
    import oracle.cv2_oracle as o
    return o.train_all(small_run.events, 4)
