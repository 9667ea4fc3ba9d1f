import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_sessionstart(session):
    """QUEST_TEST_DBG="10=1,4=1": run the session with library A/B switches (include/quest_b200.h QUEST_DBG_*) set from the start --
    how a candidate default is put under the whole suite before it becomes the default.  Unset: nothing happens."""
    spec = os.environ.get("QUEST_TEST_DBG", "")
    if spec:
        import quest_b200._kernels as K
        for kv in filter(None, spec.split(",")):
            o, v = kv.split("=")
            K.load_library().quest_debug_set_option(int(o), int(v))


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
