"""conftest of the staged copy of the reference's Python tests (oracle/ref_build/stage_pyref.sh): makes
`import quest.utils` resolve to the reference's UNMODIFIED quest/utils package in this directory while
`quest._kernels` (the pybind module the reference builds with CMake, quest/ops/setup.sh) is quest_b200._kernels.
quest/__init__.py is not staged (it imports the HF model zoo): `quest` is a bare namespace here."""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", "..", ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import quest_b200._kernels as shim  # noqa: E402

pkg = types.ModuleType("quest")
pkg.__path__ = [os.path.join(HERE, "quest")]
sys.modules["quest"] = pkg
sys.modules["quest._kernels"] = shim
pkg._kernels = shim
