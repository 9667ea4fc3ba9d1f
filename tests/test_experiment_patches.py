"""The measured-and-dropped kernel variants kept as patches under tools/experiments/  must stay applicable to
the current sources, so that their A/B figures in profiles/r02_b_experiments.md can be re-taken."""
import glob
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCHES = sorted(glob.glob(os.path.join(ROOT, "tools", "experiments", "*.patch")))


@pytest.mark.parametrize("patch", PATCHES, ids=[os.path.basename(p) for p in PATCHES])
def test_patch_applies_to_the_current_sources(patch):
    if shutil.which("git") is None:
        pytest.skip("no git")
    r = subprocess.run(["git", "apply", "--check", patch], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
