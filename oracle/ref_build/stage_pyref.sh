#!/usr/bin/env bash
# Stages the reference's OWN Python operator layer and tests, unmodified, next to the prebuilt reference kernels in
# oracle/_ref/pyref (git-ignored build output; travels to the GPU box like libquest_ref.so, where /root/reference does
# not exist), so that tests/test_reference_python_gpu.py can run them on top of quest_b200._kernels:
#   quest/utils/*.py   (imports quest._kernels: quest/utils/__init__.py:5, decode_wrapper.py:4)
#   quest/tests/*.py   (the reference's pytest files)
# plus a conftest.py of ours that aliases quest._kernels to the shim.  Authoring container only.
set -euo pipefail
cd "$(dirname "$0")"
REF=${QUEST_REFERENCE:-/root/reference}
OUT=../_ref/pyref
rm -rf "$OUT"
mkdir -p "$OUT/quest/utils" "$OUT/quest/tests"
cp "$REF"/quest/utils/*.py "$OUT/quest/utils/"
cp "$REF"/quest/tests/*.py "$OUT/quest/tests/"
cp pyref_conftest.py "$OUT/conftest.py"
echo "staged $(cd "$OUT" && pwd)"
