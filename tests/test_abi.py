"""CPU: the C-ABI library loads and exports exactly the symbols include/quest_b200.h declares; host-only
entry points (plan, workspace size) behave; ops refuse to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

import quest_b200._kernels as K

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "quest_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(quest_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_table_agree():
    assert _header_symbols() == sorted(K.SIGNATURES)


def test_library_exports_every_declared_symbol():
    lib = K.load_library()
    for name in _header_symbols():
        assert hasattr(lib, name), name
    assert lib.quest_version().decode().startswith("quest_b200") and b"sm_100a" in lib.quest_version()


def test_library_is_sm100a_only():
    out = os.popen(f"cuobjdump -lelf {K._LIB_PATH} 2>/dev/null").read()
    if not out:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out and not re.search(r"sm_(?!100a)\d+", out)


def test_decode_plan_host_only():
    lib = K.load_library()
    ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
    assert lib.quest_decode_plan(127, 32, 32, 128, 16, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
    assert ppc.value * nch.value >= 128 and (nch.value - 1) * ppc.value < 128
    assert 32 * nch.value <= 148 * 7  # one resident wave (7 CTAs of 128 threads per SM)
    # reserved | tickets | layer step's bin totals (4 KB per head) | (m, d) pairs | fp32 partial outputs, each section
    # 256-byte aligned (csrc/common.cuh)
    assert lib.quest_decode_workspace_bytes(32, 128, nch.value) == 4096 + 256 + 32 * 4096 + 32 * nch.value * 2 * 4 + 32 * nch.value * 128 * 4
    # tiny problems collapse to one chunk; bad dims are refused
    assert lib.quest_decode_plan(0, 32, 32, 128, 16, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0 and nch.value == 1
    assert lib.quest_decode_plan(10, 32, 32, 96, 16, 148, ctypes.byref(ppc), ctypes.byref(nch)) == -2
    assert lib.quest_topk_max_len() >= 65536


def test_debug_options_are_explicit_and_restored():
    """Test / A-B switches are set through the ABI (the library never reads the environment) and hand back the old value."""
    lib = K.load_library()
    assert lib.quest_debug_set_option(K.DBG_PREFILL_IMPL, 3) == 0
    with K.debug_option(K.DBG_PREFILL_IMPL, 2):
        assert lib.quest_debug_set_option(K.DBG_PREFILL_IMPL, 2) == 2
    assert lib.quest_debug_set_option(K.DBG_PREFILL_IMPL, 0) == 3
    assert lib.quest_debug_set_option(99, 1) == -1
    src = "".join(open(os.path.join(ROOT, "quest_b200", "csrc", f)).read() for f in os.listdir(os.path.join(ROOT, "quest_b200", "csrc"))
                  if f.endswith((".cu", ".cuh")))
    assert "getenv" not in src


def test_argument_errors_without_touching_the_gpu():
    lib = K.load_library()
    assert lib.quest_estimate(None, None, 4, None, None, 32, 32, 128, 16, 0, None) == -1
    assert lib.quest_topk(None, None, 0, None, None, 32, 10, 3, None) == -1
    assert lib.quest_append_decode(None, None, None, None, 1, 1, None, None, 1, 1, 32, 128, 16, None) == -1


def test_ops_fail_loudly_on_cpu_tensors():
    q = torch.zeros(1, 4, 128, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        K.apply_rope_in_place(q, q.clone(), 0, 1.0, 1e4)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        K.topk_filtering(torch.zeros(4, 8, dtype=torch.float16), torch.zeros(4, 8, dtype=torch.int32),
                         torch.zeros(4, 2, dtype=torch.float16), torch.zeros(4, 2, dtype=torch.int32), None, 2)
    w = K.BatchDecodeWithPagedKVCachePyTorchWrapper(0)
    with pytest.raises(RuntimeError, match="before begin_forward"):
        w.forward(q, q, q, q, q, 1, 0, 1.0, 1e4)
    with pytest.raises(RuntimeError, match="NHD"):
        K.BatchDecodeWithPagedKVCachePyTorchWrapper(1)


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "quest_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert "oracle" not in open(os.path.join(dirpath, f)).read(), f
