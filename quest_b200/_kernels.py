"""Drop-in for the reference's pybind module ``quest._kernels`` (reference quest/ops/csrc/bsk_ops.cu:4-20).

Same seven function names, same argument order and meaning, same ``RuntimeError`` behaviour on shape /
dtype violations (the reference's ``TORCH_CHECK`` macros, quest/ops/csrc/pytorch_extension_utils.h:37-66),
plus the ``BatchDecodeWithPagedKVCachePyTorchWrapper`` class.  Everything forwards raw device pointers to
``libquest_b200.so`` (C ABI in include/quest_b200.h) on torch's current stream.  There is no CPU path:
importing works anywhere (symbol checks), calling an op without the library or with CPU tensors raises.
"""
import ctypes
import os
from typing import Optional

import torch

# QUEST_B200_LIB selects another build of the same library (e.g. the instrumented libquest_b200_dbg.so of tools/)
_LIB_PATH = os.environ.get("QUEST_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libquest_b200.so")

_c = ctypes
_vp, _u32, _i32p, _f32, _sz = _c.c_void_p, _c.c_uint32, _c.c_void_p, _c.c_float, _c.c_size_t

# name -> (restype, argtypes); mirrors include/quest_b200.h one to one.
SIGNATURES = {
    "quest_version": (_c.c_char_p, []),
    "quest_rope_inplace": (_c.c_int, [_vp, _vp, _u32, _u32, _u32, _u32, _u32, _f32, _f32, _vp]),
    "quest_rms_norm": (_c.c_int, [_vp, _vp, _vp, _u32, _u32, _f32, _vp]),
    "quest_append_decode": (_c.c_int, [_vp, _vp, _vp, _i32p, _u32, _u32, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _vp]),
    "quest_append_prefill": (_c.c_int, [_vp, _vp, _u32, _vp, _i32p, _u32, _u32, _vp, _i32p, _u32, _u32, _u32, _u32, _vp]),
    "quest_estimate": (_c.c_int, [_vp, _vp, _u32, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _vp]),
    "quest_topk_max_len": (_u32, []),
    "quest_topk": (_c.c_int, [_vp, _i32p, _u32, _vp, _i32p, _u32, _u32, _u32, _vp]),
    "quest_decode_plan": (_c.c_int, [_u32, _u32, _u32, _u32, _u32, _u32, _c.POINTER(_u32), _c.POINTER(_u32)]),
    "quest_decode_workspace_bytes": (_sz, [_u32, _u32, _u32]),
    "quest_decode_run": (_c.c_int, [_vp, _vp, _vp, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_decode_run_partial": (_c.c_int, [_vp, _vp, _vp, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_decode_static_scratch_bytes": (_sz, []),
    "quest_decode_run_static": (_c.c_int, [_vp, _vp, _vp, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp]),
    "quest_sparse_decode_step": (_c.c_int, [_vp, _vp, _vp, _i32p, _u32, _u32, _u32, _vp, _i32p, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_decode_layer_step": (_c.c_int, [_vp, _vp, _vp, _vp, _vp, _i32p, _u32, _u32, _u32, _vp, _i32p, _u32, _u32, _vp, _i32p, _u32,
                                           _u32, _u32, _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_decode_layer_step_batch": (_c.c_int, [_vp, _vp, _vp, _vp, _vp, _u32, _sz, _sz, _u32, _vp, _sz, _i32p, _u32, _u32, _u32,
                                                 _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_rope_decode_batch": (_c.c_int, [_vp, _vp, _vp, _u32, _u32, _u32, _u32, _f32, _f32, _vp]),
    "quest_seq_desc_advance": (_c.c_int, [_vp, _u32, _u32, _vp]),
    "quest_prefill": (_c.c_int, [_vp, _vp, _u32, _vp, _i32p, _u32, _u32, _c.c_int, _u32, _u32, _u32, _u32, _vp]),
    "quest_rope_append_prefill": (_c.c_int, [_vp, _vp, _vp, _u32, _vp, _i32p, _u32, _u32, _vp, _i32p, _u32, _u32, _u32, _u32, _u32,
                                             _f32, _f32, _u32, _vp]),
    "quest_prefill_layer_step": (_c.c_int, [_vp, _vp, _vp, _vp, _u32, _vp, _i32p, _u32, _u32, _vp, _i32p, _u32, _u32, _u32, _u32,
                                            _u32, _f32, _f32, _u32, _vp]),
    "quest_metadata_to_fp8": (_c.c_int, [_vp, _i32p, _u32, _u32, _vp, _u32, _u32, _u32, _vp]),
    "quest_estimate_fp8": (_c.c_int, [_vp, _vp, _u32, _vp, _i32p, _u32, _u32, _u32, _u32, _u32, _vp]),
    "quest_decode_layer_step_fp8": (_c.c_int, [_vp, _vp, _vp, _vp, _vp, _i32p, _u32, _u32, _u32, _vp, _vp, _i32p, _u32, _u32, _vp,
                                               _i32p, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_estimate_select": (_c.c_int, [_vp, _vp, _vp, _vp, _i32p, _u32, _u32, _vp, _i32p, _u32, _u32, _u32, _vp, _vp, _u32, _i32p,
                                         _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _sz, _vp]),
    "quest_merge_states": (_c.c_int, [_vp, _vp, _u32, _u32, _vp, _vp, _u32, _u32, _vp]),
    "quest_peer_staging_bytes": (_sz, [_u32, _u32, _u32]),
    "quest_peer_allgather": (_c.c_int, [_vp, _u32, _u32, _vp, _vp, _vp, _vp, _u32, _u32, _u32, _vp]),
    "quest_device_error": (_u32, []),
    "quest_device_error_clear": (None, []),
    "quest_debug_set_option": (_c.c_int, [_c.c_int, _c.c_int]),
    "quest_launch_count": (_c.c_uint64, []),
    "quest_launch_count_reset": (None, []),
}

_lib = None


def load_library():
    """Load libquest_b200.so and bind every symbol of include/quest_b200.h.  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(
                f"{_LIB_PATH} not found: build it with quest_b200/csrc/build.sh (no CPU fallback exists)")
        lib = ctypes.CDLL(_LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the .so is stale
            fn.restype = res
            fn.argtypes = args
        if torch.cuda.is_available():
            lib.quest_device_error()  # allocates the host-mapped error words now, outside any stream capture
        _lib = lib
    return _lib


# include/quest_b200.h QUEST_DBG_*: explicit test / A-B switches (the library never reads the environment)
DBG_PDL_OFF, DBG_EST_ROWS, DBG_PREFILL_IMPL, DBG_PREFILL_P_SMEM, DBG_TOPK_NO_CLUSTER, DBG_TOPK_GENERIC, DBG_TOPK_NO_HIST, DBG_EST_PER_Q, DBG_NO_HANDOFF, DBG_NO_ROW_COMPACT = range(10)
DEVERR_PREFILL, DEVERR_PEER, DEVERR_STEP = 1, 2, 4


class debug_option:
    """``with debug_option(DBG_PREFILL_IMPL, 3): ...`` -- sets a library switch and restores the previous value."""

    def __init__(self, option, value):
        self.option, self.value = int(option), int(value)

    def __enter__(self):
        self.old = load_library().quest_debug_set_option(self.option, self.value)
        return self

    def __exit__(self, *exc):
        load_library().quest_debug_set_option(self.option, self.old)
        return False


def _check(cond, msg):
    if not cond:
        raise RuntimeError(msg)


# The checks below run on every operator call of the op-by-op path (16 tensors per layer step): messages are only built on
# failure (an f-string argument would be formatted on every call; tools/four_ops_profile.py).
def _check_input(x, name):
    if not isinstance(x, torch.Tensor):
        raise RuntimeError(f"{name} must be a tensor")
    if not x.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not x.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")


def _check_rows(x, name):
    """2-D index tensor: contiguous, or one row broadcast to all heads (stride 0 view)."""
    if not (isinstance(x, torch.Tensor) and x.is_cuda):
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if x.dim() != 2:
        raise RuntimeError(f"{name} must be a 2D tensor")
    rows, cols = x.shape
    s0, s1 = x.stride()
    if not ((cols <= 1 or s1 == 1) and (rows <= 1 or cols == 0 or s0 in (0, cols))):
        raise RuntimeError(f"{name} must be contiguous")
    return 0 if (rows > 1 and s0 == 0) else cols


def _check_half(x, op):
    if x.dtype != torch.float16:
        raise RuntimeError(f"{op} failed to dispatch with dtype {x.dtype}")


# torch.cuda.current_stream() builds a Stream object behind several device-index lookups (about 5 us a call, four calls per
# layer on the op-by-op path); the raw handle of the current device's current stream is one C call.
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_cur_device = getattr(torch._C, "_cuda_getDevice", None)


def _stream():
    if _raw_stream is not None and _cur_device is not None:
        return _raw_stream(_cur_device())
    return torch.cuda.current_stream().cuda_stream


def _rc(code, op):
    if code != 0:
        if code > 0:
            raise RuntimeError(f"{op} failed with CUDA error code {code}")
        if code == -4:
            raise RuntimeError(f"{op} refused: an earlier kernel of this library gave up on a bounded wait and its output "
                               f"was not valid (quest_device_error() = {load_library().quest_device_error()}); "
                               "quest_device_error_clear() re-arms the entry point")
        if code == -5:
            raise RuntimeError(f"{op} refused: the decode plan needs more scratch than the library's built-in array "
                               f"({load_library().quest_decode_static_scratch_bytes()} bytes); raise pages_per_chunk or use the "
                               "workspace-taking entry point")
        raise RuntimeError(f"{op} failed with argument error {code} "
                           "(-1 bad argument, -2 unsupported shape, -3 workspace too small)")


_NHD = 0


def _check_layout(layout, op):
    if int(layout) != _NHD:
        raise RuntimeError(f"{op}: only the NHD layout is built (the reference controller always uses NHD, "
                           "quest/utils/controller.py:38)")


def apply_rope_in_place(q, k, past_kv_len, rope_scale, rope_theta):
    """reference quest/ops/csrc/page.cu:212-253"""
    _check_input(q, "q"), _check_input(k, "k")
    _check(q.dim() == 3, "q must be a 3D tensor"), _check(k.dim() == 3, "k must be a 3D tensor")
    _check(q.size(0) == k.size(0), "CHECK_EQ(q.size(0), k.size(0)) failed")
    _check(q.size(2) == k.size(2), "CHECK_EQ(q.size(2), k.size(2)) failed")
    _check_half(q, "apply_rope_in_place")
    _rc(load_library().quest_rope_inplace(q.data_ptr(), k.data_ptr(), q.size(0), int(past_kv_len), q.size(1),
                                          k.size(1), q.size(2), float(rope_scale), float(rope_theta), _stream()),
        "apply_rope_in_place")


def rms_norm_forward(input, weight, output, epsilon):
    """reference quest/ops/csrc/rms_norm.cu:160-212"""
    _check_input(input, "input"), _check_input(weight, "weight"), _check_input(output, "output")
    _check_half(input, "rms_norm_forward")
    cols = input.size(-1)
    _check(weight.numel() == cols, "CHECK_EQ(weight.size(0), cols) failed")
    _rc(load_library().quest_rms_norm(input.data_ptr(), weight.data_ptr(), output.data_ptr(), input.numel() // cols,
                                      cols, float(epsilon), _stream()), "rms_norm_forward")


def topk_filtering(estimated_value, estimated_indices, d_out, indices_out, buf, page_budget):
    """reference quest/ops/csrc/topk.cu:7-46.  ``buf`` (raft scratch) is accepted and unused."""
    _check_input(estimated_value, "estimated_value")
    _check(estimated_value.dim() == 2, "estimated_value must be a 2D tensor")
    in_stride = _check_rows(estimated_indices, "estimated_indices")
    H, n = estimated_value.shape
    _check(estimated_indices.size(1) == n, "CHECK_EQ(num_pages, estimated_indices.size(1)) failed")
    _check(estimated_indices.size(0) == H, "CHECK_EQ(num_heads, estimated_indices.size(0)) failed")
    if n < page_budget:
        raise RuntimeError(f"CHECK_GE(num_pages, page_budget) failed. {n} vs {page_budget}")
    _check(estimated_indices.dtype == torch.int32, "CHECK_EQ(estimated_indices.scalar_type(), torch::kInt32) failed")
    _check(d_out.size(1) == page_budget, "CHECK_EQ(page_budget, d_out.size(1)) failed")
    _check(indices_out.size(1) == page_budget, "CHECK_EQ(page_budget, indices_out.size(1)) failed")
    _check_input(d_out, "d_out"), _check_input(indices_out, "indices_out")
    _check(indices_out.dtype == torch.int32, "indices_out must be int32")
    _check_half(estimated_value, "Top-k filtering")
    _rc(load_library().quest_topk(estimated_value.data_ptr(), estimated_indices.data_ptr(), in_stride, d_out.data_ptr(),
                                  indices_out.data_ptr(), H, n, int(page_budget), _stream()), "topk_filtering")


def _pool_dims(data, name):
    if data.dim() != 5:
        raise RuntimeError(f"{name} must be a 5D tensor")
    _, two, P, H, D = data.shape
    if two != 2:
        raise RuntimeError(f"CHECK_EQ({name}.size(1), 2) failed")
    return P, H, D  # page_size, heads, head_dim (NHD)


EST_META_STABLE, EST_META_ORDERED = 1, 2  # include/quest_b200.h QUEST_EST_*


def estimate_attn_score(q, o, metadata_data, metadata_indices, metadata_indptr, metadata_last_page_len,
                        metadata_last_page_idx, layout, flags=0):
    """reference quest/ops/csrc/estimate.cu:6-84.  Scores the first o.size(1) KV pages (the caller leaves out
    the last one: quest/utils/__init__.py:193-194).  flags: promises about the metadata (QUEST_EST_* in
    include/quest_b200.h); 0 = none, always valid."""
    _check_layout(layout, "estimate_attn_score")
    _check_input(q, "q"), _check_input(metadata_data, "metadata_data"), _check_input(metadata_indices, "metadata_indices")
    _check_input(o, "o")
    _check(q.dim() == 3, "q must be a 3D tensor"), _check(q.size(0) == 1, "CHECK_EQ(q.size(0), 1) failed")
    _check(metadata_indices.dim() == 1, "metadata_indices must be a 1D tensor")
    _check(metadata_indices.dtype == torch.int32, "CHECK_EQ(metadata_indices.scalar_type(), torch::kInt32) failed")
    P, Hkv, D = _pool_dims(metadata_data, "metadata_data")
    Hq = q.size(1)
    _check(q.size(2) == D, "CHECK_EQ(metadata_data.size(4), head_dim) failed")
    _check(Hq % Hkv == 0, "num_qo_heads must be a multiple of num_kv_heads")
    _check(o.dim() == 2 and o.size(0) == Hq, "o must be [num_heads, num_pages - 1]")
    n_entries = (metadata_indices.size(0) - 1) * P + int(metadata_last_page_len) - 1
    if o.size(1) != n_entries:
        raise RuntimeError(f"o.size(1) must equal the number of scored pages {n_entries}")
    _check_half(q, "Estimate_attn_score")
    _rc(load_library().quest_estimate(q.data_ptr(), o.data_ptr(), o.size(1), metadata_data.data_ptr(),
                                      metadata_indices.data_ptr(), Hq, Hkv, D, P, int(flags), _stream()), "estimate_attn_score")


def _append_checks(k, v, kv_data, kv_indices, kv_indptr, metadata_data, metadata_indices, metadata_indptr, layout, op):
    _check_layout(layout, op)
    for t, n in ((k, "k"), (v, "v"), (kv_data, "kv_data"), (kv_indices, "kv_indices"),
                 (metadata_data, "metadata_data"), (metadata_indices, "metadata_indices")):
        _check_input(t, n)
    _check(kv_indices.dim() == 1, "kv_indices must be a 1D tensor")
    _check(metadata_indices.dim() == 1, "metadata_indices must be a 1D tensor")
    _check(k.dim() == 3, "k must be a 3D tensor"), _check(v.dim() == 3, "v must be a 3D tensor")
    for t, n in ((kv_indices, "kv_indices"), (metadata_indices, "metadata_indices"), (kv_indptr, "kv_indptr"),
                 (metadata_indptr, "metadata_indptr")):
        if not (t.dtype == torch.int32):
            raise RuntimeError(f"CHECK_EQ({n}.scalar_type(), torch::kInt32) failed")
    P, H, D = _pool_dims(kv_data, "kv_data")
    _check(_pool_dims(metadata_data, "metadata_data") == (P, H, D), "metadata_data must share page_size/heads/dim with kv_data")
    if not (k.size(1) == H):
        raise RuntimeError(f"CHECK_EQ(kv_data.size(3), num_heads) failed. {H} vs {k.size(1)}")
    if not (k.size(2) == D):
        raise RuntimeError(f"CHECK_EQ(kv_data.size(4), head_dim) failed. {D} vs {k.size(2)}")
    _check(v.shape == k.shape, "CHECK_EQ(seq_len, v.size(0)) failed")
    _check_half(k, op)
    return P, H, D


def append_kv_cache_decode(k, v, kv_data, kv_indices, kv_indptr, kv_last_page_len, kv_last_page_idx, metadata_data,
                           metadata_indices, metadata_indptr, metadata_last_page_len, metadata_last_page_idx, layout):
    """reference quest/ops/csrc/page.cu:6-99"""
    P, H, D = _append_checks(k, v, kv_data, kv_indices, kv_indptr, metadata_data, metadata_indices, metadata_indptr,
                             layout, "Append_kv_cache_decode")
    if not (k.size(0) == 1):
        raise RuntimeError(f"CHECK_EQ(k.size(0), 1) failed. {k.size(0)} vs 1")
    _rc(load_library().quest_append_decode(k.data_ptr(), v.data_ptr(), kv_data.data_ptr(), kv_indices.data_ptr(),
                                           kv_indices.size(0), int(kv_last_page_len), metadata_data.data_ptr(),
                                           metadata_indices.data_ptr(), metadata_indices.size(0),
                                           int(metadata_last_page_len), H, D, P, _stream()), "append_kv_cache_decode")


def append_kv_cache_prefill(k, v, kv_data, kv_indices, kv_indptr, kv_last_page_len, kv_last_page_idx, metadata_data,
                            metadata_indices, metadata_indptr, metadata_last_page_len, metadata_last_page_idx, layout):
    """reference quest/ops/csrc/page.cu:101-210"""
    P, H, D = _append_checks(k, v, kv_data, kv_indices, kv_indptr, metadata_data, metadata_indices, metadata_indptr,
                             layout, "Append_kv_cache_prefill")
    if not (k.size(0) >= 2):
        raise RuntimeError(f"CHECK_GE(k.size(0), 2) failed. {k.size(0)} vs 2")
    _rc(load_library().quest_append_prefill(k.data_ptr(), v.data_ptr(), k.size(0), kv_data.data_ptr(),
                                            kv_indices.data_ptr(), kv_indices.size(0), int(kv_last_page_len),
                                            metadata_data.data_ptr(), metadata_indices.data_ptr(),
                                            metadata_indices.size(0), H, D, P, _stream()), "append_kv_cache_prefill")


def prefill_with_paged_kv_cache(q, kv_data, kv_indices, kv_last_page_len, causal, layout, allow_fp16_qk_reduction,
                                rope_scale, rope_theta):
    """reference quest/ops/csrc/batch_prefill.cu:27-117 (allocates and returns o)."""
    _check_layout(layout, "prefill_with_paged_kv_cache")
    _check_input(q, "q"), _check_input(kv_data, "kv_data"), _check_input(kv_indices, "kv_indices")
    _check(q.dim() == 3, "q must be a 3D tensor"), _check(kv_indices.dim() == 1, "kv_indices must be a 1D tensor")
    _check(kv_indices.dtype == torch.int32, "CHECK_EQ(kv_indices.scalar_type(), torch::kInt32) failed")
    P, Hkv, D = _pool_dims(kv_data, "kv_data")
    _check(q.size(2) == D, "CHECK_EQ(q.size(2), kv_data.size(4)) failed")
    _check_half(q, "BatchPrefillWithPagedKVCache")
    o = torch.empty_like(q)
    _rc(load_library().quest_prefill(q.data_ptr(), o.data_ptr(), q.size(0), kv_data.data_ptr(), kv_indices.data_ptr(),
                                     kv_indices.size(0), int(kv_last_page_len), 1 if causal else 0, q.size(1), Hkv, D,
                                     P, _stream()), "prefill_with_paged_kv_cache")
    return o


def metadata_to_fp8(metadata_data, metadata_indices, first_kv_page, end_kv_page, meta8_data):
    """e4m3 copies (max rounded up, min rounded down) of the metadata entries of KV pages [first, end): the flag-gated fp8
    metadata variant (no counterpart in the reference).  meta8_data: uint8 pool of metadata_data's shape."""
    _check_input(metadata_data, "metadata_data"), _check_input(metadata_indices, "metadata_indices"), _check_input(meta8_data, "meta8_data")
    _check(meta8_data.dtype == torch.uint8 and meta8_data.shape == metadata_data.shape, "meta8_data must be a uint8 pool of metadata_data's shape")
    P, H, D = _pool_dims(metadata_data, "metadata_data")
    _rc(load_library().quest_metadata_to_fp8(metadata_data.data_ptr(), metadata_indices.data_ptr(), int(first_kv_page),
                                             int(end_kv_page), meta8_data.data_ptr(), H, D, P, _stream()), "metadata_to_fp8")


def estimate_attn_score_fp8(q, o, meta8_data, metadata_indices, flags=0):
    """estimate_attn_score over the e4m3 metadata pool; o [num_qo_heads, scored pages] fp16."""
    _check_input(q, "q"), _check_input(o, "o"), _check_input(meta8_data, "meta8_data"), _check_input(metadata_indices, "metadata_indices")
    _check(meta8_data.dtype == torch.uint8 and meta8_data.dim() == 5, "meta8_data must be a 5D uint8 pool")
    _check_half(q, "estimate_attn_score_fp8")
    P, H, D = meta8_data.size(2), meta8_data.size(3), meta8_data.size(4)
    _rc(load_library().quest_estimate_fp8(q.data_ptr(), o.data_ptr(), o.size(1), meta8_data.data_ptr(), metadata_indices.data_ptr(),
                                          q.size(1), H, D, P, int(flags), _stream()), "estimate_attn_score_fp8")


PF_WRITE_BACK_K = 1  # include/quest_b200.h QUEST_PF_WRITE_BACK_K


def _prefill_producer_checks(q, k, v, kv_data, kv_indices, metadata_data, metadata_indices, op):
    _check_input(q, "q")
    _check(q.dim() == 3, "q must be a 3D tensor")
    P, H, D = _append_checks(k, v, kv_data, kv_indices, kv_indices, metadata_data, metadata_indices, metadata_indices, _NHD, op)
    _check(q.size(0) == k.size(0), "CHECK_EQ(q.size(0), k.size(0)) failed")
    _check(q.size(2) == D, "CHECK_EQ(q.size(2), kv_data.size(4)) failed")
    _check_half(q, op)
    return P, H, D


def rope_append_kv_cache_prefill(q, k, v, kv_data, kv_indices, kv_last_page_len, metadata_data, metadata_indices, rope_scale,
                                 rope_theta, write_back_k=False):
    """apply_rope_in_place(q, k) + append_kv_cache_prefill(k, v) in one launch (reference quest/ops/csrc/page.cu:212-253 then
    :101-210): q is rotated in place, the rotated k goes straight into the pages and their min/max metadata; k itself is
    rewritten only with write_back_k.  The RoPE position of the first new token is the cache length before the append."""
    P, H, D = _prefill_producer_checks(q, k, v, kv_data, kv_indices, metadata_data, metadata_indices, "rope_append_kv_cache_prefill")
    _rc(load_library().quest_rope_append_prefill(q.data_ptr(), k.data_ptr(), v.data_ptr(), k.size(0), kv_data.data_ptr(),
                                                 kv_indices.data_ptr(), kv_indices.size(0), int(kv_last_page_len),
                                                 metadata_data.data_ptr(), metadata_indices.data_ptr(), metadata_indices.size(0),
                                                 q.size(1), H, D, P, float(rope_scale), float(rope_theta),
                                                 PF_WRITE_BACK_K if write_back_k else 0, _stream()), "rope_append_kv_cache_prefill")


def prefill_layer_step(q, k, v, kv_data, kv_indices, kv_last_page_len, metadata_data, metadata_indices, rope_scale, rope_theta,
                       write_back_k=False):
    """The prompt's attention layer in one native call (reference quest/models/QuestAttention.py:85-99: apply_rope_in_place,
    append_kv, prefill_forward): fused RoPE + append + metadata, then causal paged prefill, chained with programmatic
    dependent launch.  Returns o like prefill_with_paged_kv_cache."""
    P, H, D = _prefill_producer_checks(q, k, v, kv_data, kv_indices, metadata_data, metadata_indices, "prefill_layer_step")
    o = torch.empty_like(q)
    _rc(load_library().quest_prefill_layer_step(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), k.size(0),
                                                kv_data.data_ptr(), kv_indices.data_ptr(), kv_indices.size(0),
                                                int(kv_last_page_len), metadata_data.data_ptr(), metadata_indices.data_ptr(),
                                                metadata_indices.size(0), q.size(1), H, D, P, float(rope_scale),
                                                float(rope_theta), PF_WRITE_BACK_K if write_back_k else 0, _stream()),
        "prefill_layer_step")
    return o


_NUM_SMS = {}


def _num_sms(device):
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _NUM_SMS:
        _NUM_SMS[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return _NUM_SMS[idx]


class BatchDecodeWithPagedKVCachePyTorchWrapper:
    """reference quest/ops/csrc/approx_attn.cu:27-151 + kernels/include/decode/decode_handler.cuh:39-244.
    The split-KV plan is pure host arithmetic and the temporary buffer is a torch tensor owned by this object
    between begin_forward and end_forward (the reference's cudaMallocAsync'ed handler buffers)."""

    def __init__(self, layout):
        _check_layout(layout, "BatchDecodeWithPagedKVCachePyTorchWrapper")
        self._plan = None
        self._ws = None

    def begin_forward(self, indptr, num_qo_heads, num_kv_heads, head_dim, page_size, empty_data, num_pages=None):
        """``num_pages`` (host int) lets a caller that already knows indptr[1] skip the device read the reference
        does in BeginForward (decode_attn.cuh:866-873)."""
        _check(empty_data.dtype == torch.float16,
               f"BatchDecodeWithPagedKVCache failed to dispatch with dtype {empty_data.dtype}")
        if num_pages is None:
            _check(indptr.dim() == 1 and indptr.dtype == torch.int32, "indptr must be a 1D int32 tensor")
            num_pages = int(indptr[-1].item())
        self._plan = (int(num_pages), int(num_qo_heads), int(num_kv_heads), int(head_dim), int(page_size))
        self._chunks = {}

    def end_forward(self):
        self._plan = None  # the workspace tensor is kept for the next begin_forward (tickets are self-resetting)

    def _planned(self, n_sel, Hq, Hkv, D, P, device):
        key = (n_sel, Hq, Hkv, D, P)
        if key not in self._chunks:
            lib = load_library()
            ppc, nch = _c.c_uint32(0), _c.c_uint32(0)
            _rc(lib.quest_decode_plan(n_sel, Hq, Hkv, D, P, _num_sms(device), _c.byref(ppc), _c.byref(nch)),
                "BatchDecodeWithPagedKVCache plan")
            self._chunks[key] = (ppc.value, nch.value, lib.quest_decode_workspace_bytes(Hq, D, nch.value))
        ppc, nch, nbytes = self._chunks[key]
        if self._ws is None or self._ws.numel() < nbytes or self._ws.device != device:
            self._ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        return ppc, nch

    def forward(self, q, o, paged_kv_data, paged_kv_indices, paged_kv_indptr, paged_kv_last_page_len,
                paged_kv_last_page_idx, rope_scale, rope_theta):
        if self._plan is None:
            raise RuntimeError("BatchDecodeWithPagedKVCache: forward() called before begin_forward()")
        _check_input(q, "q"), _check_input(paged_kv_data, "paged_kv_data")
        _check_input(o, "o")
        _check(q.dim() == 3, "q must be a 3D tensor")
        stride = _check_rows(paged_kv_indices, "paged_kv_indices")
        P, Hkv, D = _pool_dims(paged_kv_data, "paged_kv_data")
        Hq = q.size(1)
        _check(paged_kv_indices.size(0) == Hq, "CHECK_EQ(paged_kv_indices.size(0), num_qo_heads) failed")
        _check(q.size(2) == D, "CHECK_EQ(paged_kv_data.size(4), head_dim) failed")
        _check(paged_kv_indices.dtype == torch.int32, "CHECK_EQ(paged_kv_indices.scalar_type(), torch::kInt32) failed")
        _check_half(q, "BatchDecodeWithPagedKVCache")
        n_sel = self._plan[0]
        _check(n_sel <= paged_kv_indices.size(1), "indptr[1] exceeds the width of paged_kv_indices")
        ppc, _ = self._planned(n_sel, Hq, Hkv, D, P, q.device)
        _rc(load_library().quest_decode_run(q.data_ptr(), o.data_ptr(), None, paged_kv_data.data_ptr(),
                                            paged_kv_indices.data_ptr(), stride, n_sel, int(paged_kv_last_page_idx),
                                            int(paged_kv_last_page_len), Hq, Hkv, D, P, ppc, self._ws.data_ptr(),
                                            self._ws.numel(), _stream()), "BatchDecodeWithPagedKVCache")

    def layer_step_args(self, q, k, v, o, kv_data, kv_indices, kv_last_page_len, kv_last_page_idx, metadata_data,
                        metadata_indices, metadata_last_page_len, scores, topk_indices):
        """Validates a layer step and returns the argument tuple of quest_decode_layer_step (without the stream)."""
        if self._plan is None:
            raise RuntimeError("BatchDecodeWithPagedKVCache: layer_step() called before begin_forward()")
        for t_, n_ in ((q, "q"), (k, "k"), (v, "v"), (o, "o"), (kv_data, "kv_data"), (kv_indices, "kv_indices"),
                       (metadata_data, "metadata_data"), (metadata_indices, "metadata_indices"), (scores, "scores"),
                       (topk_indices, "topk_indices")):
            _check_input(t_, n_)
        P, Hkv, D = _pool_dims(kv_data, "kv_data")
        _check(_pool_dims(metadata_data, "metadata_data") == (P, Hkv, D), "metadata_data must share page_size/heads/dim with kv_data")
        Hq = q.size(1)
        n_pages, n_sel = kv_indices.size(0), self._plan[0]
        _check(q.dim() == 3 and q.size(0) == 1 and q.size(2) == D, "q must be [1, num_qo_heads, head_dim]")
        _check(k.shape == (1, Hkv, D) and v.shape == k.shape, "k, v must be [1, num_kv_heads, head_dim]")
        _check(kv_indices.dtype == torch.int32 and metadata_indices.dtype == torch.int32 and topk_indices.dtype == torch.int32,
               "page tables and topk_indices must be int32")
        _check(tuple(scores.shape) == (Hq, n_pages - 1), "scores must be [num_qo_heads, num_kv_pages - 1]")
        _check(tuple(topk_indices.shape) == (Hq, n_sel), "topk_indices must be [num_qo_heads, planned num_pages]")
        _check_half(q, "decode_layer_step")
        ppc, _ = self._planned(n_sel, Hq, Hkv, D, P, q.device)
        return (q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), kv_data.data_ptr(), kv_indices.data_ptr(), n_pages,
                int(kv_last_page_idx), int(kv_last_page_len), metadata_data.data_ptr(), metadata_indices.data_ptr(),
                metadata_indices.size(0), int(metadata_last_page_len), scores.data_ptr(), topk_indices.data_ptr(), n_sel,
                Hq, Hkv, D, P, ppc, self._ws.data_ptr(), self._ws.numel())

    def layer_step(self, *tensors_and_ints):
        """append_kv_cache_decode + estimate_attn_score + topk_filtering + forward in one C call
        (quest_decode_layer_step: the append rides on the estimate kernel).  Same tensors as those four ops take
        (see layer_step_args); begin_forward must have planned num_pages = topk_indices.size(1)."""
        _rc(load_library().quest_decode_layer_step(*self.layer_step_args(*tensors_and_ints), _stream()), "decode_layer_step")
