"""Generates tests/golden/*.npz by running the REFERENCE's own Python code (unmodified, imported from
/root/reference) on seeded inputs.  Run in the authoring container only:

    python tests/golden/make_golden.py

Inputs are drawn from numpy's PCG64 (``golden_inputs``) so the tests can rebuild them bit for bit from the
stored seed; only the reference's outputs are stored.  Harness-side shims (the reference files are never
edited): (1) stub ``quest`` / ``quest.utils`` modules so the reference test modules, which import the CUDA
extension at import time, can be loaded for their pure-torch ``_ref_*`` functions; (2)
``evaluation.quest_attention.apply_rotary_pos_emb`` is rebound to identity because the reference targets
transformers 4.45 (pyproject.toml:12) and inputs are fed post-RoPE.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def golden_inputs(seed, kv_len, H, D, Hkv=None):
    """q [1,H,D], k/v [kv_len,Hkv,D] fp16 ~ N(0,1) (same distribution as quest/tests/*.py)."""
    rng = np.random.default_rng(seed)
    Hkv = H if Hkv is None else Hkv
    q = rng.standard_normal((1, H, D), dtype=np.float32).astype(np.float16)
    k = rng.standard_normal((kv_len, Hkv, D), dtype=np.float32).astype(np.float16)
    v = rng.standard_normal((kv_len, Hkv, D), dtype=np.float32).astype(np.float16)
    return q, k, v


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    torch.manual_seed(0)
    for n in ("quest", "quest.utils"):
        sys.modules.setdefault(n, types.ModuleType(n))
    t_est = _load(f"{REF}/quest/tests/test_estimate.py", "ref_test_estimate")
    t_apx = _load(f"{REF}/quest/tests/test_approx_attention.py", "ref_test_approx")
    t_dec = _load(f"{REF}/quest/tests/test_decode_attention.py", "ref_test_decode")
    t_pre = _load(f"{REF}/quest/tests/test_prefill_attention.py", "ref_test_prefill")

    H, D, P = 4, 128, 16
    cases = {}
    # kernel-path semantics: estimate scores, forced-last-page top-k attention, dense decode, causal prefill
    for i, (kv_len, budget) in enumerate([(61, 3), (113, 4), (482, 7), (577, 15), (1011, 31)]):
        seed = 1000 + i
        q, k, v = golden_inputs(seed, kv_len, H, D)
        tq, tk, tv = (torch.from_numpy(x.astype(np.float32)).half() for x in (q, k, v))
        est = t_est._ref_cpu_estimate(tq, tk, tv, P)
        o_apx, topk = t_apx._ref_self_approx_attention(tq.clone(), tk, tv, P, budget)
        o_dense = t_dec._ref_self_attention(tq.clone(), tk, tv)
        cases[f"est_{i}"] = est.numpy()
        cases[f"apx_o_{i}"] = o_apx.float().numpy()
        cases[f"apx_topk_{i}"] = topk.reshape(H, -1).numpy() if topk is not None else np.zeros((H, 0), np.int32)
        cases[f"dense_o_{i}"] = o_dense.float().numpy()
        cases[f"meta_{i}"] = np.array([seed, kv_len, budget, H, D, P], dtype=np.int64)
    for i, (qo_len, kv_len) in enumerate([(13, 33), (47, 111), (64, 64)]):
        seed = 2000 + i
        rng = np.random.default_rng(seed)
        q = rng.standard_normal((qo_len, H, D), dtype=np.float32).astype(np.float16)
        k = rng.standard_normal((kv_len, H, D), dtype=np.float32).astype(np.float16)
        v = rng.standard_normal((kv_len, H, D), dtype=np.float32).astype(np.float16)
        ref_fn = getattr(t_pre, "_ref_self_attention")
        o = ref_fn(*(torch.from_numpy(x.astype(np.float32)).half() for x in (q, k, v)))
        cases[f"prefill_o_{i}"] = o.float().numpy()
        cases[f"prefill_meta_{i}"] = np.array([seed, qo_len, kv_len, H, D], dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "kernel_path.npz"), **cases)

    # torch path: evaluation/quest_attention.py forward with a stub attention object
    sys.path.insert(0, REF)
    import evaluation.quest_attention as qa
    qa.apply_rotary_pos_emb = lambda q, k, cos, sin, position_ids=None: (q, k)
    out = {}
    for i, (Hq, Hkv, Dh, T, budget, chunk) in enumerate([(4, 4, 64, 300, 64, 16), (8, 2, 64, 519, 128, 16),
                                                        (4, 4, 128, 97, 48, 8)]):
        seed = 3000 + i
        rng = np.random.default_rng(seed)
        q = rng.standard_normal((1, Hq, 1, Dh), dtype=np.float32)
        k = rng.standard_normal((1, Hkv, T, Dh), dtype=np.float32)
        v = rng.standard_normal((1, Hkv, T, Dh), dtype=np.float32)
        tq, tk, tv = torch.from_numpy(q), torch.from_numpy(k), torch.from_numpy(v)

        class Stub:
            pass

        s = Stub()
        s.layer_id, s.layer_idx = 2, 2
        s.num_heads, s.num_key_value_heads, s.num_key_value_groups = Hq, Hkv, Hq // Hkv
        s.head_dim, s.hidden_size = Dh, Hq * Dh
        s.token_budget, s.chunk_size = budget, chunk
        s.q_proj = lambda h: tq.transpose(1, 2).reshape(1, 1, Hq * Dh)
        s.k_proj = lambda h: tk[:, :, -1:, :].transpose(1, 2).reshape(1, 1, Hkv * Dh)
        s.v_proj = lambda h: tv[:, :, -1:, :].transpose(1, 2).reshape(1, 1, Hkv * Dh)
        s.o_proj = lambda x: x
        s.rotary_emb = lambda x, pos: (None, None)
        hidden = torch.zeros(1, 1, Hq * Dh)
        past = (tk[:, :, :-1, :], tv[:, :, :-1, :])
        o, _, _ = qa.forward(s, hidden, None, torch.tensor([[T - 1]]), past, False, False)
        out[f"o_{i}"] = o.reshape(1, 1, Hq, Dh).transpose(1, 2).numpy()
        out[f"meta_{i}"] = np.array([seed, Hq, Hkv, Dh, T, budget, chunk], dtype=np.int64)
        # and the mask helper alone
        w = torch.from_numpy(rng.standard_normal((1, Hq, 1, T), dtype=np.float32))
        out[f"mask_in_{i}"] = w.numpy()
        out[f"mask_{i}"] = qa.local_heavy_hitter_mask(w, budget, chunk).numpy()
    np.savez_compressed(os.path.join(OUT, "torch_path.npz"), **out)
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
