#!/usr/bin/env python
"""End-to-end decode throughput (tok/s) with the reference's own protocol (scripts/bench_textgen.py:41-106): prefill a
context of random embeddings, then decode `--decode_len` tokens, each step fed a random embedding, and report the mean
decode latency -- for a LongChat-7B-v1.5-32K shaped decoder (32 layers, hidden 4096, MLP 11008, 32 heads x 128, fp16,
random-init weights: there is no network for checkpoints).

This is a measurement harness, not a model zoo: the projections / MLP are plain torch matmuls (cuBLAS), everything on
the attention path is this library through the reference's operator surface (quest_b200.utils: rms_norm_forward,
apply_rope_in_place, append_kv, prefill_forward, decode_layer_step) driven exactly as quest/models/llama.py:423-439 and
QuestAttention.py:95-160 drive it: layers 0-1 run with an unlimited page budget (full attention), the others with the
token budget; `--token_budget 0` = full cache on every layer (the reference's FlashInfer-full-cache comparison point).

    python tools/e2e_textgen.py --context_len 32768 --token_budget 2048 [--decode_len 256]
Prints one JSON line: tok/s, ms per token (synchronised per token, stricter than the reference's unsynchronised timer),
device ms of the attention-path ops alone (CUDA events around them, second pass) and their share."""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200.utils as qu  # noqa: E402

L, HID, INTER, H, D, VOCAB = 32, 4096, 11008, 32, 128, 32000
P, SKIP_LAYERS, ROPE_SCALE, EPS = 16, 2, 8.0, 1e-5   # LongChat-7B-v1.5-32K: linear RoPE scaling 8 (config.json)


class Decoder:
    def __init__(self, dev, layers):
        g = torch.Generator(device=dev).manual_seed(0)

        def w(*shape):
            return (torch.randn(*shape, generator=g, device=dev, dtype=torch.float32) * 0.02).half()
        self.layers = [dict(wqkv=w(HID, 3 * HID), wo=w(HID, HID), wgu=w(HID, 2 * INTER), wd=w(INTER, HID),
                            n1=torch.ones(HID, dtype=torch.float16, device=dev), n2=torch.ones(HID, dtype=torch.float16, device=dev))
                       for _ in range(layers)]
        self.norm = torch.ones(HID, dtype=torch.float16, device=dev)
        self.lm_head = w(HID, VOCAB)
        self.attn_events = None

    def forward(self, x, ic, budget_pages, max_pages):
        """x: [n, HID] fp16 embeddings of the n new tokens.  Returns logits of the last token."""
        n = x.shape[0]
        ic.prepare_metadata(n)
        ic.set_page_budget(max_pages)            # layers 0..SKIP_LAYERS-1: unlimited budget (llama.py:430-431)
        ic.begin_forward(n)
        ev = self.attn_events
        for i, ly in enumerate(self.layers):
            if i == SKIP_LAYERS:
                ic.end_forward()
                ic.set_page_budget(budget_pages)
                ic.begin_forward(n, updateTensor=False)
            h = qu.rms_norm_forward(x.unsqueeze(0), ly["n1"], EPS).squeeze(0)
            qkv = (h @ ly["wqkv"]).view(n, 3, H, D)
            q, k, v = qkv[:, 0].contiguous(), qkv[:, 1].contiguous(), qkv[:, 2].contiguous()
            if ev is not None:
                ev[i][0].record()
            qu.apply_rope_in_place(q, k, ic.kv_cache.seqlen - n, rope_scale=ROPE_SCALE)
            if n > 1:
                qu.append_kv(k, v, ic, i)
                o = qu.prefill_forward(q, ic, i)
            else:
                o = qu.decode_layer_step(q, k, v, ic, i)
            if ev is not None:
                ev[i][1].record()
            x = x + o.view(n, HID) @ ly["wo"]
            h = qu.rms_norm_forward(x.unsqueeze(0), ly["n2"], EPS).squeeze(0)
            gu = h @ ly["wgu"]
            x = x + (torch.nn.functional.silu(gu[:, :INTER]) * gu[:, INTER:]) @ ly["wd"]
        ic.end_forward()
        last = qu.rms_norm_forward(x[-1:].unsqueeze(0), self.norm, EPS).squeeze(0)
        return last @ self.lm_head

    def decode_body(self, x, step_full, step_quest):
        """One decode token with the descriptor-driven step (BatchDecodeStep, batch 1): no host-side value changes from
        token to token, so the whole body is captured once and replayed.  x: [1, HID] static input buffer."""
        step_full.upload()
        if step_quest is not step_full:
            step_quest.upload()
        for i, ly in enumerate(self.layers):
            st = step_full if i < SKIP_LAYERS else step_quest
            h = qu.rms_norm_forward(x.unsqueeze(0), ly["n1"], EPS).squeeze(0)
            qkv = (h @ ly["wqkv"]).view(3, H, D)
            q, k, v = qkv[0:1], qkv[1:2], qkv[2:3]          # [1, H, D] each, contiguous
            st.rope(q, k, rope_scale=ROPE_SCALE)
            o = st.layer_step(q, k, v, i)
            x = x + o.view(1, HID) @ ly["wo"]
            h = qu.rms_norm_forward(x.unsqueeze(0), ly["n2"], EPS).squeeze(0)
            gu = h @ ly["wgu"]
            x = x + (torch.nn.functional.silu(gu[:, :INTER]) * gu[:, INTER:]) @ ly["wd"]
        last = qu.rms_norm_forward(x.unsqueeze(0), self.norm, EPS).squeeze(0)
        return last @ self.lm_head


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--context_len", type=int, default=32768)
    ap.add_argument("--decode_len", type=int, default=256)
    ap.add_argument("--token_budget", type=int, default=2048, help="0 = full cache on every layer")
    ap.add_argument("--layers", type=int, default=L)
    ap.add_argument("--prefill_chunk", type=int, default=8192)
    ap.add_argument("--graph", action="store_true", help="decode through ONE captured CUDA graph replayed per token "
                    "(descriptor-driven step, quest_decode_layer_step_batch) instead of eager operator calls")
    a = ap.parse_args()
    print(json.dumps(run(a)))


@torch.inference_mode()
def run(a, dev=None):
    """a: namespace with context_len, decode_len, token_budget, layers, prefill_chunk, graph.  Returns the result dict."""
    if dev is None:
        dev = torch.device("cuda", 0)
        torch.cuda.set_device(dev)
    max_seq = a.context_len + a.decode_len + 512
    max_pages = (max_seq + P - 1) // P
    budget_pages = max_pages if a.token_budget == 0 else a.token_budget // P
    model = Decoder(dev, a.layers)
    ic = qu.InferenceController(a.layers, H, D, P, budget_pages, max_seq, torch.float16, dev)

    # prefill (chunked: each chunk attends causally over everything appended so far)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    done = 0
    while done < a.context_len:
        n = min(a.prefill_chunk, a.context_len - done)
        model.forward(torch.randn(n, HID, dtype=torch.float16, device=dev), ic, budget_pages, max_pages)
        done += n
    torch.cuda.synchronize()
    prefill_s = time.perf_counter() - t0

    if a.graph:
        return graph_decode(a, model, ic, budget_pages, max_pages, prefill_s, dev)
    lat = []
    for step in range(a.decode_len):
        x = torch.randn(1, HID, dtype=torch.float16, device=dev)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        logits = model.forward(x, ic, budget_pages, max_pages)
        torch.cuda.synchronize()
        lat.append(time.perf_counter() - t0)
    assert torch.isfinite(logits.float()).all()
    warm = lat[8:] if len(lat) > 16 else lat
    ms = sum(warm) / len(warm) * 1e3

    # second pass over a few tokens: device time of the attention-path ops alone
    model.attn_events = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.layers)]
    attn_ms = []
    for step in range(8):
        model.forward(torch.randn(1, HID, dtype=torch.float16, device=dev), ic, budget_pages, max_pages)
        torch.cuda.synchronize()
        attn_ms.append(sum(e0.elapsed_time(e1) for e0, e1 in model.attn_events))
    attn = sorted(attn_ms)[len(attn_ms) // 2]
    return ({"metric": "e2e_decode_tok_per_s", "value": 1e3 / ms, "ms_per_token": ms, "prefill_s": prefill_s,
                      "context_len": a.context_len, "decode_len": a.decode_len, "token_budget": a.token_budget or "full",
                      "page_size": P, "layers": a.layers, "attention_path_ms_per_token": attn, "attention_share": attn / ms,
                      "model": "LongChat-7B-v1.5-32K shape, random init fp16, batch 1; torch matmuls + quest_b200 operators, eager",
                      "protocol": "reference scripts/bench_textgen.py (random embeddings in, mean decode latency)"})


def graph_decode(a, model, ic, budget_pages, max_pages, prefill_s, dev):
    from quest_b200.utils.batch import BatchDecodeStep
    ic.set_page_budget(max_pages)
    step_full = BatchDecodeStep([ic])            # layers 0-1: every page selected (k >= pages -> the rest of the row is -1)
    ic.set_page_budget(budget_pages)
    step_quest = step_full if budget_pages >= max_pages else BatchDecodeStep([ic])
    x = torch.zeros(1, HID, dtype=torch.float16, device=dev)

    def host_token():
        ic.prepare_metadata(1)
        step_full.update_host()
        if step_quest is not step_full:
            step_quest.update_host()
        x.copy_(torch.randn(1, HID, dtype=torch.float16, device=dev))

    host_token()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        model.decode_body(x, step_full, step_quest)     # warm-up (also token 0)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        logits = model.decode_body(x, step_full, step_quest)
    lat, dev_ms = [], []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for t in range(a.decode_len):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        host_token()
        assert step_quest.fits() and step_full.fits()
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        lat.append(time.perf_counter() - t0)
        dev_ms.append(e0.elapsed_time(e1))
    assert torch.isfinite(logits.float()).all()
    warm = lat[8:] if len(lat) > 16 else lat
    ms = sum(warm) / len(warm) * 1e3
    return ({"metric": "e2e_decode_tok_per_s", "value": 1e3 / ms, "ms_per_token": ms, "prefill_s": prefill_s,
                      "graph_device_ms_per_token": sorted(dev_ms)[len(dev_ms) // 2],
                      "context_len": a.context_len, "decode_len": a.decode_len, "token_budget": a.token_budget or "full",
                      "page_size": P, "layers": a.layers, "mode": "one CUDA graph replayed per token (records refreshed by the host)",
                      "model": "LongChat-7B-v1.5-32K shape, random init fp16, batch 1; torch matmuls + quest_b200 operators",
                      "protocol": "reference scripts/bench_textgen.py (random embeddings in, mean decode latency, host work included)"})


if __name__ == "__main__":
    main()
