"""Same-box timing of the decode-layer glue kernels against the UNMODIFIED reference CUDA kernels
(oracle/_ref/refcuda.so). Checker/measurement tool only. CUDA graph of back-to-back calls over rotating buffers,
CUDA events on the capturing stream (tests/tools/bench_vs_ref.py `timed`).

    python tests/tools/bench_glue.py > profiles/rNN_glue_vs_reference_cuda.jsonl
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from bench_vs_ref import A, impls, timed

DEV = "cuda"


def norm(tokens, hidden, iters=50):
    nbuf = max(2, int(300e6 // (tokens * hidden * 4)) + 1) if tokens * hidden * 4 > 1e6 else 64
    xs = [torch.randn(tokens, hidden, device=DEV).to(torch.bfloat16) for _ in range(nbuf)]
    rs = [torch.randn(tokens, hidden, device=DEV).to(torch.bfloat16) for _ in range(nbuf)]
    w = torch.ones(hidden, dtype=torch.bfloat16, device=DEV)
    for name, o in impls():
        us = timed(lambda i, o=o: o.fused_add_rms_norm(xs[i % nbuf], rs[i % nbuf], w, 1e-6), iters)
        print(json.dumps({"op": "fused_add_rms_norm", "tokens": tokens, "hidden": hidden, "impl": name, "us": round(us, 2),
                          "GBps": round(tokens * hidden * 2 * 4 / us / 1e3, 1)}), flush=True)


def silu(tokens, d, iters=50):
    nbuf = max(2, int(300e6 // (tokens * d * 6)) + 1) if tokens * d * 6 > 1e6 else 64
    xs = [torch.randn(tokens, 2 * d, device=DEV).to(torch.bfloat16) for _ in range(nbuf)]
    out = torch.empty(tokens, d, dtype=torch.bfloat16, device=DEV)
    for name, o in impls():
        us = timed(lambda i, o=o: o.silu_and_mul(out, xs[i % nbuf]), iters)
        print(json.dumps({"op": "silu_and_mul", "tokens": tokens, "d": d, "impl": name, "us": round(us, 2),
                          "GBps": round(tokens * d * 2 * 3 / us / 1e3, 1)}), flush=True)


def rope(tokens, nh, nkv, d, iters=50):
    nbuf = 16
    qs = [torch.randn(tokens, nh * d, device=DEV).to(torch.bfloat16) for _ in range(nbuf)]
    ks = [torch.randn(tokens, nkv * d, device=DEV).to(torch.bfloat16) for _ in range(nbuf)]
    cache = torch.randn(8192, d, device=DEV).to(torch.bfloat16)
    pos = torch.randint(0, 8192, (tokens,), device=DEV)
    for name, o in impls():
        us = timed(lambda i, o=o: o.rotary_embedding(pos, qs[i % nbuf], ks[i % nbuf], d, cache, True), iters)
        print(json.dumps({"op": "rotary_embedding", "tokens": tokens, "heads": nh, "kv_heads": nkv, "head_size": d, "impl": name,
                          "us": round(us, 2), "GBps": round(tokens * (nh + nkv) * d * 2 * 2 / us / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    A.load()
    for t in (64, 8192):
        norm(t, 4096)
        silu(t, 14336)
        rope(t, 32, 8, 128)
