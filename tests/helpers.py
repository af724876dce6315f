"""Shared builders for the GPU parity tests (inputs are made on the CPU from the integer-hash
generator so that the oracle and the CUDA path see identical bytes)."""
import numpy as np
import torch

import synth
from oracle import oracle

RTOL = {torch.float32: 1.3e-6, torch.float16: 1e-3, torch.bfloat16: 1.6e-2}


def make_attention_case(S, H, KVH, D, bs, seq_lens, dtype, kv, seed, k_scale=1.0, v_scale=1.0, extra_blocks=5):
    """Returns CPU tensors: q, key_cache, value_cache (scalar or uint8), block_tables, seq_lens."""
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    max_len = int(sl.max())
    mb = max((max_len + bs - 1) // bs, 1)
    nb = S * mb + extra_blocks
    sc = float(D) ** -0.5
    cache_esz = 1 if kv != "auto" else torch.tensor([], dtype=dtype).element_size()
    x = 16 // cache_esz
    q = synth.tensor((S, H, D), seed, dtype, sc)
    kc = synth.tensor((nb, KVH, D // x, bs, x), seed + 1, dtype, sc)
    vc = synth.tensor((nb, KVH, D, bs), seed + 2, dtype, sc)
    if kv != "auto":
        kc = oracle.fp8_encode(kc, k_scale, kv)
        vc = oracle.fp8_encode(vc, v_scale, kv)
    bt = torch.from_numpy(synth.permutation(nb, seed + 3)[: S * mb].reshape(S, mb).copy()).to(torch.int32)
    return q, kc, vc, bt, sl


def check_close(out, ref64, dtype, atol, what=""):
    out64 = out.detach().cpu().double()
    err = (out64 - ref64).abs()
    bound = atol + RTOL[dtype] * ref64.abs()
    worst = float((err - bound).max())
    assert torch.isfinite(out64).all(), f"{what}: non-finite output"
    assert worst <= 0, f"{what}: max abs err {float(err.max()):.3e} exceeds atol {atol} + rtol {RTOL[dtype]}"
    return float(err.max())
