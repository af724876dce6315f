"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

CPU restatement (torch, fp32 arithmetic with explicit roundings to the storage type) of the decode-layer glue kernels
of the reference (paths relative to /root/reference):
  rms_norm             kernels/layernorm_kernels.cu:15-43
  fused_add_rms_norm   kernels/layernorm_kernels.cu:50-137
  silu_and_mul         kernels/activation_kernels.cu:12-40,82-86
  rotary_embedding     kernels/pos_encoding_kernels.cu:10-96
Every function rounds where the CUDA kernel rounds (not where the Python `forward_native` does), so silu_and_mul and
rotary_embedding are expected bit-exact against the CUDA kernels and the norms up to the order of the fp32 sum.

Parity status: pinned - tests/test_oracle_pinned.py checks these against fixtures made by the reference's own
`forward_native` methods (tests/golden/make_golden.py make_glue -> tests/golden/glue_golden.npz; rotary bit-exact,
the others within the tolerance the reference's kernel tests use), and tests/test_gpu_layer_glue.py checks them
against the reference CUDA kernels on the GPU when oracle/_ref/refcuda.so is built.
"""
from __future__ import annotations

import torch


def _r(x: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """Round an fp32 tensor to the storage type and come back to fp32."""
    return x.to(dtype).float()


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    dt = x.dtype
    f = x.float()
    rs = torch.rsqrt(f.pow(2).sum(-1, keepdim=True) / x.shape[-1] + eps)
    return (_r(f * rs, dt) * weight.float()).to(dt)


def fused_add_rms_norm(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, eps: float):
    """Returns (normed, new_residual); the kernel writes them over `input` and `residual`."""
    dt = x.dtype
    z = (x.float() + residual.float()).to(dt)  # scalar_t add
    return rms_norm(z, weight, eps), z


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    dt = x.dtype
    d = x.shape[-1] // 2
    g, u = x[..., :d].float(), x[..., d:].float()
    act = _r(g / (1.0 + torch.exp(-g)), dt)
    return (act * u).to(dt)


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool):
    """Returns rotated copies of (query, key); tensors are [tokens, heads * head_size]."""
    dt = query.dtype
    rot = cos_sin_cache.shape[1]
    cs = cos_sin_cache[positions.flatten().long()].float()
    cos, sin = cs[:, None, : rot // 2], cs[:, None, rot // 2:]

    def one(t: torch.Tensor) -> torch.Tensor:
        shape = t.shape
        t = t.reshape(positions.numel(), -1, head_size).clone()
        r = t[..., :rot].float()
        x, y = (r[..., : rot // 2], r[..., rot // 2:]) if is_neox else (r[..., 0::2], r[..., 1::2])
        ox = _r(_r(x * cos, dt) - _r(y * sin, dt), dt)
        oy = _r(_r(y * cos, dt) + _r(x * sin, dt), dt)
        o = torch.cat((ox, oy), -1) if is_neox else torch.stack((ox, oy), -1).flatten(-2)
        t[..., :rot] = o.to(dt)
        return t.reshape(shape)

    return one(query), one(key)
