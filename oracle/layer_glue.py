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


# ---- the other activations of kernels/activation_kernels.cu:42-62,111-115,183-205 ----
def _act(name: str, x: torch.Tensor, threshold: float = 0.0) -> torch.Tensor:
    """act(x) rounded to x.dtype the way the CUDA function of that name rounds."""
    dt, f = x.dtype, x.float()
    if name == "silu":
        return (f / (1.0 + torch.exp(-f))).to(dt)
    if name == "gelu":
        return (f * 0.5 * (1.0 + torch.erf(f * 0.7071067811865476))).to(dt)
    if name == "gelu_tanh":
        beta = torch.tensor(1.4142135623730951 * 1.1283791670955126 * 0.5, dtype=torch.float32)
        inner = beta * (f + torch.tensor(0.044715, dtype=torch.float32) * (f * f * f))
        return (0.5 * f * (1.0 + torch.tanh(inner))).to(dt)
    if name == "fatrelu":
        return torch.where(f > threshold, f, torch.zeros_like(f)).to(dt)
    if name == "gelu_new":
        x3 = _r(_r(f * f, dt) * f, dt)
        s = _r(f + _r(0.044715 * x3, dt), dt)
        t = _r(torch.tanh(_r(0.79788456 * s, dt)), dt)
        return (_r(_r(torch.tensor(0.5), dt) * f, dt) * _r(1.0 + t, dt)).to(dt)
    if name == "gelu_fast":
        arg = _r(_r(f * 0.79788456, dt) * _r(1.0 + _r(_r(0.044715 * f, dt) * f, dt), dt), dt)
        t = _r(torch.tanh(arg), dt)
        return (_r(0.5 * f, dt) * _r(1.0 + t, dt)).to(dt)
    if name == "gelu_quick":
        return (f / (1.0 + torch.exp(-1.702 * f))).to(dt)
    raise ValueError(name)


def act_and_mul(name: str, x: torch.Tensor, threshold: float = 0.0) -> torch.Tensor:
    """silu_and_mul / mul_and_silu / gelu_and_mul / gelu_tanh_and_mul / fatrelu_and_mul."""
    dt = x.dtype
    d = x.shape[-1] // 2
    a, b = x[..., :d], x[..., d:]
    if name == "mul_and_silu":
        return (a.float() * _act("silu", b).float()).to(dt)
    fn = {"silu_and_mul": "silu", "gelu_and_mul": "gelu", "gelu_tanh_and_mul": "gelu_tanh", "fatrelu_and_mul": "fatrelu"}[name]
    return (_act(fn, a, threshold).float() * b.float()).to(dt)


def activation(name: str, x: torch.Tensor) -> torch.Tensor:
    """gelu_new / gelu_fast / gelu_quick."""
    return _act(name, x)


def advance_step_flashattn(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids, input_positions, seq_lens,
                           slot_mapping, block_tables):
    """kernels/prepare_inputs/advance_step.cu:12-66, in place on CPU tensors (plain loop: integer work)."""
    for i in range(num_queries, num_seqs):
        input_tokens[i], input_positions[i], slot_mapping[i] = 0, 0, -1
    for q in range(num_queries):
        input_tokens[q] = sampled_token_ids[q, 0]
        pos = int(seq_lens[q])
        seq_lens[q] = pos + 1
        input_positions[q] = pos
        slot_mapping[q] = int(block_tables[q, pos // block_size]) * block_size + pos % block_size


def advance_step_flashinfer(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids, input_positions, seq_lens,
                            slot_mapping, block_tables, paged_kv_indices, paged_kv_indptr, paged_kv_last_page_len,
                            block_table_bound):
    """kernels/prepare_inputs/advance_step.cu:91-167 (the three kernels in launch order), in place on CPU tensors."""
    flat = sampled_token_ids.reshape(-1)
    for i in range(num_queries, num_seqs):
        input_tokens[i], input_positions[i], slot_mapping[i] = 0, 0, -1
    for q in range(num_queries):
        input_tokens[q] = flat[q]
        pos = int(seq_lens[q])
        seq_lens[q] = pos + 1
        input_positions[q] = pos
        paged_kv_last_page_len[q] = pos % block_size + 1
        slot_mapping[q] = int(block_tables[q, pos // block_size]) * block_size + pos % block_size
        block_table_bound[q] = -(-(pos + 1) // block_size)
    paged_kv_indptr[0] = 0
    total = 0
    for q in range(num_queries):
        total += int(block_table_bound[q])
        paged_kv_indptr[q + 1] = total
    for i in range(num_queries + 1, num_seqs + 1):
        paged_kv_indptr[i] = total
    for q in range(num_queries):
        start, n = int(paged_kv_indptr[q]), int(block_table_bound[q])
        paged_kv_indices[start:start + n] = block_tables[q, :n]
