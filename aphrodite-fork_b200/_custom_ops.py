"""Host-side mirror of the reference's Python op wrappers for the decode hot path.

Same names, argument order, defaults and error behaviour as
/root/reference/aphrodite/_custom_ops.py:65-123 (paged attention), :909-929 (gptq_marlin_gemm),
:1444-1456 (reshape_and_cache), :1504-1508 (convert_fp8); they resolve `torch.ops._C.*` lazily at
call time, so inside a live Aphrodite the reference's own wrappers reach the same kernels.
"""
from __future__ import annotations

from typing import Optional, Union

import torch

from ._loader import load


def _ops():
    load()
    return torch.ops._C, torch.ops._C_cache_ops


def paged_attention_v1(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    _ops()[0].paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                                 block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
                                 blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                                 blocksparse_head_sliding_step)


def paged_attention_v2(
    out: torch.Tensor,
    exp_sum: torch.Tensor,
    max_logits: torch.Tensor,
    tmp_out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    _ops()[0].paged_attention_v2(out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads,
                                 scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                                 kv_cache_dtype, k_scale, v_scale, tp_rank, blocksparse_local_blocks,
                                 blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step)


def reshape_and_cache(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    _ops()[1].reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale)


def reshape_and_cache_flash(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                            slot_mapping: torch.Tensor, kv_cache_dtype: str, k_scale: torch.Tensor,
                            v_scale: torch.Tensor) -> None:
    """aphrodite/_custom_ops.py:1459-1473 (NHD caches [num_blocks, block_size, num_heads, head_size])."""
    _ops()[1].reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale)


def copy_blocks(key_caches, value_caches, block_mapping: torch.Tensor) -> None:
    """aphrodite/_custom_ops.py:1489-1492."""
    _ops()[1].copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """aphrodite/_custom_ops.py:1499-1501 (block_mapping is a CPU tensor)."""
    _ops()[1].swap_blocks(src, dst, block_mapping)


def convert_fp8(output: torch.Tensor, input: torch.Tensor, scale: float = 1.0, kv_dtype: str = "fp8") -> None:
    _ops()[1].convert_fp8(output, input, scale, kv_dtype)


def gptq_marlin_gemm(a: torch.Tensor,
                     b_q_weight: torch.Tensor,
                     b_scales: torch.Tensor,
                     b_zeros: torch.Tensor,
                     g_idx: torch.Tensor,
                     perm: torch.Tensor,
                     workspace: torch.Tensor,
                     b_q_type: Union[int, "object"],
                     size_m: int,
                     size_n: int,
                     size_k: int,
                     is_k_full: bool,
                     has_zp: bool = False,
                     use_atomic_add: bool = False,
                     use_fp32_reduce: bool = False,
                     is_zp_float: bool = False) -> torch.Tensor:
    # the reference passes a ScalarType and forwards `.id`; plain ints are accepted too
    type_id = b_q_type.id if hasattr(b_q_type, "id") else int(b_q_type)
    return _ops()[0].gptq_marlin_gemm(a, b_q_weight, b_scales, b_zeros, g_idx, perm, workspace, type_id, size_m,
                                      size_n, size_k, is_k_full, has_zp, use_atomic_add, use_fp32_reduce,
                                      is_zp_float)


# aphrodite/_custom_ops.py:868-878 (checkpoint format -> Marlin wire format, run once per weight at load)
def gptq_marlin_repack(b_q_weight: torch.Tensor, perm: torch.Tensor, size_k: int, size_n: int,
                       num_bits: int) -> torch.Tensor:
    return _ops()[0].gptq_marlin_repack(b_q_weight, perm, size_k, size_n, num_bits)


def awq_marlin_repack(b_q_weight: torch.Tensor, size_k: int, size_n: int, num_bits: int) -> torch.Tensor:
    return _ops()[0].awq_marlin_repack(b_q_weight, size_k, size_n, num_bits)


# aphrodite/_custom_ops.py:881-906: one repack per expert, same loop as the reference
def gptq_marlin_moe_repack(b_q_weight: torch.Tensor, perm: torch.Tensor, size_k: int, size_n: int,
                           num_bits: int) -> torch.Tensor:
    num_experts = b_q_weight.shape[0]
    assert size_k % 16 == 0
    output = torch.empty((num_experts, size_k // 16, size_n * (num_bits // 2)), device=b_q_weight.device,
                         dtype=b_q_weight.dtype)
    for e in range(num_experts):
        output[e] = gptq_marlin_repack(b_q_weight[e], perm[e], size_k, size_n, num_bits)
    return output


def awq_marlin_moe_repack(b_q_weight: torch.Tensor, perm: torch.Tensor, size_k: int, size_n: int,
                          num_bits: int) -> torch.Tensor:
    num_experts = b_q_weight.shape[0]
    assert size_k % 16 == 0
    output = torch.empty((num_experts, size_k // 16, size_n * (num_bits // 2)), device=b_q_weight.device,
                         dtype=b_q_weight.dtype)
    for e in range(num_experts):
        output[e] = awq_marlin_repack(b_q_weight[e], size_k, size_n, num_bits)
    return output


# aphrodite/_custom_ops.py:36-37,178-219 (decode-layer glue)
def silu_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].silu_and_mul(out, x)


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    # the reference copies query / key to contiguous buffers and back around its kernel (_custom_ops.py:186-192);
    # this kernel takes the token stride, so slices of the qkv projection are rotated in place
    _ops()[0].rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    _ops()[0].rms_norm(out, input, weight, epsilon)


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    _ops()[0].fused_add_rms_norm(input, residual, weight, epsilon)


# aphrodite/_custom_ops.py:36-72: the other activations of kernels/activation_kernels.cu
def mul_and_silu(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].mul_and_silu(out, x)


def gelu_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].gelu_and_mul(out, x)


def gelu_tanh_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].gelu_tanh_and_mul(out, x)


def fatrelu_and_mul(out: torch.Tensor, x: torch.Tensor, threshold: float = 0.0) -> None:
    _ops()[0].fatrelu_and_mul(out, x, threshold)


def gelu_fast(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].gelu_fast(out, x)


def gelu_new(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].gelu_new(out, x)


def gelu_quick(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops()[0].gelu_quick(out, x)


# aphrodite/_custom_ops.py:195-208
def batched_rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                             cos_sin_cache: torch.Tensor, is_neox: bool, rot_dim: int,
                             cos_sin_cache_offsets: torch.Tensor) -> None:
    _ops()[0].batched_rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim,
                                       cos_sin_cache_offsets)


# aphrodite/_custom_ops.py:222-232
def advance_step_flashattn(num_seqs: int, num_queries: int, block_size: int, input_tokens: torch.Tensor,
                           sampled_token_ids: torch.Tensor, input_positions: torch.Tensor, seq_lens: torch.Tensor,
                           slot_mapping: torch.Tensor, block_tables: torch.Tensor) -> None:
    """Advance a step on GPU for existing inputs for a multi-step runner"""
    return _ops()[0].advance_step_flashattn(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids,
                                            input_positions, seq_lens, slot_mapping, block_tables)


# aphrodite/_custom_ops.py:236-251
def advance_step_flashinfer(num_seqs: int, num_queries: int, block_size: int, input_tokens: torch.Tensor,
                            sampled_token_ids: torch.Tensor, input_positions: torch.Tensor, seq_lens: torch.Tensor,
                            slot_mapping: torch.Tensor, block_tables: torch.Tensor, paged_kv_indices: torch.Tensor,
                            paged_kv_indptr: torch.Tensor, paged_kv_last_page_len: torch.Tensor,
                            block_table_bound: torch.Tensor) -> None:
    return _ops()[0].advance_step_flashinfer(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids,
                                             input_positions, seq_lens, slot_mapping, block_tables, paged_kv_indices,
                                             paged_kv_indptr, paged_kv_last_page_len, block_table_bound)


# aphrodite/_custom_ops.py:1475-1518: the MLA cache ops
def concat_and_cache_mla(kv_c: torch.Tensor, k_pe: torch.Tensor, kv_cache: torch.Tensor, slot_mapping: torch.Tensor,
                         kv_cache_dtype: str, scale: torch.Tensor) -> None:
    _ops()[1].concat_and_cache_mla(kv_c, k_pe, kv_cache, slot_mapping, kv_cache_dtype, scale)


def copy_blocks_mla(kv_caches: list[torch.Tensor], block_mapping: torch.Tensor) -> None:
    _ops()[1].copy_blocks_mla(kv_caches, block_mapping)


def gather_cache(src_cache: torch.Tensor, dst: torch.Tensor, block_table: torch.Tensor, cu_seq_lens: torch.Tensor,
                 batch_size: int, seq_starts: Optional[torch.Tensor] = None) -> None:
    _ops()[1].gather_cache(src_cache, dst, block_table, cu_seq_lens, batch_size, seq_starts)
