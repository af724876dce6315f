"""Host side of the Marlin weight format: the helpers a GPTQ / AWQ checkpoint loader needs around
`gptq_marlin_repack` / `awq_marlin_repack` / `gptq_marlin_gemm`, with the names, arguments and error behaviour of the
reference's `aphrodite/quantization/utils/marlin_utils.py` (:30-420) and the column packers of `quant_utils.py` (:485-540),
so that its quantisation layers (`gptq_marlin.py`, `awq_marlin.py`) find what they import.

Everything here is index arithmetic on torch tensors and runs on whatever device the checkpoint tensors live on (the
reference's packers go through numpy on the host). The permutations are written in closed form:

* group scales  - inside every block of 64 columns, output position p holds column (p % 8) * 8 + p // 8 (the order in which
  the MMA fragments of four 16-column tiles want their scales);
* channel scales - inside every block of 32 columns, output position p holds column
  2 * (p // 8) + (p % 2) + 8 * ((p % 8) // 2);
* zero points   - the group-scale order, then inside every `32 / num_bits` columns the dequantiser's interleave
  (4 bit: 0 2 4 6 1 3 5 7; 8 bit: 0 2 1 3), then packed along the columns into int32.

Pinned against the reference's own Python by tests/test_marlin_utils.py (fixtures made by tests/golden/make_marlin_utils_golden.py).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from .tp import (GPTQ_MARLIN_MAX_PARALLEL, GPTQ_MARLIN_MIN_THREAD_K, GPTQ_MARLIN_MIN_THREAD_N, GPTQ_MARLIN_TILE,  # noqa: F401
                 check_marlin_supports_shape, marlin_is_k_full, marlin_make_workspace, marlin_repeat_scales_on_all_ranks,
                 verify_marlin_supports_shape)

MARLIN_SUPPORTED_GROUP_SIZES = [-1, 32, 64, 128]
USE_FP32_REDUCE_DEFAULT = True


@dataclass(frozen=True)
class ScalarType:
    """The four weight types of this path, with the id the reference's ScalarType computes (scalar_type.py:133-159:
    exponent / mantissa / signed / bias / finite-only / NaN representation packed LSB-first as 8 / 8 / 1 / 32 / 1 / 8 bits)."""
    name: str
    size_bits: int
    bias: int

    @property
    def id(self) -> int:
        return (self.size_bits << 8) | ((self.bias & 0xFFFFFFFF) << 17) | (1 << 50)

    def has_bias(self) -> bool:
        return self.bias != 0

    def __str__(self) -> str:
        return self.name


class scalar_types:
    uint4 = ScalarType("uint4", 4, 0)
    uint8 = ScalarType("uint8", 8, 0)
    uint4b8 = ScalarType("uint4b8", 4, 8)
    uint8b128 = ScalarType("uint8b128", 8, 128)


def query_marlin_supported_quant_types(has_zp: bool, device_capability: Optional[int] = None) -> List[ScalarType]:
    """marlin_utils.py:30-49. AWQ style (unsigned + run-time zero point) or GPTQ style (unsigned with a symmetric bias)."""
    if device_capability is None:
        device_capability = 100  # this library is sm_100a only
    if device_capability < 80:
        return []
    return [scalar_types.uint4, scalar_types.uint8] if has_zp else [scalar_types.uint4b8, scalar_types.uint8b128]


def _check_marlin_supported(quant_type, group_size: Optional[int], has_zp: bool,
                            device_capability: Optional[int] = None) -> Tuple[bool, Optional[str]]:
    supported = query_marlin_supported_quant_types(has_zp, device_capability)
    if getattr(quant_type, "id", quant_type) not in [t.id for t in supported]:
        return (False, f"Marlin does not support weight_bits = {quant_type}. Only types = {supported} are supported (for "
                       f"group_size = {group_size}, device_capability = {device_capability}, zp = {has_zp}).")
    if group_size is None or group_size not in MARLIN_SUPPORTED_GROUP_SIZES:
        return (False, f"Marlin does not support group_size = {group_size}. Only group_sizes = "
                       f"{MARLIN_SUPPORTED_GROUP_SIZES} are supported.")
    return True, None


def check_marlin_supported(quant_type, group_size: int, has_zp: bool = False, device_capability: Optional[int] = None) -> bool:
    return _check_marlin_supported(quant_type, group_size, has_zp, device_capability)[0]


def verify_marlin_supported(quant_type, group_size: int, has_zp: bool = False) -> None:
    ok, msg = _check_marlin_supported(quant_type, group_size, has_zp)
    if not ok:
        raise ValueError(msg)


def marlin_make_empty_g_idx(device) -> torch.Tensor:
    return torch.nn.Parameter(torch.empty(0, dtype=torch.int, device=device), requires_grad=False)


def marlin_make_empty_zp(device) -> torch.Tensor:
    return torch.nn.Parameter(torch.empty(0, dtype=torch.int, device=device), requires_grad=False)


def marlin_sort_g_idx(g_idx: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Rows sorted by quantisation group (act-order): (sorted g_idx, the int32 permutation that sorts it)."""
    order = torch.argsort(g_idx).to(torch.int)
    return g_idx[order], order


def get_scale_perms() -> Tuple[List[int], List[int]]:
    """(order of 64 group-scale columns, order of 32 channel-scale columns) - marlin_utils.py:204-213 as closed forms."""
    grouped = [(p % 8) * 8 + p // 8 for p in range(64)]
    single = [2 * (p // 8) + (p % 2) + 8 * ((p % 8) // 2) for p in range(32)]
    return grouped, single


def _permute_blocks(t: torch.Tensor, order: List[int]) -> torch.Tensor:
    idx = torch.tensor(order, dtype=torch.long, device=t.device)
    return t.reshape(-1, len(order)).index_select(1, idx)


def marlin_permute_scales(s: torch.Tensor, size_k: int, size_n: int, group_size: int) -> torch.Tensor:
    grouped, single = get_scale_perms()
    order = grouped if (group_size < size_k and group_size != -1) else single
    return _permute_blocks(s, order).reshape(-1, size_n).contiguous()


def marlin_moe_permute_scales(s: torch.Tensor, size_k: int, size_n: int, group_size: int) -> torch.Tensor:
    return torch.stack([marlin_permute_scales(s[e], size_k, size_n, group_size) for e in range(s.shape[0])])


def get_pack_factor(num_bits: int) -> int:
    assert 32 % num_bits == 0, f"Unsupported num_bits = {num_bits}"
    return 32 // num_bits


def pack_cols(q_w: torch.Tensor, num_bits: int, size_k: int, size_n: int) -> torch.Tensor:
    """[size_k, size_n] values below 2^num_bits -> [size_k, size_n / pack] int32, value i of a word in bits [i*num_bits, ...)."""
    assert q_w.shape == (size_k, size_n)
    pack = get_pack_factor(num_bits)
    assert size_n % pack == 0
    shifts = torch.arange(pack, dtype=torch.int64, device=q_w.device) * num_bits
    words = (q_w.to(torch.int64).reshape(size_k, size_n // pack, pack) << shifts).sum(dim=-1)
    words = torch.where(words >= 2**31, words - 2**32, words)  # the bit pattern as int32
    return words.to(torch.int32).contiguous()


def unpack_cols(packed_q_w: torch.Tensor, num_bits: int, size_k: int, size_n: int) -> torch.Tensor:
    pack = get_pack_factor(num_bits)
    assert size_n % pack == 0
    assert packed_q_w.shape == (size_k, size_n // pack), \
        "packed_q_w.shape = {} size_k = {}, size_n = {} pack_Factor = {}".format(packed_q_w.shape, size_k, size_n, pack)
    shifts = torch.arange(pack, dtype=torch.int64, device=packed_q_w.device) * num_bits
    words = packed_q_w.to(torch.int64) & 0xFFFFFFFF
    vals = (words.unsqueeze(-1) >> shifts) & ((1 << num_bits) - 1)
    return vals.reshape(size_k, size_n).to(torch.int32).contiguous()


def _zp_interleave(num_bits: int) -> List[int]:
    if num_bits == 4:
        return [0, 2, 4, 6, 1, 3, 5, 7]
    if num_bits == 8:
        return [0, 2, 1, 3]
    raise Exception("num_bits must be 4 or 8, got {}".format(num_bits))


def marlin_zero_points(zp: torch.Tensor, size_k: int, size_n: int, num_bits: int) -> torch.Tensor:
    """Unpacked zero points [groups, size_n] -> the packed layout the GEMM's dequantiser reads (`size_k` = number of rows of zp)."""
    grouped, _ = get_scale_perms()
    zp = _permute_blocks(zp, grouped)
    zp = _permute_blocks(zp, _zp_interleave(num_bits)).reshape(-1, size_n).contiguous()
    return pack_cols(zp, num_bits, size_k, size_n)


def awq_to_marlin_zero_points(q_zp_packed: torch.Tensor, size_k: int, size_n: int, num_bits: int) -> torch.Tensor:
    """AWQ checkpoints pack their zero points along the columns, already interleaved for AWQ's own dequantiser: unpack,
    undo that interleave, then lay them out for Marlin."""
    q_zp = unpack_cols(q_zp_packed, num_bits, size_k, size_n)
    inter = _zp_interleave(num_bits)
    undo = sorted(range(len(inter)), key=inter.__getitem__)
    q_zp = _permute_blocks(q_zp, undo).reshape(-1, size_n).contiguous()
    return marlin_zero_points(q_zp, size_k, size_n, num_bits)


def moe_awq_to_marlin_zero_points(q_zp_packed: torch.Tensor, size_k: int, size_n: int, num_bits: int) -> torch.Tensor:
    return torch.stack([awq_to_marlin_zero_points(q_zp_packed[e], size_k, size_n, num_bits) for e in range(q_zp_packed.shape[0])])


def should_use_atomic_add_reduce(m: int, n: int, k: int, device, dtype) -> bool:
    """The reference can reduce its split-K partials with atomicAdd (marlin_utils.py:305-320). This library merges them in
    shared memory of a thread-block cluster or in rank order through a scratch buffer - deterministic either way - so the
    flag is accepted by `gptq_marlin_gemm` and never asked for."""
    return False


def _apply(input: torch.Tensor, weight, weight_scale, weight_zp, g_idx, g_idx_sort_indices, workspace, wtype,
           output_size_per_partition: int, input_size_per_partition: int, has_zp: bool, is_k_full: bool, bias,
           use_fp32_reduce: bool) -> torch.Tensor:
    from . import _custom_ops as ops
    x = input.reshape(-1, input.shape[-1])
    out = ops.gptq_marlin_gemm(x, weight, weight_scale, weight_zp, g_idx, g_idx_sort_indices, workspace, wtype,
                               size_m=x.shape[0], size_n=output_size_per_partition, size_k=input_size_per_partition,
                               is_k_full=is_k_full, has_zp=has_zp, use_atomic_add=False, use_fp32_reduce=use_fp32_reduce,
                               is_zp_float=False)
    if bias is not None:
        out.add_(bias)
    return out.reshape(input.shape[:-1] + (output_size_per_partition,))


def apply_gptq_marlin_linear(input: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor, weight_zp: torch.Tensor,
                             g_idx: torch.Tensor, g_idx_sort_indices: torch.Tensor, workspace: torch.Tensor, wtype,
                             output_size_per_partition: int, input_size_per_partition: int, has_zp: bool, is_k_full: bool,
                             bias: Optional[torch.Tensor] = None, use_fp32_reduce: bool = USE_FP32_REDUCE_DEFAULT) -> torch.Tensor:
    """marlin_utils.py:322-367: y = x @ dequant(W) (+ bias) for any leading shape of x."""
    return _apply(input, weight, weight_scale, weight_zp, g_idx, g_idx_sort_indices, workspace, wtype,
                  output_size_per_partition, input_size_per_partition, has_zp, is_k_full, bias, use_fp32_reduce)


def apply_awq_marlin_linear(input: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor, weight_zp: torch.Tensor,
                            g_idx: torch.Tensor, g_idx_sort_indices: torch.Tensor, workspace: torch.Tensor, quant_type,
                            output_size_per_partition: int, input_size_per_partition: int, bias: Optional[torch.Tensor] = None,
                            use_fp32_reduce: bool = USE_FP32_REDUCE_DEFAULT) -> torch.Tensor:
    """marlin_utils.py:369-420: the AWQ flavour - zero points, never act-order."""
    return _apply(input, weight, weight_scale, weight_zp, g_idx, g_idx_sort_indices, workspace, quant_type,
                  output_size_per_partition, input_size_per_partition, True, True, bias, use_fp32_reduce)
