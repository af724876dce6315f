"""Tensor-parallel plumbing of the decode hot path: how heads and Marlin-format weights are split
over ranks, and the one exchange step (all_reduce after the row-parallel GEMMs) over `torch.distributed`.

Mirrors, with the same names and error behaviour where a name exists in the reference:
  * /root/reference/aphrodite/quantization/utils/marlin_utils.py:96-122 (verify_marlin_supports_shape),
    :160-185 (marlin_make_workspace, marlin_is_k_full, marlin_repeat_scales_on_all_ranks);
  * /root/reference/aphrodite/modeling/layers/linear.py:474-484 (column parallel: split N, no exchange),
    :1279-1283 (row parallel: split K, all_reduce of the partial outputs);
  * /root/reference/aphrodite/modeling/models/llama.py:118-127 (q heads split, kv heads split or replicated).

The reference shards the checkpoint tensors and repacks each shard; here a tensor that is ALREADY in the Marlin wire
format is sliced, which gives the same bytes because every permutation of the format is local to a 16 x 64 tile of
weights (64 columns of scales, 8 int32 columns of zero points) and shard boundaries fall on those tiles.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.distributed as dist

GPTQ_MARLIN_MIN_THREAD_N = 64
GPTQ_MARLIN_MIN_THREAD_K = 128
GPTQ_MARLIN_MAX_PARALLEL = 16
GPTQ_MARLIN_TILE = 16


def verify_marlin_supports_shape(output_size_per_partition: int, input_size_per_partition: int, input_size: int,
                                 group_size: int) -> None:
    if output_size_per_partition % GPTQ_MARLIN_MIN_THREAD_N != 0:
        raise ValueError(f"Weight output_size_per_partition = {output_size_per_partition} is not divisible by "
                         f" min_thread_n = {GPTQ_MARLIN_MIN_THREAD_N}. Consider reducing tensor_parallel_size or "
                         "running with --quantization gptq.")
    if input_size_per_partition % GPTQ_MARLIN_MIN_THREAD_K != 0:
        raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition} is not divisible by "
                         f"min_thread_k = {GPTQ_MARLIN_MIN_THREAD_K}. Consider reducing tensor_parallel_size or "
                         "running with --quantization gptq.")
    if group_size < input_size and input_size_per_partition % group_size != 0:
        raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition} is not divisible by "
                         f"group_size = {group_size}. Consider reducing tensor_parallel_size or running with "
                         "--quantization gptq.")


def check_marlin_supports_shape(output_size_per_partition: int, input_size_per_partition: int, input_size: int,
                                group_size: int) -> Tuple[bool, Optional[str]]:
    try:
        verify_marlin_supports_shape(output_size_per_partition, input_size_per_partition, input_size, group_size)
    except ValueError as e:
        return False, str(e)
    return True, None


def marlin_make_workspace(output_size_per_partition: int, device) -> torch.Tensor:
    n = (output_size_per_partition // GPTQ_MARLIN_MIN_THREAD_N) * GPTQ_MARLIN_MAX_PARALLEL
    return torch.zeros(n, dtype=torch.int, device=device, requires_grad=False)


def marlin_is_k_full(act_order: bool, is_row_parallel: bool) -> bool:
    return (not act_order) or (act_order and not is_row_parallel)


def marlin_repeat_scales_on_all_ranks(act_order: bool, group_size: int, is_row_parallel: bool) -> bool:
    return act_order or (group_size == -1 and is_row_parallel)


def shard_heads(num_heads: int, num_kv_heads: int, tp_size: int) -> Tuple[int, int]:
    """(q heads, kv heads) held by one rank; kv heads are replicated once tp_size exceeds their number."""
    if num_heads % tp_size != 0:
        raise ValueError(f"total_num_heads = {num_heads} is not divisible by tensor_parallel_size = {tp_size}")
    if num_kv_heads >= tp_size:
        if num_kv_heads % tp_size != 0:
            raise ValueError(f"total_num_kv_heads = {num_kv_heads} is not divisible by tensor_parallel_size = {tp_size}")
    elif tp_size % num_kv_heads != 0:
        raise ValueError(f"tensor_parallel_size = {tp_size} is not divisible by total_num_kv_heads = {num_kv_heads}")
    return num_heads // tp_size, max(1, num_kv_heads // tp_size)


@dataclass
class MarlinShard:
    """One rank's slice of a W4A16 weight in the Marlin wire format."""
    q_weight: torch.Tensor            # [K_l / 16, N_l * 2] int32
    scales: torch.Tensor              # [K_l / g, N_l] or [1, N_l]
    zeros: Optional[torch.Tensor]     # [K_l / g, N_l / 8] int32 or None
    size_k: int
    size_n: int
    group_size: int                   # -1 = channelwise
    is_row_parallel: bool

    def workspace(self) -> torch.Tensor:
        return marlin_make_workspace(self.size_n, self.q_weight.device)


def shard_marlin_weight(q_weight: torch.Tensor, scales: torch.Tensor, zeros: Optional[torch.Tensor], size_k: int,
                        size_n: int, group_size: int, tp_size: int, tp_rank: int, parallel: str) -> MarlinShard:
    """Slice a Marlin-format weight for `tp_rank` of `tp_size`: parallel = "column" splits N, "row" splits K."""
    if parallel not in ("column", "row"):
        raise ValueError(f"parallel must be 'column' or 'row', got {parallel!r}")
    if not 0 <= tp_rank < tp_size:
        raise ValueError(f"tp_rank = {tp_rank} outside [0, {tp_size})")
    row = parallel == "row"
    if (size_n if not row else size_k) % tp_size != 0:
        raise ValueError(f"{'size_k' if row else 'size_n'} = {size_k if row else size_n} is not divisible by "
                         f"tensor_parallel_size = {tp_size}")
    k_l = size_k // tp_size if row else size_k
    n_l = size_n if row else size_n // tp_size
    eff_group = size_k if group_size == -1 else group_size
    verify_marlin_supports_shape(n_l, k_l, size_k, eff_group)
    if q_weight.shape != (size_k // GPTQ_MARLIN_TILE, size_n * GPTQ_MARLIN_TILE // 8):
        raise ValueError(f"q_weight has shape {tuple(q_weight.shape)}, expected "
                         f"{(size_k // GPTQ_MARLIN_TILE, size_n * GPTQ_MARLIN_TILE // 8)}")
    if row:
        r0 = tp_rank * k_l // GPTQ_MARLIN_TILE
        qw = q_weight[r0:r0 + k_l // GPTQ_MARLIN_TILE]
        if marlin_repeat_scales_on_all_ranks(False, group_size, True):
            sc, zp = scales, zeros
        else:
            g0 = tp_rank * k_l // group_size
            sc = scales[g0:g0 + k_l // group_size]
            zp = None if zeros is None else zeros[g0:g0 + k_l // group_size]
    else:
        c0 = tp_rank * n_l
        qw = q_weight[:, c0 * 2:(c0 + n_l) * 2]
        sc = scales[:, c0:c0 + n_l]
        zp = None if zeros is None else zeros[:, c0 // 8:(c0 + n_l) // 8]
    return MarlinShard(qw.contiguous(), sc.contiguous(), None if zp is None else zp.contiguous(), k_l, n_l, group_size, row)


class TPGroup:
    """The exchange step: all_reduce(sum) of the row-parallel partial outputs, in place, on the current stream.

    world_size 1 never touches torch.distributed. On GPUs small messages (<= 8 MiB, the reference's limit,
    custom_all_reduce.py:53) go through the one-shot peer-memory kernel of csrc/all_reduce.cu when
    `enable_custom_all_reduce()` succeeded; everything else - and every CPU test (gloo) - goes through the process
    group's own all_reduce (NCCL over NVLink / NVSwitch on GPUs). APH_CUSTOM_AR=0 keeps NCCL for everything.
    """

    def __init__(self, group: Optional[dist.ProcessGroup] = None):
        if dist.is_available() and dist.is_initialized():
            self.group = group if group is not None else dist.group.WORLD
            self.world_size = dist.get_world_size(self.group)
            self.rank = dist.get_rank(self.group)
        else:
            self.group, self.world_size, self.rank = None, 1, 0
        self.custom = None
        self.custom_error: Optional[str] = None

    @classmethod
    def single(cls) -> "TPGroup":
        """A group of one rank whatever torch.distributed says (the unsharded model next to a sharded one)."""
        g = cls.__new__(cls)
        g.group, g.world_size, g.rank, g.custom, g.custom_error = None, 1, 0, None, None
        return g

    def enable_custom_all_reduce(self, device, max_size: int = 8 * 1024 * 1024) -> bool:
        """Set up the peer-memory all-reduce; returns False (and keeps NCCL) when the world size or the platform does
        not support it - the reason is kept in `custom_error`."""
        import os
        if self.world_size == 1 or os.environ.get("APH_CUSTOM_AR", "1") == "0":
            return False
        try:
            from .custom_all_reduce import OneShotAllReduce
            self.custom = OneShotAllReduce(self.group, device, max_size)
        except Exception as e:  # no NVLink peer access, unsupported world size, ...: NCCL serves the exchange
            self.custom, self.custom_error = None, f"{type(e).__name__}: {e}"
        return self.custom is not None

    def describe(self, x: torch.Tensor) -> str:
        """Which exchange a message like `x` takes (for benchmark records)."""
        if self.world_size == 1:
            return "none"
        if self.custom is None or not self.custom.should_use(x):
            return "NCCL" + (f" ({self.custom_error})" if self.custom_error else "")
        kind = self.custom.mode(x.numel() * x.element_size()).replace("_", "-")
        return f"{kind} peer-memory kernel fused with the residual add + RMSNorm (csrc/all_reduce.cu)"

    def all_reduce(self, x: torch.Tensor) -> torch.Tensor:
        if self.world_size > 1:
            if self.custom is not None and self.custom.should_use(x):
                return self.custom.all_reduce(x)
            dist.all_reduce(x, op=dist.ReduceOp.SUM, group=self.group)
        return x

    def all_reduce_add_rms_norm(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
        """In place: residual += all_reduce(x); x = rms_norm(residual) * weight - what every row-parallel output of a
        decoder layer goes through (llama.py:270-292). One fused kernel over peer memory when eligible
        (APH_FUSED_AR_NORM=0 keeps the two steps apart), else all_reduce + the fused_add_rms_norm op."""
        import os
        from . import _custom_ops as ops
        if (self.world_size > 1 and self.custom is not None and os.environ.get("APH_FUSED_AR_NORM", "1") != "0"
                and self.custom.should_fuse_norm(x, residual, weight)):
            self.custom.all_reduce_add_rms_norm(x, residual, weight, epsilon)
            return
        self.all_reduce(x)
        ops.fused_add_rms_norm(x, residual, weight, epsilon)
