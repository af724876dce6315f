"""A Llama decode step built from this library's kernels only, tensor-parallel over `torch.distributed`.

This is the caller side of SURVEY.md 3.1 reduced to what the hot path needs: per layer
    h   = fused_add_rms_norm(x, residual)                      (rms_norm for the first layer)
    qkv = W4A16 GEMM (column parallel)  ->  rotary_embedding(q, k)  ->  reshape_and_cache(k, v)
                                        ->  paged attention (this rank's heads)
    o   = W4A16 GEMM (row parallel)     ->  all_reduce + fused_add_rms_norm(o, residual)  (one kernel over peer memory)
    gu  = W4A16 GEMM (column parallel)  ->  silu_and_mul
    dn  = W4A16 GEMM (row parallel)     ->  all_reduce + the next layer's fused_add_rms_norm
and a final norm: 11 kernels per layer, every one launched with programmatic dependent launch, no
torch elementwise kernels in between. Weights are synthetic. One step can be captured in a CUDA graph, the NCCL
all_reduces included, which is how the reference runs decode (`cuda_communicator.py`).

Reference call sites mirrored: modeling/models/llama.py:100-215 (layer structure and head split),
modeling/layers/linear.py:474-484,1279-1283, attention/ops/paged_attn.py:60-160.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

from . import _custom_ops as ops
from .paged_attn import PagedAttention
from .tp import (MarlinShard, TPGroup, marlin_make_workspace, shard_heads, shard_marlin_weight,
                 verify_marlin_supports_shape)

SCALAR_TYPE_UINT4B8 = 1125899907892224  # kernels/core/scalar_type.hpp:100-138


@dataclass
class DecodeConfig:
    hidden: int
    intermediate: int
    num_heads: int
    num_kv_heads: int
    head_size: int
    num_layers: int
    group_size: int = 128
    block_size: int = 16
    kv_cache_dtype: str = "fp8"  # "auto" | "fp8" | "fp8_e4m3" | "fp8_e5m2"
    dtype: torch.dtype = torch.bfloat16
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0

    # benchmark_shapes.py:3-8 / :57-62
    @staticmethod
    def llama3_8b(num_layers: int = 32, **kw) -> "DecodeConfig":
        return DecodeConfig(4096, 14336, 32, 8, 128, num_layers, **kw)

    @staticmethod
    def llama3_70b(num_layers: int = 80, **kw) -> "DecodeConfig":
        return DecodeConfig(8192, 28672, 64, 8, 128, num_layers, **kw)


def _synthetic_marlin(size_k: int, size_n: int, group_size: int, dtype, device, gen, row_parallel: bool) -> MarlinShard:
    """Random bytes in the Marlin layout: any int32 pattern is a valid packed 4-bit tile."""
    groups = 1 if group_size == -1 else size_k // group_size
    qw = torch.randint(-2**31, 2**31 - 1, (size_k // 16, size_n * 2), dtype=torch.int32, device=device, generator=gen)
    sc = (torch.rand(groups, size_n, device=device, generator=gen) * 0.0005 + 0.00075).to(dtype)  # layer gain < 1
    return MarlinShard(qw, sc, None, size_k, size_n, group_size, row_parallel)


class MarlinLinear:
    """apply() of the reference's GPTQMarlinLinearMethod (gptq_marlin.py:330-360) for one shard."""

    def __init__(self, shard: MarlinShard):
        self.w = shard
        self.workspace = marlin_make_workspace(shard.size_n, shard.q_weight.device)
        self.empty = torch.empty(0, dtype=torch.int32, device=shard.q_weight.device)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        w = self.w
        zeros = self.empty if w.zeros is None else w.zeros
        return ops.gptq_marlin_gemm(x, w.q_weight, w.scales, zeros, self.empty, self.empty, self.workspace,
                                    SCALAR_TYPE_UINT4B8, x.shape[0], w.size_n, w.size_k, True, w.zeros is not None, False,
                                    True, False)


class DecodeLayer:
    def __init__(self, cfg: DecodeConfig, tp: TPGroup, num_blocks: int, device, gen, weights_from: "Optional[DecodeLayer]" = None):
        self.cfg, self.tp = cfg, tp
        t = tp.world_size
        self.h, self.kvh = shard_heads(cfg.num_heads, cfg.num_kv_heads, t)
        d = cfg.head_size
        self.q_size, self.kv_size = self.h * d, self.kvh * d
        inter_l = cfg.intermediate // t
        g = cfg.group_size
        for n_l, k_l, k in ((self.q_size + 2 * self.kv_size, cfg.hidden, cfg.hidden), (cfg.hidden, self.q_size, cfg.num_heads * d),
                            (2 * inter_l, cfg.hidden, cfg.hidden), (cfg.hidden, inter_l, cfg.intermediate)):
            verify_marlin_supports_shape(n_l, k_l, k, k if g == -1 else g)
        mk = lambda k, n, row: MarlinLinear(_synthetic_marlin(k, n, g, cfg.dtype, device, gen, row))
        if weights_from is not None:  # another batch size over the same (immutable) weights
            self.qkv, self.o, self.gate_up, self.down = weights_from.qkv, weights_from.o, weights_from.gate_up, weights_from.down
        else:
            self.qkv = mk(cfg.hidden, self.q_size + 2 * self.kv_size, False)
            self.o = mk(self.q_size, cfg.hidden, True)
            self.gate_up = mk(cfg.hidden, 2 * inter_l, False)
            self.down = mk(inter_l, cfg.hidden, True)
        fp8 = cfg.kv_cache_dtype != "auto"
        shape = PagedAttention.get_kv_cache_shape(num_blocks, cfg.block_size, self.kvh, d)
        if fp8:
            self.kv_cache = torch.randint(0, 120, shape, dtype=torch.uint8, device=device, generator=gen)
        else:
            self.kv_cache = (torch.rand(shape, device=device, generator=gen) - 0.5).to(cfg.dtype)
        self.key_cache, self.value_cache = PagedAttention.split_kv_cache(self.kv_cache, self.kvh, d)
        self.k_scale = torch.ones((), dtype=torch.float32, device=device)
        self.v_scale = torch.ones((), dtype=torch.float32, device=device)
        self.scale = d ** -0.5
        ln = lambda: (1.0 + 0.1 * torch.randn(cfg.hidden, device=device, generator=gen)).to(cfg.dtype)
        self.input_layernorm, self.post_attention_layernorm = ln(), ln()

    @classmethod
    def shard_of(cls, full: "DecodeLayer", tp: TPGroup) -> "DecodeLayer":
        """This rank's shard of an UNSHARDED layer, cut the way the reference cuts a checkpoint (linear.py:474-484 column
        parallel with q / k / v and gate / up split separately, :1279-1283 row parallel; llama.py:118-127 heads): the
        sharded step must reproduce the unsharded one up to the summation order of the all-reduce."""
        cfg, t, r = full.cfg, tp.world_size, tp.rank
        self = cls.__new__(cls)
        self.cfg, self.tp = cfg, tp
        self.h, self.kvh = shard_heads(cfg.num_heads, cfg.num_kv_heads, t)
        d, g = cfg.head_size, cfg.group_size
        self.q_size, self.kv_size = self.h * d, self.kvh * d
        kv_parts = min(t, cfg.num_kv_heads)  # kv heads are split t ways, or replicated once t exceeds their number
        kv_rank = r * kv_parts // t           # which of the kv_parts shards this rank holds

        def columns(w: MarlinShard, c0: int, c1: int) -> MarlinShard:  # a block of output columns (multiples of 64)
            return MarlinShard(w.q_weight[:, 2 * c0:2 * c1].contiguous(), w.scales[:, c0:c1].contiguous(), None, w.size_k,
                               c1 - c0, g, False)

        def column_parallel(blocks) -> MarlinLinear:  # [(block, parts, part index)] -> this rank's columns, concatenated
            sh = [shard_marlin_weight(b.q_weight, b.scales, None, b.size_k, b.size_n, g, parts, idx, "column") for b, parts, idx in blocks]
            return MarlinLinear(MarlinShard(torch.cat([x.q_weight for x in sh], 1).contiguous(), torch.cat([x.scales for x in sh], 1).contiguous(),
                                            None, sh[0].size_k, sum(x.size_n for x in sh), g, False))

        def row_parallel(w: MarlinShard) -> MarlinLinear:
            return MarlinLinear(shard_marlin_weight(w.q_weight, w.scales, None, w.size_k, w.size_n, g, t, r, "row"))

        fq, fkv = full.q_size, full.kv_size
        wq = full.qkv.w
        self.qkv = column_parallel([(columns(wq, 0, fq), t, r), (columns(wq, fq, fq + fkv), kv_parts, kv_rank),
                                    (columns(wq, fq + fkv, fq + 2 * fkv), kv_parts, kv_rank)])
        self.o = row_parallel(full.o.w)
        wg, inter = full.gate_up.w, cfg.intermediate
        self.gate_up = column_parallel([(columns(wg, 0, inter), t, r), (columns(wg, inter, 2 * inter), t, r)])
        self.down = row_parallel(full.down.w)
        k0 = kv_rank * self.kvh
        kc = full.key_cache[:, k0:k0 + self.kvh].contiguous()
        vc = full.value_cache[:, k0:k0 + self.kvh].contiguous()
        self.kv_cache = torch.stack([kc.reshape(kc.shape[0], -1), vc.reshape(vc.shape[0], -1)]).contiguous()
        self.key_cache, self.value_cache = PagedAttention.split_kv_cache(self.kv_cache, self.kvh, d)
        self.k_scale, self.v_scale, self.scale = full.k_scale, full.v_scale, full.scale
        self.input_layernorm, self.post_attention_layernorm = full.input_layernorm, full.post_attention_layernorm
        return self

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor], positions, cos_sin_cache, slot_mapping,
                block_tables, seq_lens, max_seq_len: int):
        """modeling/models/llama.py:270-292: returns (mlp output, residual stream)."""
        cfg, d = self.cfg, self.cfg.head_size
        if residual is None:
            residual = x
            h = torch.empty_like(x)
            ops.rms_norm(h, x, self.input_layernorm, cfg.rms_norm_eps)
        else:  # x is the previous layer's row-parallel partial: its all-reduce is fused with this add + norm
            self.tp.all_reduce_add_rms_norm(x, residual, self.input_layernorm, cfg.rms_norm_eps)
            h = x
        qkv = self.qkv(h)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        ops.rotary_embedding(positions, q, k, d, cos_sin_cache, True)  # in place on the slices of qkv
        q = q.reshape(-1, self.h, d)
        PagedAttention.write_to_paged_cache(k.reshape(-1, self.kvh, d), v.reshape(-1, self.kvh, d), self.key_cache,
                                            self.value_cache, slot_mapping, cfg.kv_cache_dtype, self.k_scale, self.v_scale)
        a = PagedAttention.forward_decode(q, self.key_cache, self.value_cache, block_tables, seq_lens, max_seq_len,
                                          cfg.kv_cache_dtype, self.kvh, self.scale, None, self.k_scale, self.v_scale,
                                          tp_rank=self.tp.rank)
        o = self.o(a.reshape(-1, self.q_size))
        self.tp.all_reduce_add_rms_norm(o, residual, self.post_attention_layernorm, cfg.rms_norm_eps)
        gu = self.gate_up(o)
        act = torch.empty(gu.shape[0], gu.shape[1] // 2, dtype=gu.dtype, device=gu.device)
        ops.silu_and_mul(act, gu)
        return self.down(act), residual  # partial: reduced by the next all_reduce_add_rms_norm


class DecodeModel:
    """`num_layers` layers over one set of paged-cache metadata; `step()` runs one decode step for the whole batch."""

    def __init__(self, cfg: DecodeConfig, batch: int, ctx: int, tp: Optional[TPGroup] = None, device="cuda", seed: int = 0,
                 weights_from: "Optional[DecodeModel]" = None):
        self.cfg, self.batch, self.ctx = cfg, batch, ctx
        self.tp = tp if tp is not None else TPGroup()
        if self.tp.world_size > 1 and self.tp.custom is None and torch.device(device).type == "cuda":
            self.tp.enable_custom_all_reduce(device)  # peer-memory all-reduce for the 1 MiB messages; NCCL if unavailable
        gen = torch.Generator(device=device).manual_seed(seed)
        bs = cfg.block_size
        blocks_per_seq = (ctx + bs - 1) // bs
        num_blocks = batch * blocks_per_seq
        perm = torch.randperm(num_blocks, device=device, generator=gen).to(torch.int32)
        self.block_tables = perm.view(batch, blocks_per_seq).contiguous()
        self.seq_lens = torch.full((batch,), ctx, dtype=torch.int32, device=device)
        last = self.block_tables[:, (ctx - 1) // bs].to(torch.int64)
        self.slot_mapping = last * bs + (ctx - 1) % bs  # the new token is the last position of each sequence
        self.layers: List[DecodeLayer] = [
            DecodeLayer(cfg, self.tp, num_blocks, device, gen, weights_from.layers[i] if weights_from is not None else None)
            for i in range(cfg.num_layers)]
        self.x = (torch.randn(batch, cfg.hidden, device=device, generator=gen) * 0.05).to(cfg.dtype)
        self.norm = (1.0 + 0.1 * torch.randn(cfg.hidden, device=device, generator=gen)).to(cfg.dtype)
        self.positions = torch.full((batch,), ctx - 1, dtype=torch.int64, device=device)
        # modeling/layers/rotary_embedding.py:114-131: [max_position, rot_dim] = cos half | sin half
        inv_freq = 1.0 / (cfg.rope_theta ** (torch.arange(0, cfg.head_size, 2, dtype=torch.float, device=device) / cfg.head_size))
        freqs = torch.einsum("i,j->ij", torch.arange(ctx, dtype=torch.float, device=device), inv_freq)
        self.cos_sin_cache = torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(cfg.dtype)
        self.out = torch.empty_like(self.x)
        self.graph: Optional[torch.cuda.CUDAGraph] = None

    @classmethod
    def shard_of(cls, full: "DecodeModel", tp: TPGroup) -> "DecodeModel":
        """The tensor-parallel model whose ranks together hold exactly `full` (an unsharded model on this device)."""
        self = cls.__new__(cls)
        for k in ("cfg", "batch", "ctx", "block_tables", "seq_lens", "slot_mapping", "x", "norm", "positions", "cos_sin_cache"):
            setattr(self, k, getattr(full, k))
        self.tp = tp
        if tp.world_size > 1 and tp.custom is None and full.x.is_cuda:
            tp.enable_custom_all_reduce(full.x.device)
        self.layers = [DecodeLayer.shard_of(l, tp) for l in full.layers]
        self.out = torch.empty_like(full.x)
        self.graph = None
        return self

    def _forward(self) -> None:
        x, residual = self.x.clone(), None  # the layers work in place on the residual stream
        for layer in self.layers:
            x, residual = layer.forward(x, residual, self.positions, self.cos_sin_cache, self.slot_mapping, self.block_tables,
                                        self.seq_lens, self.ctx)
        self.tp.all_reduce_add_rms_norm(x, residual, self.norm, self.cfg.rms_norm_eps)
        self.out.copy_(x)

    def capture(self) -> None:
        """Capture one step (kernels and all_reduces) in a CUDA graph after a warm-up on a side stream."""
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                self._forward()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._forward()

    def step(self) -> torch.Tensor:
        if self.graph is not None:
            self.graph.replay()
        else:
            self._forward()
        return self.out

    def launches_per_step(self) -> int:
        """Kernels of this repo per step: per layer 4 GEMMs, 2 norms, rotary, cache write, attention + its split-K
        reduce, silu_and_mul; one final norm."""
        return self.cfg.num_layers * 11 + 1

    def kv_bytes_per_step(self) -> int:
        l0 = self.layers[0]
        return 2 * self.batch * self.ctx * l0.kvh * self.cfg.head_size * l0.kv_cache.element_size() * self.cfg.num_layers

    def weight_bytes(self) -> int:
        tot = 0
        for l in self.layers:
            for lin in (l.qkv, l.o, l.gate_up, l.down):
                tot += lin.w.q_weight.numel() * 4 + lin.w.scales.numel() * 2
        return tot
