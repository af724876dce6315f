"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

CPU restatement of the Marlin W4A16/W8A16 wire format and of the reference quantiser.
Parity status: pinned - tests/test_oracle_pinned.py compares every function here against
golden vectors produced by importing the reference's own Python code
(tests/golden/make_golden.py -> tests/golden/marlin_*.npz).

What each function follows (paths relative to /root/reference):
  weight_perm / pack_weights   aphrodite/quantization/utils/marlin_utils_test.py:30-92
  permute_scales               aphrodite/quantization/utils/marlin_utils.py:204-225
  pack_zero_points             aphrodite/quantization/utils/marlin_utils.py:246-265
  quantize                     aphrodite/quantization/utils/quant_utils.py:242-329
  scalar type ids              aphrodite/scalar_type.py:133-159
  gptq_pack / awq_pack         aphrodite/quantization/utils/quant_utils.py:461-569
  gptq/awq_marlin_repack       kernels/quantization/gptq_marlin/gptq_marlin_repack.cu:236-322,
                               awq_marlin_repack.cu:189-268; contract from tests/kernels/test_marlin_gemm.py:66-176

The layout is restated as closed-form index arithmetic (not as the reference's list
building), and the inverse maps (unpack_*) are derived here: they are what the GPU kernel's
understanding of the format is tested against.
"""
from __future__ import annotations

import numpy as np
import torch

TILE = 16

# ScalarType ids (exponent, mantissa, signed, bias, finite_values_only, nan_repr packed
# LSB-first as 8/8/1/32/1/8 bits; scalar_type.py:133-159).


def scalar_type_id(exponent: int, mantissa: int, signed: bool, bias: int) -> int:
    val, off = 0, 0
    for v, w in ((exponent, 8), (mantissa, 8), (int(signed), 1), (bias & 0xFFFFFFFF, 32), (0, 1), (1, 8)):
        val |= (int(v) & ((1 << w) - 1)) << off
        off += w
    return val


UINT4 = scalar_type_id(0, 4, False, 0)
UINT4B8 = scalar_type_id(0, 4, False, 8)
UINT8 = scalar_type_id(0, 8, False, 0)
UINT8B128 = scalar_type_id(0, 8, False, 128)


def _interleave(num_bits: int) -> np.ndarray:
    if num_bits == 4:
        return np.array([0, 2, 4, 6, 1, 3, 5, 7])
    if num_bits == 8:
        return np.array([0, 2, 1, 3])
    raise ValueError(f"num_bits must be 4 or 8, got {num_bits}")


def weight_perm(num_bits: int) -> np.ndarray:
    """perm[p] = source index inside a 16(k) x 64(n) block stored as 4 row-major 16x16 tiles.

    Output position p = lane*32 + tile*8 + e: lane in [0,32) owns, for each of the 4 tiles,
    the 8 values an mma.m16n8k16 B fragment pair needs: rows {2(lane%4), +1, +8, +9} x
    cols {lane/4, lane/4+8}.
    """
    p = np.arange(1024)
    lane, tile, e = p // 32, (p % 32) // 8, p % 8
    block, ridx = e // 4, e % 4
    row = 2 * (lane % 4) + (ridx % 2) + 8 * (ridx // 2)
    col = lane // 4 + 8 * block
    src = 256 * tile + 16 * row + col
    il = _interleave(num_bits)
    return src.reshape(-1, len(il))[:, il].ravel()


def _tile_weights(q: np.ndarray) -> np.ndarray:
    k, n = q.shape
    assert k % TILE == 0 and n % 64 == 0, (k, n)
    return q.reshape(k // TILE, TILE, n // TILE, TILE).transpose(0, 2, 1, 3).reshape(k // TILE, n * TILE)


def pack_weights(q: np.ndarray, num_bits: int) -> np.ndarray:
    """q [K,N] unsigned ints (bias already added) -> b_q_weight [K/16, N*16/pack] int32."""
    k, n = q.shape
    t = _tile_weights(np.asarray(q, dtype=np.uint32))
    t = t.reshape(-1, 1024)[:, weight_perm(num_bits)].reshape(t.shape)
    pack = 32 // num_bits
    out = np.zeros((t.shape[0], t.shape[1] // pack), dtype=np.uint32)
    for i in range(pack):
        out |= t[:, i::pack] << np.uint32(num_bits * i)
    return out.astype(np.int32)


def unpack_weights(b: np.ndarray, size_k: int, size_n: int, num_bits: int) -> np.ndarray:
    """Inverse of pack_weights: b_q_weight -> q [K,N] (uint32)."""
    pack = 32 // num_bits
    u = np.asarray(b).astype(np.uint32)
    assert u.shape == (size_k // TILE, size_n * TILE // pack), u.shape
    t = np.zeros((u.shape[0], u.shape[1] * pack), dtype=np.uint32)
    mask = np.uint32((1 << num_bits) - 1)
    for i in range(pack):
        t[:, i::pack] = (u >> np.uint32(num_bits * i)) & mask
    inv = np.argsort(weight_perm(num_bits))
    t = t.reshape(-1, 1024)[:, inv].reshape(t.shape)
    return t.reshape(size_k // TILE, size_n // TILE, TILE, TILE).transpose(0, 2, 1, 3).reshape(size_k, size_n)


def gptq_pack(q: np.ndarray, num_bits: int) -> np.ndarray:
    """GPTQ checkpoint format: q [K,N] -> [K/pack, N] int32, row pack*r+i in bits [num_bits*i, ...)."""
    pack = 32 // num_bits
    u = np.asarray(q, dtype=np.uint32)
    out = np.zeros((u.shape[0] // pack, u.shape[1]), dtype=np.uint32)
    for i in range(pack):
        out |= u[i::pack, :] << np.uint32(num_bits * i)
    return out.astype(np.int32)


def gptq_unpack(w: np.ndarray, num_bits: int) -> np.ndarray:
    pack = 32 // num_bits
    u = np.asarray(w).astype(np.uint32)
    q = np.zeros((u.shape[0] * pack, u.shape[1]), dtype=np.uint32)
    for i in range(pack):
        q[i::pack, :] = (u >> np.uint32(num_bits * i)) & np.uint32((1 << num_bits) - 1)
    return q


def awq_pack(q: np.ndarray, num_bits: int) -> np.ndarray:
    """AWQ checkpoint format: q [K,N] -> [K, N/pack] int32; bit position i of a word holds column
    pack*(n/pack) + interleave[i]."""
    pack = 32 // num_bits
    u = np.asarray(q, dtype=np.uint32)
    k, n = u.shape
    u = u.reshape(k, n // pack, pack)[:, :, _interleave(num_bits)]
    out = np.zeros((k, n // pack), dtype=np.uint32)
    for i in range(pack):
        out |= u[:, :, i] << np.uint32(num_bits * i)
    return out.astype(np.int32)


def awq_unpack(w: np.ndarray, num_bits: int) -> np.ndarray:
    pack = 32 // num_bits
    u = np.asarray(w).astype(np.uint32)
    k, nw = u.shape
    q = np.zeros((k, nw, pack), dtype=np.uint32)
    il = _interleave(num_bits)
    for i in range(pack):
        q[:, :, il[i]] = (u >> np.uint32(num_bits * i)) & np.uint32((1 << num_bits) - 1)
    return q.reshape(k, nw * pack)


def gptq_marlin_repack(w: np.ndarray, perm, size_k: int, size_n: int, num_bits: int) -> np.ndarray:
    """What the repack kernel must produce: the Marlin packing of the unpacked matrix, rows taken in `perm`
    order when act-order sort indices are given (test_marlin_gemm.py:97-118)."""
    q = gptq_unpack(w, num_bits)
    assert q.shape == (size_k, size_n), q.shape
    if perm is not None and len(perm) > 0:
        q = q[np.asarray(perm, dtype=np.int64), :]
    return pack_weights(q, num_bits)


def awq_marlin_repack(w: np.ndarray, size_k: int, size_n: int, num_bits: int) -> np.ndarray:
    q = awq_unpack(w, num_bits)
    assert q.shape == (size_k, size_n), q.shape
    return pack_weights(q, num_bits)


def _scale_perms():
    i = np.arange(64)
    grouped = (i % 8) * 8 + i // 8  # transpose of the 8x8 index grid
    j = np.array([0, 1, 8, 9, 16, 17, 24, 25])
    single = (2 * np.arange(4)[:, None] + j[None, :]).ravel()
    return grouped, single


def permute_scales(s: np.ndarray, size_k: int, size_n: int, group_size: int) -> np.ndarray:
    grouped, single = _scale_perms()
    if group_size < size_k and group_size != -1:
        out = s.reshape(-1, 64)[:, grouped]
    else:
        out = s.reshape(-1, 32)[:, single]
    return np.ascontiguousarray(out.reshape(-1, size_n))


def unpermute_scales(s: np.ndarray, size_k: int, size_n: int, group_size: int) -> np.ndarray:
    grouped, single = _scale_perms()
    if group_size < size_k and group_size != -1:
        out = s.reshape(-1, 64)[:, np.argsort(grouped)]
    else:
        out = s.reshape(-1, 32)[:, np.argsort(single)]
    return np.ascontiguousarray(out.reshape(-1, size_n))


def pack_zero_points(zp: np.ndarray, num_groups: int, size_n: int, num_bits: int) -> np.ndarray:
    """zp [groups, N] ints -> Marlin b_zeros [groups, N/pack] int32."""
    grouped, _ = _scale_perms()
    il = _interleave(num_bits)
    z = np.asarray(zp, dtype=np.uint32).reshape(-1, 64)[:, grouped]
    z = z.reshape(-1, len(il))[:, il].reshape(num_groups, size_n)
    pack = 32 // num_bits
    out = np.zeros((num_groups, size_n // pack), dtype=np.uint32)
    for i in range(pack):
        out |= z[:, i::pack] << np.uint32(num_bits * i)
    return out.astype(np.int32)


def unpack_zero_points(bz: np.ndarray, num_groups: int, size_n: int, num_bits: int) -> np.ndarray:
    pack = 32 // num_bits
    u = np.asarray(bz).astype(np.uint32)
    z = np.zeros((num_groups, size_n), dtype=np.uint32)
    mask = np.uint32((1 << num_bits) - 1)
    for i in range(pack):
        z[:, i::pack] = (u >> np.uint32(num_bits * i)) & mask
    grouped, _ = _scale_perms()
    il = _interleave(num_bits)
    z = z.reshape(-1, len(il))[:, np.argsort(il)]
    z = z.reshape(-1, 64)[:, np.argsort(grouped)]
    return z.reshape(num_groups, size_n)


def quantize(w: torch.Tensor, num_bits: int, group_size: int, zero_points: bool = False):
    """Reference quantiser, arithmetic in w's own dtype like the reference.

    Returns (w_ref [K,N] dtype of w, q [K,N] int32 with bias added, s [groups,N], zp or None).
    Symmetric types are uint4b8 / uint8b128 (bias 8 / 128); zero-point types uint4 / uint8.
    """
    size_k, size_n = w.shape
    dtype = w.dtype
    if group_size == -1:
        group_size = size_k
    g = w
    if group_size < size_k:
        g = w.reshape(-1, group_size, size_n).permute(1, 0, 2).reshape(group_size, -1)
    max_val = torch.max(g, 0, keepdim=True).values
    min_val = torch.min(g, 0, keepdim=True).values
    if zero_points:
        max_q, min_q, bias = (1 << num_bits) - 1, 0, 0
        s = (max_val - min_val).clamp(min=1e-5) / max_q
        zp = torch.round(torch.abs(min_val / s)).clamp(min_q, max_q).int()
    else:
        bias = 1 << (num_bits - 1)
        max_q, min_q = bias - 1, -bias
        s = torch.max((max_val / max_q).abs(), (min_val / min_q).abs())
        zp = None
    q = torch.round(g / s).int() + (zp if zero_points else 0)
    q = torch.clamp(q, min_q, max_q)
    w_ref = (q - (zp if zero_points else 0)).to(dtype) * s
    q = q + bias
    if group_size < size_k:
        def back(t):
            return t.reshape(group_size, -1, size_n).permute(1, 0, 2).reshape(size_k, size_n).contiguous()
        q, w_ref = back(q), back(w_ref)
    s = s.reshape(-1, size_n).contiguous()
    if zp is not None:
        zp = zp.reshape(-1, size_n).contiguous()
    return w_ref, q, s, zp


def marlin_quantize(w: torch.Tensor, num_bits: int, group_size: int, zero_points: bool = False):
    """w [K,N] float -> (w_ref, b_q_weight int32, b_scales, b_zeros or None) in Marlin layout."""
    size_k, size_n = w.shape
    gs = size_k if group_size == -1 else group_size
    w_ref, q, s, zp = quantize(w, num_bits, gs, zero_points)
    b = torch.from_numpy(pack_weights(q.numpy().astype(np.uint32), num_bits))
    s_np = s.view(torch.int16).numpy() if s.dtype in (torch.bfloat16, torch.float16) else s.numpy()
    sp = torch.from_numpy(permute_scales(s_np, size_k, size_n, gs))
    if s.dtype in (torch.bfloat16, torch.float16):
        sp = sp.view(s.dtype)
    bz = None
    if zp is not None:
        bz = torch.from_numpy(pack_zero_points(zp.numpy(), size_k // gs, size_n, num_bits))
    return w_ref, b, sp, bz


def dequantize(b: torch.Tensor, s: torch.Tensor, bz, size_k: int, size_n: int, num_bits: int,
               group_size: int, has_zp: bool) -> torch.Tensor:
    """Decode Marlin tensors back to w_ref [K,N] in s.dtype: (q - zp_or_bias) * s per group."""
    gs = size_k if group_size == -1 else group_size
    q = torch.from_numpy(unpack_weights(b.numpy(), size_k, size_n, num_bits).astype(np.int32))
    raw = s.view(torch.int16).numpy() if s.dtype in (torch.bfloat16, torch.float16) else s.numpy()
    su = torch.from_numpy(unpermute_scales(raw, size_k, size_n, gs))
    if s.dtype in (torch.bfloat16, torch.float16):
        su = su.view(s.dtype)
    if has_zp:
        z = torch.from_numpy(unpack_zero_points(bz.numpy(), size_k // gs, size_n, num_bits).astype(np.int32))
        sub = z.repeat_interleave(gs, dim=0)
    else:
        sub = 1 << (num_bits - 1)
    return (q - sub).to(s.dtype) * su.repeat_interleave(gs, dim=0)
