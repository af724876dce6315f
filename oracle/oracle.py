"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path
(aphrodite-fork_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may use it, and only as the checker or the timed CPU baseline.

ctypes front-end over oracle/liboracle.so (oracle_kernels.cpp) taking torch CPU tensors, plus a
loader for the unmodified reference CPU backend in oracle/_ref (built by oracle/Makefile).
Parity status: pinned (tests/test_oracle_pinned.py).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_KV = {"auto": 0, "fp8": 1, "fp8_e4m3": 1, "fp8_e5m2": 2}
_lib = None


def build(force: bool = False) -> None:
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(os.path.join(_HERE, "liboracle.so")):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    if os.path.exists("/root/reference/kernels/cpu/attention.cpp"):
        subprocess.check_call(["make", "-C", _HERE, "-j4", "ref"], stdout=subprocess.DEVNULL)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _lib = ctypes.CDLL(path)
    return _lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def kv_code(kv_cache_dtype: str) -> int:
    return _KV[kv_cache_dtype]


def fp8_encode(src: torch.Tensor, scale: float, kv_cache_dtype: str) -> torch.Tensor:
    src = src.contiguous()
    dst = torch.empty(src.shape, dtype=torch.uint8)
    lib().oracle_fp8_encode(_p(src), _DT[src.dtype], _p(dst), ctypes.c_int64(src.numel()),
                            ctypes.c_float(scale), _KV[kv_cache_dtype])
    return dst


def fp8_decode(src: torch.Tensor, dtype: torch.dtype, scale: float, kv_cache_dtype: str) -> torch.Tensor:
    src = src.contiguous()
    dst = torch.empty(src.shape, dtype=dtype)
    lib().oracle_fp8_decode(_p(src), _p(dst), _DT[dtype], ctypes.c_int64(src.numel()),
                            ctypes.c_float(scale), _KV[kv_cache_dtype])
    return dst


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale):
    """In-place on key_cache / value_cache (CPU tensors), same argument meaning as
    aphrodite/_custom_ops.py:1444-1456."""
    assert key.stride(-1) == 1 and key.stride(1) == key.shape[2]
    nb, kvh, dx, bs, x = key_cache.shape
    lib().oracle_reshape_and_cache(
        _p(key), _p(value), _p(key_cache), _p(value_cache), _p(slot_mapping.contiguous()),
        ctypes.c_int64(slot_mapping.numel()), ctypes.c_int64(key.stride(0)), ctypes.c_int64(value.stride(0)),
        kvh, key.shape[2], bs, x, _DT[key.dtype], _KV[kv_cache_dtype],
        ctypes.c_float(float(k_scale)), ctypes.c_float(float(v_scale)))


def paged_attention(query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                    alibi_slopes=None, kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0, tp_rank=0,
                    blocksparse_local_blocks=0, blocksparse_vert_stride=0, blocksparse_block_size=64,
                    blocksparse_head_sliding_step=0):
    """Returns (out_f64 [S,H,D] double, out rounded to query.dtype)."""
    S, H, D = query.shape
    assert query.stride(2) == 1 and query.stride(1) == D
    x = key_cache.shape[4]
    out64 = torch.empty(S, H, D, dtype=torch.float64)
    out = torch.empty(S, H, D, dtype=query.dtype)
    bt = block_tables.to(torch.int32).contiguous()
    sl = seq_lens.to(torch.int32).contiguous()
    al = alibi_slopes.float().contiguous() if alibi_slopes is not None else None
    lib().oracle_paged_attention(
        _p(out64), _p(out), _p(query), _p(key_cache), _p(value_cache), S, H, num_kv_heads, D, block_size, x,
        ctypes.c_int64(query.stride(0)), ctypes.c_int64(key_cache.stride(0)), ctypes.c_int64(key_cache.stride(1)),
        ctypes.c_float(scale), _p(bt), bt.shape[1], _p(sl), _p(al), _DT[query.dtype], _KV[kv_cache_dtype],
        ctypes.c_float(float(k_scale)), ctypes.c_float(float(v_scale)), tp_rank, blocksparse_local_blocks,
        blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step)
    return out64, out


def gemm(a: torch.Tensor, w_ref: torch.Tensor) -> torch.Tensor:
    """a [M,K] @ w_ref [K,N] in double (w_ref = dequantised weights, same dtype as a)."""
    M, K = a.shape
    N = w_ref.shape[1]
    assert a.stride(1) == 1 and a.dtype == w_ref.dtype
    w_ref = w_ref.contiguous()
    c = torch.empty(M, N, dtype=torch.float64)
    lib().oracle_gemm(_p(c), _p(a), _p(w_ref), ctypes.c_int64(M), ctypes.c_int64(N), ctypes.c_int64(K),
                      ctypes.c_int64(a.stride(0)), _DT[a.dtype])
    return c


_refcpu = None


def refcpu():
    """torch.ops._refcpu = the UNMODIFIED reference CPU backend (oracle/_ref), or None."""
    global _refcpu
    if _refcpu is None:
        flags = open("/proc/cpuinfo").read()
        isa = "avx512bf16" if "avx512_bf16" in flags else ("avx512" if "avx512f" in flags else "avx2")
        path = os.path.join(_HERE, "_ref", f"refcpu_{isa}.so")
        if not os.path.exists(path):
            return None
        torch.ops.load_library(path)
        _refcpu = torch.ops._refcpu
    return _refcpu


_refcuda = None


def refcuda():
    """torch.ops._refcuda = the UNMODIFIED reference CUDA kernels (oracle/_ref/refcuda.so, built by
    `make -C oracle refcuda` from the sources under /root/reference), or None when not built or
    when there is no GPU. Same-hardware parity target and "kernel to beat"; checker only."""
    global _refcuda
    if _refcuda is None:
        path = os.path.join(_HERE, "_ref", "refcuda.so")
        if not os.path.exists(path):  # attention + cache only (linked before gptq_marlin.cu finished compiling)
            path = os.path.join(_HERE, "_ref", "refcuda_partial.so")
        if not os.path.exists(path) or not torch.cuda.is_available():
            return None
        torch.ops.load_library(path)
        _refcuda = torch.ops._refcuda
    return _refcuda


# ---- the rest of kernels/cache_kernels.cu: index restatements in torch on the CPU ----
def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale):
    """In-place on NHD caches [num_blocks, block_size, num_heads, head_size] (CPU tensors). Follows
    cache_kernels.cu:264-306: slot < 0 skipped, block = slot / block_size, offset = slot % block_size,
    cache[block, offset, h, e] = cvt(key[t, h, e]); the FP8 conversion is the write codec (fp8_encode)."""
    bs = key_cache.shape[1]
    if kv_cache_dtype != "auto":
        key = fp8_encode(key.contiguous(), float(k_scale), kv_cache_dtype)
        value = fp8_encode(value.contiguous(), float(v_scale), kv_cache_dtype)
    for t in range(slot_mapping.numel()):  # num_tokens = slot_mapping.size(0) (cache_kernels.cu:423-431)
        slot = int(slot_mapping[t])
        if slot < 0:
            continue
        key_cache[slot // bs, slot % bs] = key[t]
        value_cache[slot // bs, slot % bs] = value[t]


def copy_blocks(key_caches, value_caches, block_mapping):
    """cache_kernels.cu:71-93: for every layer and (src, dst) pair, in list order, dst block = src block."""
    for kc, vc in zip(key_caches, value_caches):
        for src, dst in block_mapping.tolist():
            kc[dst] = kc[src].clone()
            vc[dst] = vc[src].clone()


def swap_blocks(src, dst, block_mapping):
    """cache_kernels.cu:25-67: dst[dst_block] = src[src_block] for every pair (whole blocks, padding included)."""
    for s, d in block_mapping.tolist():
        dst[d] = src[s]


def concat_and_cache_mla(kv_c, k_pe, kv_cache, slot_mapping, kv_cache_dtype, scale):
    """kernels/cache_kernels.cu:308-348: kv_cache[slot] = [kv_c | k_pe], FP8-encoded when the cache is FP8. In place."""
    bs = kv_cache.shape[1]
    row = torch.cat((kv_c, k_pe), dim=-1)
    if kv_cache_dtype != "auto":
        row = fp8_encode(row, float(scale), kv_cache_dtype)
    for t in range(slot_mapping.shape[0]):
        s = int(slot_mapping[t])
        if s >= 0:
            kv_cache[s // bs, s % bs] = row[t].view(kv_cache.dtype) if row.dtype != kv_cache.dtype else row[t]


def copy_blocks_mla(kv_caches, block_mapping):
    """kernels/cache_kernels.cu:100-117: sequential (src, dst) block copies inside every layer's joint cache."""
    for c in kv_caches:
        for s, d in block_mapping.tolist():
            c[d] = c[s].clone()


def gather_cache(src_cache, dst, block_table, cu_seq_lens, batch_size, seq_starts=None):
    """kernels/cache_kernels.cu:582-660: the paged rows of every sequence, contiguous in dst. In place."""
    bs = src_cache.shape[1]
    for b in range(batch_size):
        start, end = int(cu_seq_lens[b]), int(cu_seq_lens[b + 1])
        off = int(seq_starts[b]) // bs if seq_starts is not None else 0
        for t in range(end - start):
            dst[start + t] = src_cache[int(block_table[b, off + t // bs]), t % bs]
