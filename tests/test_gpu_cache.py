"""GPU parity: reshape_and_cache / convert_fp8 through the registered torch ops (which call the
C ABI) against the CPU oracle - BIT-EXACT, including the FP8 bytes. Mirrors
/root/reference/tests/kernels/test_cache.py:112-206 and :400-433."""
import os

import numpy as np
import pytest
import torch

import synth
from oracle import oracle

pytestmark = pytest.mark.gpu
DT = {0: torch.float32, 1: torch.float16, 2: torch.bfloat16}


def _ops():
    import aphrodite_fork_b200 as A
    return A.ops


def _prefill_slots(T, bs, nb, seed):
    """Slot mapping of a prefill: runs of consecutive slots (whole blocks in a shuffled order), with a ragged first
    run, one padding token and one run broken in the middle, so that both the transposed group path and the
    per-element path of the kernel are taken inside one call."""
    blocks = synth.permutation(nb, seed)
    slots = []
    start = 3  # the first sequence starts in the middle of a block
    for b in blocks:
        slots.extend(int(b) * bs + o for o in range(start, bs))
        start = 0
        if len(slots) >= T:
            break
    slots = torch.tensor(slots[:T], dtype=torch.int64)
    if T > 40:
        slots[17] = -1
        slots[33], slots[34] = slots[34].clone(), slots[33].clone()
    return slots


def _run_case(T, KVH, D, bs, nb, dtype, kv, k_scale, v_scale, seed, neg_frac=0.05, strided=True, prefill=False):
    ops = _ops()
    cache_dt = torch.uint8 if kv != "auto" else dtype
    x = 16 // torch.tensor([], dtype=cache_dt).element_size()
    if strided:
        qkv = synth.tensor((T, 3, KVH, D), seed, dtype, 3.0)
        key, value = qkv[:, 1], qkv[:, 2]
    else:
        key = synth.tensor((T, KVH, D), seed, dtype, 3.0)
        value = synth.tensor((T, KVH, D), seed + 7, dtype, 3.0)
    slots = torch.from_numpy(synth.permutation(nb * bs, seed + 1)[:T].copy()).to(torch.int64)
    if prefill:
        slots = _prefill_slots(T, bs, nb, seed + 1)
    nneg = 0 if prefill else int(T * neg_frac)
    if nneg:
        slots[torch.from_numpy(synth.permutation(T, seed + 2)[:nneg].copy())] = -1
    if kv == "auto":
        kc = synth.tensor((nb, KVH, D // x, bs, x), seed + 3, dtype, 1.0)
        vc = synth.tensor((nb, KVH, D, bs), seed + 4, dtype, 1.0)
    else:
        kc = torch.from_numpy((synth.hash_uniform(nb * KVH * D * bs, seed + 3, 0, 255)).astype(np.uint8)).reshape(nb, KVH, D // x, bs, x)
        vc = torch.from_numpy((synth.hash_uniform(nb * KVH * D * bs, seed + 4, 0, 255)).astype(np.uint8)).reshape(nb, KVH, D, bs)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    oracle.reshape_and_cache(key, value, kc_ref, vc_ref, slots, kv, k_scale, v_scale)
    dev = "cuda"
    if strided:
        qkv_d = qkv.to(dev)
        key_d, value_d = qkv_d[:, 1], qkv_d[:, 2]
    else:
        key_d, value_d = key.to(dev), value.to(dev)
    kc_d, vc_d = kc.to(dev), vc.to(dev)
    ks = torch.tensor(k_scale, dtype=torch.float32, device=dev)
    vs = torch.tensor(v_scale, dtype=torch.float32, device=dev)
    ops.reshape_and_cache(key_d, value_d, kc_d, vc_d, slots.to(dev), kv, ks, vs)
    torch.cuda.synchronize()
    iv = {1: torch.uint8, 2: torch.int16, 4: torch.int32}[kc.element_size()]
    assert torch.equal(kc_d.cpu().view(iv), kc_ref.view(iv)), "key cache differs"
    assert torch.equal(vc_d.cpu().view(iv), vc_ref.view(iv)), "value cache differs"


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("D", [64, 80, 128, 256])
@pytest.mark.parametrize("bs", [8, 16, 32])
def test_reshape_and_cache_sweep(kv, dtype, D, bs):
    _run_case(42, 8, D, bs, 64, dtype, kv, 1.0, 1.0, seed=D + bs)


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("bs", [8, 16, 32])
def test_reshape_and_cache_prefill_runs(kv, dtype, bs):
    # consecutive slots (the transposed 8-token group path) mixed with ragged / broken / padded groups
    _run_case(203, 4, 128, bs, 203 // bs + 4, dtype, kv, 0.5 if kv != "auto" else 1.0, 0.25 if kv != "auto" else 1.0,
              seed=bs + 11, prefill=True)
    _run_case(64, 8, 64, bs, 64 // bs + 2, dtype, kv, 1.0, 1.0, seed=bs + 12, prefill=True, strided=False)


@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
def test_reshape_and_cache_scaled_fp8(kv):
    _run_case(64, 8, 128, 16, 128, torch.bfloat16, kv, 0.05, 0.02, seed=5)
    _run_case(64, 8, 128, 16, 128, torch.bfloat16, kv, 3.0, 7.5, seed=6)  # saturates to the max finite value


def test_reshape_and_cache_baseline_shapes():
    # C2 decode (64 tokens, 8 kv heads x 128) and a prefill-sized batch, contiguous and strided
    _run_case(64, 8, 128, 16, 4096, torch.bfloat16, "fp8", 1.0, 1.0, seed=11)
    _run_case(8192, 8, 128, 16, 1024, torch.bfloat16, "fp8", 1.0, 1.0, seed=12, neg_frac=0.01)
    _run_case(8192, 8, 128, 16, 1024, torch.bfloat16, "auto", 1.0, 1.0, seed=13, neg_frac=0.01, strided=False)
    _run_case(8, 12, 64, 16, 256, torch.bfloat16, "auto", 1.0, 1.0, seed=14)  # C1 (OPT-125m)


def test_reshape_and_cache_edge_cases():
    _run_case(1, 1, 32, 16, 4, torch.float16, "auto", 1.0, 1.0, seed=21, neg_frac=0)
    _run_case(3, 2, 64, 16, 4, torch.float16, "fp8", 1.0, 1.0, seed=22, neg_frac=1.0)  # every slot padded
    ops = _ops()
    e = torch.empty(0, 2, 64, dtype=torch.float16, device="cuda")
    kc = torch.zeros(2, 2, 8, 16, 8, dtype=torch.float16, device="cuda")
    vc = torch.zeros(2, 2, 64, 16, dtype=torch.float16, device="cuda")
    one = torch.tensor(1.0, device="cuda")
    ops.reshape_and_cache(e, e, kc, vc, torch.empty(0, dtype=torch.int64, device="cuda"), "auto", one, one)
    with pytest.raises(RuntimeError):
        ops.reshape_and_cache(e, e, kc, vc, torch.empty(0, dtype=torch.int64, device="cuda"), "int8", one, one)


def test_reshape_and_cache_golden_reference_cpu():
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cache_golden.npz"))
    ops = _ops()
    for ci in range(int(g["num_cases"][0])):
        T, KVH, D, bs, nb, dtc = [int(v) for v in g[f"r{ci}_meta"]]
        dt = DT[dtc]
        x = 16 // torch.tensor([], dtype=dt).element_size()
        qkv = synth.tensor((T, 3, KVH, D), 500 + ci, dt, 2.0).cuda()
        slots = torch.from_numpy(g[f"r{ci}_slots"]).cuda()
        kc = synth.tensor((nb, KVH, D // x, bs, x), 700 + ci, dt, 2.0).cuda()
        vc = synth.tensor((nb, KVH, D, bs), 800 + ci, dt, 2.0).cuda()
        one = torch.tensor(1.0, device="cuda")
        ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots, "auto", one, one)
        iv = torch.int16 if dt != torch.float32 else torch.int32
        rk = torch.from_numpy(g[f"r{ci}_kc1"])
        rv = torch.from_numpy(g[f"r{ci}_vc1"])
        assert torch.equal(kc.cpu().view(iv).reshape(-1), rk.view(iv).reshape(-1)), ci
        assert torch.equal(vc.cpu().view(iv).reshape(-1), rv.view(iv).reshape(-1)), ci


@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_convert_fp8_bit_exact_and_roundtrip(kv, dtype):
    ops = _ops()
    lim = 224.0 if kv == "fp8" else 5000.0
    src = synth.tensor((3, 1000, 7), 31, dtype, lim)
    for scale in (1.0, 0.37):
        ref = oracle.fp8_encode(src, scale, kv)
        dst = torch.empty(src.shape, dtype=torch.uint8, device="cuda")
        ops.convert_fp8(dst, src.cuda(), scale, kv)
        assert torch.equal(dst.cpu(), ref)
        back = torch.empty(src.shape, dtype=dtype, device="cuda")
        ops.convert_fp8(back, dst, scale, kv)
        ref_back = oracle.fp8_decode(ref, dtype, scale, kv)
        iv = torch.int16 if dtype != torch.float32 else torch.int32
        assert torch.equal(back.cpu().view(iv), ref_back.view(iv))
        # the reference's own round-trip tolerance (test_cache.py:400-433): atol 1e-3, rtol 0.1
        if kv == "fp8" and scale == 1.0:
            assert torch.allclose(back.cpu().float(), src.float(), atol=1e-3, rtol=0.1)
