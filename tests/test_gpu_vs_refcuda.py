"""GPU parity against the UNMODIFIED reference CUDA kernels on the same inputs, same B200.

oracle/_ref/refcuda.so is /root/reference's own paged_attention_v1/v2.cu, cache_kernels.cu and
gptq_marlin.cu compiled for sm_100a by `make -C oracle refcuda` (checker only; registered under the
private namespace torch.ops._refcuda). north_star: "Outputs must match the reference's own kernels
on the same inputs: bit-exact for the block-table and slot-mapping cache writes, and within a stated
bf16 tolerance for floating point". Tolerances:
  cache writes / convert_fp8      torch.equal on the cache bytes
  attention                       |ours - ref| <= atol + rtol*|ref|, atol 1e-3 (16-bit KV) / 1e-2 (FP8),
                                  rtol 1.6e-2 bf16 / 1e-3 f16; and our error against the fp64 oracle
                                  <= the reference kernel's error + one ulp of the scalar dtype
  gptq_marlin_gemm                max|ours - ref| <= 1e-2 * max|ref| (bf16) / 2e-3 (f16)
Skipped (not failed) when refcuda.so has not been built."""
import pytest
import torch

import synth
from helpers import RTOL, make_attention_case
from oracle import marlin_format as mf
from oracle import oracle

pytestmark = pytest.mark.gpu


def _ref():
    r = oracle.refcuda()
    if r is None:
        pytest.skip("oracle/_ref/refcuda.so not built (make -C oracle refcuda)")
    return r


def _ops():
    import aphrodite_fork_b200 as A
    return A.ops


def ragged(S, max_len, seed):
    lens = [int(v) % max_len + 1 for v in synth.permutation(100003, seed)[:S]]
    lens[-1] = max_len
    return lens


# ---------------------------------------------------------------- cache writes: bit-exact
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(64, 8, 128, 16), (37, 12, 64, 16), (19, 4, 80, 32), (23, 2, 256, 8), (2048, 8, 128, 16)])
def test_reshape_and_cache_equals_reference_cuda(kv, dtype, shape):
    ref, ops = _ref(), _ops()
    T, KVH, D, bs = shape
    dev = "cuda"
    nb = (T + bs - 1) // bs + 3
    esz = 1 if kv != "auto" else torch.tensor([], dtype=dtype).element_size()
    cdt = torch.uint8 if kv != "auto" else dtype
    x = 16 // esz
    qkv = synth.tensor((T, 3, KVH, D), T + D, dtype, 3.0).to(dev)  # strided slices of a fused qkv, as the caller passes
    key, value = qkv[:, 1], qkv[:, 2]
    slots = torch.from_numpy(synth.permutation(nb * bs, T + 1)[:T].copy()).to(torch.int64)
    if T == 2048:  # prefill: consecutive slots
        slots = torch.arange(5, 5 + T, dtype=torch.int64)
    slots[::17] = -1
    slots = slots.to(dev)
    ks = torch.tensor(0.37 if kv != "auto" else 1.0, device=dev)
    vs = torch.tensor(0.011 if kv != "auto" else 1.0, device=dev)
    caches = []
    for fn in (ref.reshape_and_cache, ops.reshape_and_cache):
        kc = torch.full((nb, KVH, D // x, bs, x), 7, dtype=cdt, device=dev)
        vc = torch.full((nb, KVH, D, bs), 9, dtype=cdt, device=dev)
        fn(key, value, kc, vc, slots, kv, ks, vs)
        caches.append((kc, vc))
    torch.cuda.synchronize()
    assert torch.equal(caches[0][0].view(torch.uint8), caches[1][0].view(torch.uint8)), "key cache differs"
    assert torch.equal(caches[0][1].view(torch.uint8), caches[1][1].view(torch.uint8)), "value cache differs"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_convert_fp8_equals_reference_cuda(dtype):
    ref, ops = _ref(), _ops()
    src = synth.tensor((6, 4, 16, 16, 8), 5, dtype, 40.0).cuda()
    src.view(-1)[:4] = torch.tensor([1e6, -1e6, 0.0, 448.0], dtype=dtype)
    a = torch.zeros(src.shape, dtype=torch.uint8, device="cuda")
    b = torch.zeros_like(a)
    ref.convert_fp8(a, src, 0.25, "fp8")
    ops.convert_fp8(b, src, 0.25, "fp8")
    assert torch.equal(a, b)
    back_a, back_b = torch.empty_like(src), torch.empty_like(src)
    ref.convert_fp8(back_a, a, 0.25, "fp8")
    ops.convert_fp8(back_b, a, 0.25, "fp8")
    assert torch.equal(back_a.view(torch.uint8), back_b.view(torch.uint8))


# ---------------------------------------------------------------- attention
def _attention_both(S, H, KVH, D, bs, lens, dtype, kv, seed, version, k_scale=1.0, v_scale=1.0, alibi=False):
    ref, ops = _ref(), _ops()
    dev = "cuda"
    q, kc, vc, bt, sl = make_attention_case(S, H, KVH, D, bs, lens, dtype, kv, seed, k_scale, v_scale)
    truth, _ = oracle.paged_attention(q, kc, vc, KVH, D ** -0.5, bt, sl, bs, kv_cache_dtype=kv, k_scale=k_scale,
                                      v_scale=v_scale,
                                      alibi_slopes=(synth.tensor((H,), seed + 9, torch.float32, 0.5).abs() if alibi else None))
    al = synth.tensor((H,), seed + 9, torch.float32, 0.5).abs().to(dev) if alibi else None
    qd, kcd, vcd, btd, sld = q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), sl.to(dev)
    ks, vs = torch.tensor(k_scale, device=dev), torch.tensor(v_scale, device=dev)
    max_len = max(lens)
    outs = []
    for o in (ref, ops):
        out = torch.empty(S, H, D, dtype=dtype, device=dev)
        if version == 1:
            o.paged_attention_v1(out, qd, kcd, vcd, KVH, D ** -0.5, btd, sld, bs, max_len, al, kv, ks, vs, 0, 0, 0, 64, 0)
        else:
            P = (max_len + 511) // 512
            es = torch.empty(S, H, P, dtype=torch.float32, device=dev)
            ml = torch.empty_like(es)
            tmp = torch.empty(S, H, P, D, dtype=dtype, device=dev)
            o.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, KVH, D ** -0.5, btd, sld, bs, max_len, al, kv, ks, vs,
                                 0, 0, 0, 64, 0)
        outs.append(out.cpu().double())
    r, m = outs
    atol = 1e-3 if kv == "auto" else 1e-2
    d = (m - r).abs()
    assert float((d - (atol + RTOL[dtype] * r.abs())).max()) <= 0, \
        f"ours vs reference CUDA: max abs diff {float(d.max()):.3e}"
    ulp = float(truth.abs().max()) * (2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11)
    err_ref, err_me = float((r - truth).abs().max()), float((m - truth).abs().max())
    assert err_me <= err_ref + ulp, f"our error vs fp64 oracle {err_me:.3e} > reference kernel's {err_ref:.3e} + ulp {ulp:.1e}"
    return float(d.max()), err_ref, err_me


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("heads", [(32, 8), (12, 12), (8, 1)])
def test_paged_attention_matches_reference_cuda(version, kv, dtype, heads):
    H, KVH = heads
    D = 64 if H == 12 else 128
    _attention_both(6, H, KVH, D, 16, ragged(6, 1300, H + D), dtype, kv, seed=H + 3, version=version)


@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
def test_paged_attention_scales_alibi_match_reference_cuda(kv):
    _attention_both(4, 32, 8, 128, 16, ragged(4, 900, 4), torch.bfloat16, kv, seed=42, version=2, k_scale=0.05,
                    v_scale=0.02)
    _attention_both(4, 16, 4, 128, 16, ragged(4, 700, 5), torch.bfloat16, kv, seed=43, version=1, alibi=True)


@pytest.mark.parametrize("D,bs", [(80, 16), (96, 8), (112, 32), (192, 16), (256, 16), (32, 32)])
def test_paged_attention_head_block_sizes_match_reference_cuda(D, bs):
    _attention_both(3, 8, 2, D, bs, ragged(3, 600, D), torch.bfloat16, "auto", seed=D, version=2)
    _attention_both(3, 8, 2, D, bs, ragged(3, 600, D + 1), torch.float16, "fp8", seed=D + 1, version=1)


def test_paged_attention_c2_shape_matches_reference_cuda():
    # BASELINE configs[1] geometry at a reduced batch (8 x 4096 ctx), v2
    _attention_both(8, 32, 8, 128, 16, [4096] * 8, torch.bfloat16, "fp8", seed=77, version=2)


# ---------------------------------------------------------------- W4A16 GEMM
def _gemm_both(M, K, N, gs, dtype, zp=False, seed=1):
    ref, ops = _ref(), _ops()
    dev = "cuda"
    w = synth.tensor((K, N), seed, dtype, 1.0)
    w_ref, b, s, bz = mf.marlin_quantize(w, 4, gs, zp)
    a = synth.tensor((M, K), seed + 100, dtype, 1.0)
    truth = oracle.gemm(a, w_ref)
    empty_i = torch.empty(0, dtype=torch.int32, device=dev)
    zeros = bz.to(dev) if bz is not None else empty_i
    tid = mf.UINT4 if zp else mf.UINT4B8
    outs = []
    for o in (ref, ops):
        ws = torch.zeros(max(N // 64 * 16, 148 * 16), dtype=torch.int32, device=dev)
        try:
            out = o.gptq_marlin_gemm(a.to(dev), b.to(dev), s.to(dev), zeros, empty_i, empty_i, ws, tid, M, N, K, True, zp,
                                     False, True, False)
        except NotImplementedError:
            pytest.skip("refcuda built without gptq_marlin.cu (partial library)")
        torch.cuda.synchronize()
        assert int(ws.abs().sum()) == 0
        outs.append(out.cpu().double())
    r, m = outs
    scale = float(r.abs().max())
    d = float((m - r).abs().max())
    assert d <= (1e-2 if dtype == torch.bfloat16 else 2e-3) * scale, f"ours vs reference CUDA: {d:.3e} (max|ref| {scale:.3e})"
    # "truth" is a @ w_ref with w_ref = (q - zp) * s ROUNDED to the weight dtype (quant_utils.py:298-301); the reference
    # kernel reproduces that rounding bit for bit in its dequant (hmul2), so its error is output rounding only
    # (~1.4e-3 mean-rel in bf16). Our M <= 16 kernel applies s to the fp32 group sum without rounding w_ref, which
    # differs from this truth by one extra bf16 rounding per weight (measured 2.2e-3). Both are far inside the
    # reference test's own bar (mean-rel < 0.04, test_marlin_gemm.py:254); assert 2x the reference's error + 1e-3.
    e_ref = float((r - truth).abs().mean() / truth.abs().mean())
    e_me = float((m - truth).abs().mean() / truth.abs().mean())
    assert e_me <= 2.0 * e_ref + 1e-3, f"mean-rel error vs truth: ours {e_me:.3e}, reference {e_ref:.3e}"
    return d / scale, e_ref, e_me


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("gs", [-1, 128])
@pytest.mark.parametrize("M", [1, 7, 16, 33, 64, 256])
def test_gptq_marlin_gemm_matches_reference_cuda(dtype, gs, M):
    _gemm_both(M, 1024, 2048, gs, dtype, seed=M)


@pytest.mark.parametrize("M", [3, 64])
@pytest.mark.parametrize("gs", [32, 64, 128])
def test_awq_marlin_gemm_matches_reference_cuda(M, gs):
    _gemm_both(M, 1024, 1024, gs, torch.float16, zp=True, seed=M + gs)
    _gemm_both(M, 512, 1536, gs, torch.bfloat16, zp=True, seed=M + gs + 1)


@pytest.mark.parametrize("M", [1, 16, 256])
def test_gptq_marlin_gemm_c3_shape_matches_reference_cuda(M):
    _gemm_both(M, 4096, 6144, 128, torch.bfloat16, seed=9)
