"""Pins the oracle (oracle/) against the reference: golden vectors generated from the reference's
own Python code and from its unmodified CPU backend (tests/golden/make_golden.py), and - when
oracle/_ref is present - against the reference CPU backend run live."""
import os

import numpy as np
import pytest
import torch

import synth
from oracle import marlin_format as mf
from oracle import oracle

G = os.path.join(os.path.dirname(__file__), "golden")
DT = {0: torch.float32, 1: torch.float16, 2: torch.bfloat16}


def _t(a, dt):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.view(dt) if dt in (torch.float16, torch.bfloat16) else t


def test_scalar_type_ids_and_perms():
    g = np.load(os.path.join(G, "marlin_golden.npz"))
    assert mf.UINT4 == int(g["id_uint4"][0]) and mf.UINT4B8 == int(g["id_uint4b8"][0])
    assert mf.UINT8 == int(g["id_uint8"][0]) and mf.UINT8B128 == int(g["id_uint8b128"][0])
    assert mf.UINT4B8 == 1125899907892224
    assert np.array_equal(mf.weight_perm(4), g["perm4"])
    assert np.array_equal(mf.weight_perm(8), g["perm8"])
    sp, sps = mf._scale_perms()
    assert np.array_equal(sp, g["scale_perm"]) and np.array_equal(sps, g["scale_perm_single"])


def test_marlin_quantize_matches_reference_python():
    g = np.load(os.path.join(G, "marlin_golden.npz"))
    for ci in range(int(g["num_cases"][0])):
        k, n, bits, gs, zp, dtc = [int(v) for v in g[f"c{ci}_meta"]]
        dt = DT[dtc]
        w = _t(g[f"c{ci}_w"], dt)
        w_ref, b, s, bz = mf.marlin_quantize(w, bits, gs, bool(zp))
        assert torch.equal(w_ref.view(torch.int16), _t(g[f"c{ci}_w_ref"], dt).view(torch.int16)), ci
        assert np.array_equal(b.numpy(), g[f"c{ci}_b_q_weight"]), ci
        assert np.array_equal(s.view(torch.int16).numpy(), g[f"c{ci}_b_scales"]), ci
        if zp:
            assert np.array_equal(bz.numpy(), g[f"c{ci}_b_zeros"]), ci
        # the inverse maps decode the reference's tensors back to the reference's w_ref
        back = mf.dequantize(torch.from_numpy(g[f"c{ci}_b_q_weight"]), _t(g[f"c{ci}_b_scales"], dt),
                             torch.from_numpy(g[f"c{ci}_b_zeros"]) if zp else None, k, n, bits, gs, bool(zp))
        assert torch.equal(back.view(torch.int16), _t(g[f"c{ci}_w_ref"], dt).view(torch.int16)), ci


def _attn_case(g, ci):
    S, H, KVH, D, max_len, dtc, alibi, ver, nb = [int(v) for v in g[f"a{ci}_meta"]]
    dt = DT[dtc]
    x = 16 // torch.tensor([], dtype=dt).element_size()
    sc = float(D) ** -0.5
    q = synth.tensor((S, H, D), 200 + ci, dt, sc)
    kc = synth.tensor((nb, KVH, D // x, 16, x), 300 + ci, dt, sc)
    vc = synth.tensor((nb, KVH, D, 16), 400 + ci, dt, sc)
    sl = torch.from_numpy(g[f"a{ci}_seq_lens"])
    bt = torch.from_numpy(g[f"a{ci}_block_tables"])
    al = torch.from_numpy(g[f"a{ci}_alibi"]) if alibi else None
    ref = _t(g[f"a{ci}_out"], dt)
    return dict(S=S, H=H, KVH=KVH, D=D, max_len=max_len, dt=dt, ver=ver, q=q, kc=kc, vc=vc, sl=sl, bt=bt,
                al=al, ref=ref, sc=sc)


def test_attention_oracle_vs_reference_cpu_golden():
    g = np.load(os.path.join(G, "attention_golden.npz"))
    for ci in range(int(g["num_cases"][0])):
        c = _attn_case(g, ci)
        o64, o = oracle.paged_attention(c["q"], c["kc"], c["vc"], c["KVH"], c["sc"], c["bt"], c["sl"], 16,
                                        alibi_slopes=c["al"])
        ref = c["ref"].double()
        # the reference CPU backend accumulates in fp32 and rounds once to the scalar dtype;
        # reference test tolerance: atol 1e-3 (tests/kernels/test_attention.py:323-326) + dtype rtol
        rtol = {torch.float32: 1.3e-6, torch.float16: 1e-3, torch.bfloat16: 1.6e-2}[c["dt"]]
        err = (o64 - ref).abs()
        assert bool((err <= 1e-3 + rtol * ref.abs()).all()), (ci, float(err.max()))


def test_cache_oracle_vs_reference_cpu_golden():
    g = np.load(os.path.join(G, "cache_golden.npz"))
    for ci in range(int(g["num_cases"][0])):
        T, KVH, D, bs, nb, dtc = [int(v) for v in g[f"r{ci}_meta"]]
        dt = DT[dtc]
        x = 16 // torch.tensor([], dtype=dt).element_size()
        qkv = synth.tensor((T, 3, KVH, D), 500 + ci, dt, 2.0)
        slots = torch.from_numpy(g[f"r{ci}_slots"])
        kc = synth.tensor((nb, KVH, D // x, bs, x), 700 + ci, dt, 2.0)
        vc = synth.tensor((nb, KVH, D, bs), 800 + ci, dt, 2.0)
        oracle.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots, "auto", 1.0, 1.0)
        i = torch.int16 if dt != torch.float32 else torch.int32
        assert torch.equal(kc.view(i), _t(g[f"r{ci}_kc1"], dt).view(i)), ci
        assert torch.equal(vc.view(i), _t(g[f"r{ci}_vc1"], dt).view(i)), ci


@pytest.mark.skipif(oracle.refcpu() is None, reason="oracle/_ref not built")
def test_oracle_vs_live_reference_cpu():
    ref = oracle.refcpu()
    S, H, KVH, D, L = 3, 8, 2, 128, 333
    dt = torch.bfloat16
    sc = D ** -0.5
    nb = S * 21 + 1
    q = synth.tensor((S, H, D), 1, dt, sc)
    kc = synth.tensor((nb, KVH, D // 8, 16, 8), 2, dt, sc)
    vc = synth.tensor((nb, KVH, D, 16), 3, dt, sc)
    bt = torch.from_numpy(synth.permutation(nb, 4)[: S * 21].reshape(S, 21)).to(torch.int32)
    sl = torch.tensor([17, 200, L], dtype=torch.int32)
    o = torch.empty_like(q)
    one = torch.tensor(1.0)
    ref.paged_attention_v1(o, q, kc, vc, KVH, sc, bt, sl, 16, L, None, "auto", one, one, 0, 0, 0, 64, 0)
    o64, _ = oracle.paged_attention(q, kc, vc, KVH, sc, bt, sl, 16)
    assert float((o64 - o.double()).abs().max()) < 1e-3


def test_fp8_codec_known_answers():
    # E4M3: 448 is the max finite; saturate-to-finite; 1.0 = 0x38; E5M2: 1.0 = 0x3C, max 57344
    x = torch.tensor([0.0, 1.0, -1.0, 448.0, 1e6, -1e6, 2.0 ** -9, 0.3], dtype=torch.float32)
    e4 = oracle.fp8_encode(x, 1.0, "fp8_e4m3")
    assert e4.tolist()[:7] == [0x00, 0x38, 0xB8, 0x7E, 0x7E, 0xFE, 0x01]
    e5 = oracle.fp8_encode(x, 1.0, "fp8_e5m2")
    assert e5.tolist()[:3] == [0x00, 0x3C, 0xBC] and e5.tolist()[4] == 0x7B
    # every byte decodes and re-encodes to itself (idempotence; NaN codes excluded)
    allb = torch.arange(256, dtype=torch.uint8)
    for kv, nan in (("fp8_e4m3", {0x7F, 0xFF}), ("fp8_e5m2", set(range(0x7C, 0x80)) | set(range(0xFC, 0x100)))):
        dec = oracle.fp8_decode(allb, torch.float32, 1.0, kv)
        re = oracle.fp8_encode(dec, 1.0, kv)
        ok = [int(re[i]) == i for i in range(256) if i not in nan]
        assert all(ok), kv
    # scaled: encode(x, s) == encode(x / s, 1)
    y = synth.tensor((1000,), 9, torch.bfloat16, 3.0)
    assert torch.equal(oracle.fp8_encode(y, 0.05, "fp8"), oracle.fp8_encode((y.float() / 0.05), 1.0, "fp8"))


def test_gemm_oracle():
    a = synth.tensor((5, 64), 11, torch.bfloat16, 1.0)
    w = synth.tensor((64, 32), 12, torch.bfloat16, 1.0)
    c = oracle.gemm(a, w)
    assert torch.allclose(c, a.double() @ w.double(), atol=1e-12)


def test_copy_blocks_oracle_vs_live_reference_cpu():
    """oracle.copy_blocks against the UNMODIFIED reference CPU backend (kernels/cpu/cache.cpp:136-152), bit-exact."""
    ref = oracle.refcpu()
    if ref is None or not hasattr(ref, "copy_blocks"):
        pytest.skip("oracle/_ref not built")
    layers, nb = 3, 24
    kcs = [synth.tensor((nb, 4, 8, 16, 8), 70 + l, torch.bfloat16, 1.0) for l in range(layers)]
    vcs = [synth.tensor((nb, 4, 64, 16), 80 + l, torch.bfloat16, 1.0) for l in range(layers)]
    mapping = torch.tensor([[1, 20], [5, 21], [1, 22], [7, 3]], dtype=torch.int64)
    ka, va = [t.clone() for t in kcs], [t.clone() for t in vcs]
    kb, vb = [t.clone() for t in kcs], [t.clone() for t in vcs]
    oracle.copy_blocks(ka, va, mapping)
    ref.copy_blocks(kb, vb, mapping)
    for l in range(layers):
        assert torch.equal(ka[l].view(torch.int16), kb[l].view(torch.int16))
        assert torch.equal(va[l].view(torch.int16), vb[l].view(torch.int16))


def test_repack_oracle_matches_reference_python():
    """gptq_pack / awq_pack and the repack contract against fixtures made by the reference's own Python
    (tests/golden/make_golden.py make_repack)."""
    from oracle import marlin_format as mf
    g = np.load(os.path.join(G, "marlin_repack_golden.npz"))
    for ci in range(int(g["num_cases"][0])):
        k, n, bits, act = (int(v) for v in g[f"c{ci}_meta"])
        q = g[f"c{ci}_q"]
        np.testing.assert_array_equal(mf.gptq_pack(q, bits), g[f"c{ci}_gptq"])
        np.testing.assert_array_equal(mf.awq_pack(q, bits), g[f"c{ci}_awq"])
        np.testing.assert_array_equal(mf.gptq_unpack(g[f"c{ci}_gptq"], bits), q)
        np.testing.assert_array_equal(mf.awq_unpack(g[f"c{ci}_awq"], bits), q)
        np.testing.assert_array_equal(mf.gptq_marlin_repack(g[f"c{ci}_gptq"], None, k, n, bits), g[f"c{ci}_marlin"])
        np.testing.assert_array_equal(mf.awq_marlin_repack(g[f"c{ci}_awq"], k, n, bits), g[f"c{ci}_marlin"])
        if act:
            np.testing.assert_array_equal(
                mf.gptq_marlin_repack(g[f"c{ci}_gptq"], g[f"c{ci}_sort_indices"], k, n, bits), g[f"c{ci}_marlin_sorted"])


def _rope_cache(g, pre, dt):
    """Rebuild a [max_position, rot_dim] table holding the rows the fixture carries."""
    rows = torch.from_numpy(g[pre + "cache_rows"])
    part = _t(g[pre + "cache"], dt)
    cache = torch.zeros(4096, part.shape[1], dtype=dt)
    cache[rows] = part
    return cache


def test_glue_oracle_matches_reference_forward_native():
    """oracle/layer_glue.py against the reference's RMSNorm / SiluAndMul / RotaryEmbedding forward_native
    (tests/golden/make_golden.py make_glue). Rotary embedding rounds at the same places in both and must be
    bit-exact; the norms and SiLU round to the storage type at different places in the CUDA kernels (which the
    oracle follows) and in forward_native, so they carry the tolerance of the reference's own kernel tests
    (tests/kernels/test_layernorm.py: atol = rtol = 1e-2)."""
    from oracle import layer_glue as lg
    g = np.load(os.path.join(G, "glue_golden.npz"))
    for ci in range(int(g["num_norm"][0])):
        pre = f"norm{ci}_"
        dt = DT[int(g[pre + "meta"][2])]
        x, res, w = _t(g[pre + "x"], dt), _t(g[pre + "res"], dt), _t(g[pre + "w"], dt)
        eps = float(g[pre + "eps"][0])
        torch.testing.assert_close(lg.rms_norm(x, w, eps), _t(g[pre + "out"], dt), atol=1e-2, rtol=1e-2)
        o, r = lg.fused_add_rms_norm(x, res, w, eps)
        torch.testing.assert_close(o, _t(g[pre + "fused_out"], dt), atol=1e-2, rtol=1e-2)
        if dt == torch.float32:
            torch.testing.assert_close(r, _t(g[pre + "fused_res"], dt), atol=1e-6, rtol=1e-6)
        else:
            assert torch.equal(r, _t(g[pre + "fused_res"], dt))  # one rounding of the same fp32 sum
        torch.testing.assert_close(lg.silu_and_mul(_t(g[pre + "silu_x"], dt)), _t(g[pre + "silu_out"], dt), atol=1e-3,
                                   rtol=1e-2)
        sx = _t(g[pre + "silu_x"], dt)
        # tests/kernels/test_activation.py compares the CUDA kernels with forward_native at atol = rtol = 1e-2 for
        # the gelu_new / gelu_fast / gelu_quick family and at the default bf16 / fp16 tolerances for the gated ones
        for name in ("mul_and_silu", "gelu_and_mul", "gelu_tanh_and_mul"):
            torch.testing.assert_close(lg.act_and_mul(name, sx), _t(g[pre + name + "_out"], dt), atol=1e-3, rtol=1.6e-2)
        torch.testing.assert_close(lg.act_and_mul("fatrelu_and_mul", sx, 0.3), _t(g[pre + "fatrelu_and_mul_out"], dt), atol=0,
                                   rtol=0)
        for name in ("gelu_new", "gelu_fast", "gelu_quick"):
            torch.testing.assert_close(lg.activation(name, sx), _t(g[pre + name + "_out"], dt), atol=1e-2, rtol=1e-2)
    for ci in range(int(g["num_rope"][0])):
        pre = f"rope{ci}_"
        t, nh, nkv, d, rot, neox, dc = (int(v) for v in g[pre + "meta"])
        dt = DT[dc]
        q, k = lg.rotary_embedding(torch.from_numpy(g[pre + "pos"]), _t(g[pre + "q"], dt), _t(g[pre + "k"], dt), d,
                                   _rope_cache(g, pre, dt), bool(neox))
        if dt == torch.float32:
            torch.testing.assert_close(q, _t(g[pre + "q_out"], dt), atol=1e-6, rtol=1e-6)
            torch.testing.assert_close(k, _t(g[pre + "k_out"], dt), atol=1e-6, rtol=1e-6)
        else:
            assert torch.equal(q, _t(g[pre + "q_out"], dt)) and torch.equal(k, _t(g[pre + "k_out"], dt)), ci



def test_advance_step_flashinfer_oracle_known_answer():
    """Hand-worked case of prepare_inputs/advance_step.cu:91-167: block size 4, three sequences of which two run.
    Lengths 3 and 8 become 4 and 9: positions 3 and 8, pages 1 and 3, last-page fill 4 and 1."""
    import torch
    from oracle import layer_glue as lg
    tok, pos, slots = torch.tensor([5, 6, 7]), torch.tensor([9, 9, 9]), torch.tensor([1, 1, 1])
    sampled, lens = torch.tensor([[11], [12]]), torch.tensor([3, 8, 2], dtype=torch.int32)
    bt = torch.tensor([[10, 11, 12], [20, 21, 22], [30, 31, 32]], dtype=torch.int32)
    indices, indptr = torch.full((8,), -1, dtype=torch.int32), torch.full((4,), -1, dtype=torch.int32)
    last, bound = torch.full((3,), -1, dtype=torch.int32), torch.full((3,), -1, dtype=torch.int32)
    lg.advance_step_flashinfer(3, 2, 4, tok, sampled, pos, lens, slots, bt, indices, indptr, last, bound)
    assert tok.tolist() == [11, 12, 0] and pos.tolist() == [3, 8, 0] and lens.tolist() == [4, 9, 2]
    assert slots.tolist() == [10 * 4 + 3, 22 * 4 + 0, -1]
    assert last.tolist() == [4, 1, -1] and bound.tolist() == [1, 3, -1]
    assert indptr.tolist() == [0, 1, 4, 4] and indices.tolist() == [10, 20, 21, 22, -1, -1, -1, -1]
