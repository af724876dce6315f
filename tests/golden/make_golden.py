"""Generate the golden fixtures under tests/golden/ FROM THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference and oracle/_ref built by
`make -C oracle ref`):   python tests/golden/make_golden.py

Sources of truth:
  * Marlin format / quantiser / ScalarType ids: the reference's own Python files, loaded by path
    (aphrodite/quantization/utils/{quant_utils,marlin_utils,marlin_utils_test}.py,
    aphrodite/scalar_type.py) with stub modules for the unrelated imports they pull in.
  * paged_attention_v1/v2 and reshape_and_cache: the UNMODIFIED reference CPU backend
    (kernels/cpu/attention.cpp, kernels/cpu/cache.cpp) compiled into oracle/_ref.
The fixtures are what travels to the GPU box; /root/reference does not.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def load_reference_python():
    _stub("aphrodite")
    st = _load("aphrodite.scalar_type", f"{REF}/aphrodite/scalar_type.py")
    _stub("aphrodite.common")
    _stub("aphrodite.common.envs")
    sys.modules["aphrodite.common"].envs = sys.modules["aphrodite.common.envs"]
    _stub("aphrodite._custom_ops")
    sys.modules["aphrodite"]._custom_ops = sys.modules["aphrodite._custom_ops"]
    _stub("aphrodite.modeling")
    _stub("aphrodite.modeling.layers")
    _stub("aphrodite.modeling.layers.linear", LinearBase=object)
    _stub("aphrodite.platforms", current_platform=None)
    _stub("aphrodite.quantization")
    _stub("aphrodite.quantization.qqq", MARLIN_QQQ_SUPPORTED_NUM_BITS=[4])
    _stub("aphrodite.quantization.utils")
    base = f"{REF}/aphrodite/quantization/utils"
    qu = _load("aphrodite.quantization.utils.quant_utils", f"{base}/quant_utils.py")
    mu = _load("aphrodite.quantization.utils.marlin_utils", f"{base}/marlin_utils.py")
    mt = _load("aphrodite.quantization.utils.marlin_utils_test", f"{base}/marlin_utils_test.py")
    return st, qu, mu, mt


def raw(t: torch.Tensor) -> np.ndarray:
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.contiguous().view(torch.int16).numpy()
    return t.contiguous().numpy()


def make_marlin(st, qu, mu, mt):
    out = {}
    ids = st.scalar_types
    out["id_uint4"] = np.array([ids.uint4.id], dtype=np.int64)
    out["id_uint4b8"] = np.array([ids.uint4b8.id], dtype=np.int64)
    out["id_uint8"] = np.array([ids.uint8.id], dtype=np.int64)
    out["id_uint8b128"] = np.array([ids.uint8b128.id], dtype=np.int64)
    out["perm4"] = mt.get_weight_perm(4).numpy()
    out["perm8"] = mt.get_weight_perm(8).numpy()
    sp, sps = mu.get_scale_perms()
    out["scale_perm"] = np.array(sp)
    out["scale_perm_single"] = np.array(sps)
    cases = []
    g = torch.Generator().manual_seed(0)
    # (K, N, bits, group, zero_points, dtype)
    for ci, (k, n, bits, gs, zp, dt) in enumerate([
        (128, 64, 4, 128, False, torch.bfloat16),
        (256, 128, 4, 128, False, torch.bfloat16),
        (256, 128, 4, -1, False, torch.float16),
        (256, 64, 4, 32, False, torch.float16),
        (128, 128, 8, 64, False, torch.float16),
        (256, 128, 4, 128, True, torch.float16),
        (128, 64, 8, -1, False, torch.bfloat16),
        (256, 64, 4, 64, True, torch.bfloat16),
    ]):
        w = torch.randn(k, n, generator=g).to(dt)
        if zp:
            qt = ids.uint4 if bits == 4 else ids.uint8
            w_ref, bq, s, bz = mt.awq_marlin_quantize(w, qt, gs)
        else:
            qt = ids.uint4b8 if bits == 4 else ids.uint8b128
            w_ref, bq, s, _gi, _si, _rp = mt.marlin_quantize(w, qt, gs, False)
            bz = None
        pre = f"c{ci}_"
        out[pre + "meta"] = np.array([k, n, bits, gs, int(zp), {torch.bfloat16: 2, torch.float16: 1}[dt]])
        out[pre + "w"] = raw(w)
        out[pre + "w_ref"] = raw(w_ref)
        out[pre + "b_q_weight"] = bq.numpy()
        out[pre + "b_scales"] = raw(s)
        if bz is not None:
            out[pre + "b_zeros"] = bz.numpy()
        cases.append(ci)
    out["num_cases"] = np.array([len(cases)])
    np.savez_compressed(os.path.join(HERE, "marlin_golden.npz"), **out)
    print("marlin_golden.npz:", len(out), "arrays")


def make_repack(st, qu, mu, mt):
    """Checkpoint formats and the Marlin packing they must repack to, all from the reference's Python
    (quant_utils.gptq_pack / awq_pack / sort_weights, marlin_utils_test.marlin_weights): the contract of
    tests/kernels/test_marlin_gemm.py:66-176."""
    out = {}
    g = torch.Generator().manual_seed(1)
    # (K, N, bits, act_order)
    cases = [(128, 64, 4, False), (256, 192, 4, True), (128, 128, 8, False), (256, 64, 8, True), (64, 256, 4, False)]
    for ci, (k, n, bits, act) in enumerate(cases):
        q = torch.randint(0, 1 << bits, (k, n), generator=g, dtype=torch.int32)
        pre = f"c{ci}_"
        out[pre + "meta"] = np.array([k, n, bits, int(act)])
        out[pre + "q"] = q.numpy()
        out[pre + "gptq"] = qu.gptq_pack(q, bits, k, n).numpy()
        out[pre + "awq"] = qu.awq_pack(q, bits, k, n).numpy()
        perm = mt.get_weight_perm(bits)
        out[pre + "marlin"] = mt.marlin_weights(q, k, n, bits, perm).numpy()
        if act:
            g_idx = torch.randint(0, k // 32, (k,), generator=g, dtype=torch.int32)
            q_sorted, _, sort_indices = qu.sort_weights(q, g_idx)
            out[pre + "sort_indices"] = sort_indices.numpy()
            out[pre + "marlin_sorted"] = mt.marlin_weights(q_sorted, k, n, bits, perm).numpy()
    out["num_cases"] = np.array([len(cases)])
    np.savez_compressed(os.path.join(HERE, "marlin_repack_golden.npz"), **out)
    print("marlin_repack_golden.npz:", len(out), "arrays")


def make_marlin_utils(st, qu, mu, mt):
    """Host helpers of the Marlin format (marlin_utils.py:188-300, quant_utils.py:485-540) evaluated by the reference's own
    Python: scale / zero-point permutations, column packers, g_idx sort, ScalarType ids, the support predicates."""
    out = {}
    g = torch.Generator().manual_seed(7)
    sp, sps = mu.get_scale_perms()
    out["scale_perm"], out["scale_perm_single"] = np.array(sp), np.array(sps)
    # (size_k, size_n, group_size, num_bits)
    cases = [(256, 128, 128, 4), (256, 192, 64, 4), (128, 256, 32, 8), (512, 64, -1, 4), (256, 320, 128, 8)]
    for ci, (k, n, gs, bits) in enumerate(cases):
        pre = f"c{ci}_"
        groups = 1 if gs == -1 else k // gs
        s = torch.randn(groups, n, generator=g)
        zp = torch.randint(0, 1 << bits, (groups, n), generator=g, dtype=torch.int32)
        out[pre + "meta"] = np.array([k, n, gs, bits, groups])
        out[pre + "s"], out[pre + "zp"] = s.numpy(), zp.numpy()
        out[pre + "s_marlin"] = mu.marlin_permute_scales(s, k, n, gs).numpy()
        out[pre + "zp_marlin"] = mu.marlin_zero_points(zp, groups, n, bits).numpy()
        packed = qu.pack_cols(zp, bits, groups, n)
        out[pre + "zp_packed"] = packed.numpy()
        out[pre + "zp_unpacked"] = qu.unpack_cols(packed, bits, groups, n).numpy()
        # an AWQ checkpoint's zero points: interleaved for AWQ's dequantiser, then packed along the columns
        inter = np.array([0, 2, 4, 6, 1, 3, 5, 7]) if bits == 4 else np.array([0, 2, 1, 3])
        awq = zp.reshape(-1, len(inter))[:, inter].reshape(groups, n).contiguous()
        awq_packed = qu.pack_cols(awq, bits, groups, n)
        out[pre + "awq_zp_packed"] = awq_packed.numpy()
        out[pre + "awq_zp_marlin"] = mu.awq_to_marlin_zero_points(awq_packed, groups, n, bits).numpy()
    g_idx = torch.randint(0, 8, (256,), generator=g, dtype=torch.int32)
    srt, order = mu.marlin_sort_g_idx(g_idx)
    out["g_idx"], out["g_idx_sorted"], out["g_idx_order"] = g_idx.numpy(), srt.numpy(), order.numpy()
    moe_s = torch.randn(3, 2, 128, generator=g)
    out["moe_s"], out["moe_s_marlin"] = moe_s.numpy(), mu.marlin_moe_permute_scales(moe_s, 256, 128, 128).numpy()
    for name in ("uint4", "uint8", "uint4b8", "uint8b128"):
        out["id_" + name] = np.array([getattr(st.scalar_types, name).id], dtype=np.int64)
    # the support predicates: (type, group size, has_zp) -> supported?
    rows = []
    for ti, name in enumerate(("uint4", "uint8", "uint4b8", "uint8b128")):
        for gs in (-1, 32, 64, 128, 256):
            for has_zp in (False, True):
                rows.append([ti, gs, int(has_zp), int(mu.check_marlin_supported(getattr(st.scalar_types, name), gs, has_zp, 100))])
    out["supported"] = np.array(rows)
    shapes = []
    for o, i, full, gs in ((4096, 4096, 4096, 128), (4100, 4096, 4096, 128), (4096, 4000, 4096, 128), (1024, 1792, 14336, 128),
                           (1024, 1856, 14848, 128), (512, 1024, 8192, -1), (512, 1088, 8704, 64)):
        ok, _ = mu.check_marlin_supports_shape(o, i, full, gs)
        shapes.append([o, i, full, gs, int(ok)])
    out["shapes"] = np.array(shapes)
    out["num_cases"] = np.array([len(cases)])
    np.savez_compressed(os.path.join(HERE, "marlin_utils_golden.npz"), **out)
    print("marlin_utils_golden.npz:", len(out), "arrays")


def load_reference_layers():
    """The reference's RMSNorm / SiluAndMul / RotaryEmbedding modules, loaded by path with the platform pinned to
    'not CUDA' so that only their forward_native code is used."""
    import torch.nn as nn

    class CustomOp(nn.Module):
        @classmethod
        def register(cls, name):
            return lambda c: c

    class _Platform:
        def __getattr__(self, name):
            return lambda *a, **k: False

    _stub("aphrodite.modeling._custom_op", CustomOp=CustomOp)
    sys.modules["aphrodite.platforms"].current_platform = _Platform()
    _stub("aphrodite.common.utils", LazyDict=dict)
    _stub("aphrodite.distributed", divide=None, get_tensor_model_parallel_rank=None, get_tensor_model_parallel_world_size=None)
    _stub("aphrodite.modeling.utils", set_weight_attrs=None)
    base = f"{REF}/aphrodite/modeling/layers"
    ln = _load("aphrodite.modeling.layers.layernorm", f"{base}/layernorm.py")
    act = _load("aphrodite.modeling.layers.activation", f"{base}/activation.py")
    rope = _load("aphrodite.modeling.layers.rotary_embedding", f"{base}/rotary_embedding.py")
    return ln, act, rope


@torch.no_grad()
def make_glue():
    """Inputs and forward_native outputs of the decode-layer glue (reference tests: tests/kernels/test_layernorm.py,
    test_activation.py, test_pos_encoding.py compare the CUDA kernels against exactly these methods)."""
    ln, act, rope = load_reference_layers()
    out = {}
    g = torch.Generator().manual_seed(2)
    cases = [(7, 512, torch.bfloat16), (3, 4096, torch.float16), (5, 768, torch.float32), (4, 8192, torch.bfloat16)]
    for ci, (t, h, dt) in enumerate(cases):
        layer = ln.RMSNorm(h).to(dt)
        layer.weight.data.normal_(mean=1.0, std=0.1, generator=g)
        x = (torch.randn(t, h, generator=g) * (1 / (2 * h)) ** 0.5 * 30).to(dt)
        res = (torch.randn(t, h, generator=g) * (1 / (2 * h)) ** 0.5 * 30).to(dt)
        pre = f"norm{ci}_"
        out[pre + "meta"] = np.array([t, h, {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[dt]])
        out[pre + "x"], out[pre + "res"], out[pre + "w"] = raw(x), raw(res), raw(layer.weight.data)
        out[pre + "eps"] = np.array([layer.variance_epsilon])
        out[pre + "out"] = raw(layer.forward_native(x))
        o2, r2 = layer.forward_native(x, res)
        out[pre + "fused_out"], out[pre + "fused_res"] = raw(o2), raw(r2)
        xs = torch.randn(t, 2 * h, generator=g).to(dt)
        out[pre + "silu_x"] = raw(xs)
        out[pre + "silu_out"] = raw(act.SiluAndMul().forward_native(xs))
        out[pre + "mul_and_silu_out"] = raw(act.MulAndSilu().forward_native(xs))
        out[pre + "gelu_and_mul_out"] = raw(act.GeluAndMul("none").forward_native(xs))
        out[pre + "gelu_tanh_and_mul_out"] = raw(act.GeluAndMul("tanh").forward_native(xs))
        out[pre + "fatrelu_and_mul_out"] = raw(act.FatreluAndMul(0.3).forward_native(xs))
        out[pre + "gelu_new_out"] = raw(act.NewGELU().forward_native(xs))
        out[pre + "gelu_fast_out"] = raw(act.FastGELU().forward_native(xs))
        out[pre + "gelu_quick_out"] = raw(act.QuickGELU().forward_native(xs))
    # (tokens, heads, kv heads, head size, rotary dim, neox, dtype)
    rcases = [(9, 8, 2, 128, 128, True, torch.bfloat16), (5, 4, 4, 64, 32, False, torch.float16),
              (6, 6, 2, 96, 96, False, torch.bfloat16), (4, 8, 1, 128, 64, True, torch.float32)]
    for ci, (t, nh, nkv, d, rot, neox, dt) in enumerate(rcases):
        layer = rope.RotaryEmbedding(d, rot, 4096, 10000, neox, dt)
        pos = torch.randint(0, 4096, (t,), generator=g)
        q = torch.randn(t, nh * d, generator=g).to(dt)
        k = torch.randn(t, nkv * d, generator=g).to(dt)
        pre = f"rope{ci}_"
        out[pre + "meta"] = np.array([t, nh, nkv, d, rot, int(neox), {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[dt]])
        out[pre + "pos"], out[pre + "q"], out[pre + "k"] = pos.numpy(), raw(q), raw(k)
        out[pre + "cache"] = raw(layer.cos_sin_cache[pos.unique()])  # only the rows that are used travel
        out[pre + "cache_rows"] = pos.unique().numpy()
        qo, ko = layer.forward_native(pos, q.clone(), k.clone())
        out[pre + "q_out"], out[pre + "k_out"] = raw(qo), raw(ko)
    out["num_norm"], out["num_rope"] = np.array([len(cases)]), np.array([len(rcases)])
    np.savez_compressed(os.path.join(HERE, "glue_golden.npz"), **out)
    print("glue_golden.npz:", len(out), "arrays")


def load_refcpu():
    flags = open("/proc/cpuinfo").read()
    isa = "avx512bf16" if "avx512_bf16" in flags else ("avx512" if "avx512f" in flags else "avx2")
    torch.ops.load_library(os.path.join(ROOT, "oracle", "_ref", f"refcpu_{isa}.so"))
    return torch.ops._refcpu


def make_attention(ref):
    out = {}
    g = torch.Generator().manual_seed(0)
    # (S, H, KVH, D, max_len, dtype, alibi, version)
    cases = [
        (3, 8, 8, 64, 100, torch.bfloat16, False, 1),
        (4, 8, 2, 128, 700, torch.bfloat16, False, 2),
        (2, 4, 4, 128, 200, torch.float16, True, 1),
        (3, 16, 2, 128, 1100, torch.float16, False, 2),
        (2, 4, 1, 64, 50, torch.float32, False, 1),
        (2, 12, 12, 64, 512, torch.bfloat16, False, 1),
        (2, 8, 2, 256, 300, torch.bfloat16, True, 2),
        (2, 6, 2, 96, 130, torch.float16, False, 1),
    ]
    bs = 16
    for ci, (S, H, KVH, D, max_len, dt, alibi, ver) in enumerate(cases):
        x = 16 // torch.tensor([], dtype=dt).element_size()
        sl = torch.randint(1, max_len + 1, (S,), generator=g, dtype=torch.int32)
        sl[-1] = max_len
        mb = (max_len + bs - 1) // bs
        nb = S * mb + 3
        bt = torch.from_numpy(synth.permutation(nb, 100 + ci)[: S * mb].reshape(S, mb)).to(torch.int32)
        sc = float(D) ** -0.5
        q = synth.tensor((S, H, D), 200 + ci, dt, sc)
        kc = synth.tensor((nb, KVH, D // x, bs, x), 300 + ci, dt, sc)
        vc = synth.tensor((nb, KVH, D, bs), 400 + ci, dt, sc)
        slopes = torch.randn(H, generator=g).float() if alibi else None
        o = torch.empty_like(q)
        one = torch.tensor(1.0)
        if ver == 1:
            ref.paged_attention_v1(o, q, kc, vc, KVH, sc, bt, sl, bs, max_len, slopes, "auto", one, one,
                                   0, 0, 0, 64, 0)
        else:
            P = (max_len + 511) // 512
            es = torch.empty(S, H, P)
            ml = torch.empty(S, H, P)
            tmp = torch.empty(S, H, P, D, dtype=dt)
            ref.paged_attention_v2(o, es, ml, tmp, q, kc, vc, KVH, sc, bt, sl, bs, max_len, slopes, "auto",
                                   one, one, 0, 0, 0, 64, 0)
        pre = f"a{ci}_"
        out[pre + "meta"] = np.array([S, H, KVH, D, max_len, {torch.float32: 0, torch.float16: 1,
                                                               torch.bfloat16: 2}[dt], int(alibi), ver, nb])
        out[pre + "seq_lens"] = sl.numpy()
        out[pre + "block_tables"] = bt.numpy()
        if alibi:
            out[pre + "alibi"] = slopes.numpy()
        out[pre + "out"] = raw(o)
    out["num_cases"] = np.array([len(cases)])
    np.savez_compressed(os.path.join(HERE, "attention_golden.npz"), **out)
    print("attention_golden.npz:", len(out), "arrays")


def make_cache(ref):
    out = {}
    g = torch.Generator().manual_seed(0)
    cases = [(42, 8, 64, 16, 8, torch.bfloat16), (42, 4, 128, 16, 4, torch.float32),
             (17, 2, 80, 16, 4, torch.float32), (9, 2, 256, 16, 2, torch.bfloat16)]
    for ci, (T, KVH, D, bs, nb, dt) in enumerate(cases):
        x = 16 // torch.tensor([], dtype=dt).element_size()
        qkv = synth.tensor((T, 3, KVH, D), 500 + ci, dt, 2.0)
        key, value = qkv[:, 1], qkv[:, 2]
        slots = torch.from_numpy(synth.permutation(nb * bs, 600 + ci)[:T]).to(torch.int64)
        slots[T // 2] = -1
        kc = synth.tensor((nb, KVH, D // x, bs, x), 700 + ci, dt, 2.0)
        vc = synth.tensor((nb, KVH, D, bs), 800 + ci, dt, 2.0)
        kc0, vc0 = kc.clone(), vc.clone()
        one = torch.tensor(1.0)
        ref.reshape_and_cache(key, value, kc, vc, slots, "auto", one, one)
        pre = f"r{ci}_"
        out[pre + "meta"] = np.array([T, KVH, D, bs, nb, {torch.float32: 0, torch.float16: 1,
                                                           torch.bfloat16: 2}[dt]])
        out[pre + "slots"] = slots.numpy()
        # only the touched blocks change; store the whole result (small pools)
        out[pre + "kc1"] = raw(kc)
        out[pre + "vc1"] = raw(vc)
    out["num_cases"] = np.array([len(cases)])
    np.savez_compressed(os.path.join(HERE, "cache_golden.npz"), **out)
    print("cache_golden.npz:", len(out), "arrays")


if __name__ == "__main__":
    torch.manual_seed(0)
    mods = load_reference_python()
    if sys.argv[1:] == ["repack"]:
        make_repack(*mods)
        sys.exit(0)
    if sys.argv[1:] == ["marlin_utils"]:
        make_marlin_utils(*mods)
        sys.exit(0)
    if sys.argv[1:] == ["glue"]:
        make_glue()
        sys.exit(0)
    make_marlin(*mods)
    make_repack(*mods)
    make_marlin_utils(*mods)
    make_glue()
    ref = load_refcpu()
    make_attention(ref)
    make_cache(ref)
