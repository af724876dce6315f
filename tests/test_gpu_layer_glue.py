"""Decode-layer glue on the GPU (rms_norm, fused_add_rms_norm, silu_and_mul, rotary_embedding): against the
reference-Python fixtures, the oracle (oracle/layer_glue.py) and - when oracle/_ref/refcuda.so is built - the
reference's own CUDA kernels. Shapes follow /root/reference/tests/kernels/test_layernorm.py,
test_activation.py and test_pos_encoding.py.

Tolerances: silu_and_mul and rotary_embedding are elementwise with the reference's roundings -> bit-exact against
the reference CUDA kernels and the oracle (rotary; SiLU up to the libm/CUDA expf difference on the CPU oracle).
The norms differ from the reference only through the order of the fp32 sum of squares: at most 2 ulp of the
storage type on a tiny fraction of elements."""
import os

import numpy as np
import pytest
import torch

import aphrodite_fork_b200 as A
from oracle import layer_glue as lg
from oracle import oracle

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEV = "cuda"
DT = {0: torch.float32, 1: torch.float16, 2: torch.bfloat16}
DTYPES = [torch.bfloat16, torch.float16, torch.float32]


def _t(a, dt):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return (t.view(dt) if dt in (torch.float16, torch.bfloat16) else t).to(DEV)


def _ulp_close(a: torch.Tensor, b: torch.Tensor, frac: float = 0.01):
    """Equal except for at most `frac` of the elements, which may differ by up to two units in the last place (one ulp
    in the rounding of x * rsqrt, carried through the multiplication by a weight > 1)."""
    if a.dtype == torch.float32:
        torch.testing.assert_close(a, b, atol=1e-6, rtol=1e-5)
        return
    ia, ib = a.view(torch.int16).int(), b.view(torch.int16).int()
    d = (ia - ib).abs()
    # values in the far tails of the GELUs (x < -4: 1 + erf(x) cancels to a few fp32 ulps, so libm and libdevice agree to
    # ~1e-7 absolutely but not relatively) are compared absolutely: differences below 1e-5 on activations of order 1
    tiny = (a.float() - b.float()).abs() <= 1e-5
    assert bool(((d <= 2) | tiny).all()), int(d[~tiny].max())
    assert float(((d > 0) & ~tiny).float().mean()) <= frac, float((d > 0).float().mean())


def _ref():
    r = oracle.refcuda()
    return r if r is not None and hasattr(r, "rms_norm") else None


def test_glue_golden_fixtures():
    g = np.load(os.path.join(G, "glue_golden.npz"))
    for ci in range(int(g["num_norm"][0])):
        pre = f"norm{ci}_"
        dt = DT[int(g[pre + "meta"][2])]
        x, res, w = _t(g[pre + "x"], dt), _t(g[pre + "res"], dt), _t(g[pre + "w"], dt)
        eps = float(g[pre + "eps"][0])
        out = torch.empty_like(x)
        A.ops.rms_norm(out, x, w, eps)
        torch.testing.assert_close(out, _t(g[pre + "out"], dt), atol=1e-2, rtol=1e-2)
        xi, ri = x.clone(), res.clone()
        A.ops.fused_add_rms_norm(xi, ri, w, eps)
        torch.testing.assert_close(xi, _t(g[pre + "fused_out"], dt), atol=1e-2, rtol=1e-2)
        torch.testing.assert_close(ri, _t(g[pre + "fused_res"], dt), atol=1e-6, rtol=1e-6)
        sx = _t(g[pre + "silu_x"], dt)
        so = torch.empty(sx.shape[0], sx.shape[1] // 2, dtype=dt, device=DEV)
        A.ops.silu_and_mul(so, sx)
        torch.testing.assert_close(so, _t(g[pre + "silu_out"], dt), atol=1e-3, rtol=1e-2)
    for ci in range(int(g["num_rope"][0])):
        pre = f"rope{ci}_"
        t, nh, nkv, d, rot, neox, dc = (int(v) for v in g[pre + "meta"])
        dt = DT[dc]
        cache = torch.zeros(4096, rot, dtype=dt, device=DEV)
        cache[torch.from_numpy(g[pre + "cache_rows"]).to(DEV)] = _t(g[pre + "cache"], dt)
        q, k = _t(g[pre + "q"], dt), _t(g[pre + "k"], dt)
        A.ops.rotary_embedding(torch.from_numpy(g[pre + "pos"]).to(DEV), q, k, d, cache, bool(neox))
        if dt == torch.float32:
            torch.testing.assert_close(q, _t(g[pre + "q_out"], dt), atol=1e-6, rtol=1e-6)
            torch.testing.assert_close(k, _t(g[pre + "k_out"], dt), atol=1e-6, rtol=1e-6)
        else:
            assert torch.equal(q, _t(g[pre + "q_out"], dt)) and torch.equal(k, _t(g[pre + "k_out"], dt)), ci


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("tokens,hidden", [(1, 4096), (7, 768), (64, 4096), (83, 5120), (64, 8192), (4096, 4096), (5, 8199),
                                           (3, 40000)])
def test_rms_norm_and_fused_add(tokens, hidden, dtype):
    gen = torch.Generator(device=DEV).manual_seed(tokens + hidden)
    x = torch.randn(tokens, hidden, device=DEV, generator=gen).to(dtype)
    res = torch.randn(tokens, hidden, device=DEV, generator=gen).to(dtype)
    w = (1 + 0.1 * torch.randn(hidden, device=DEV, generator=gen)).to(dtype)
    out = torch.empty_like(x)
    A.ops.rms_norm(out, x, w, 1e-6)
    xi, ri = x.clone(), res.clone()
    A.ops.fused_add_rms_norm(xi, ri, w, 1e-6)
    want = lg.rms_norm(x.cpu(), w.cpu(), 1e-6)
    wo, wr = lg.fused_add_rms_norm(x.cpu(), res.cpu(), w.cpu(), 1e-6)
    _ulp_close(out.cpu(), want)
    _ulp_close(xi.cpu(), wo)
    assert torch.equal(ri.cpu(), wr)
    ref = _ref()
    if ref is not None:
        ro = torch.empty_like(x)
        ref.rms_norm(ro, x, w, 1e-6)
        rx, rr = x.clone(), res.clone()
        ref.fused_add_rms_norm(rx, rr, w, 1e-6)
        _ulp_close(out, ro)
        _ulp_close(xi, rx)
        assert torch.equal(ri, rr)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("tokens,d", [(1, 14336), (7, 512), (64, 14336), (83, 13824), (2048, 14336), (5, 1001)])
def test_silu_and_mul(tokens, d, dtype):
    gen = torch.Generator(device=DEV).manual_seed(tokens + d)
    x = (torch.randn(tokens, 2 * d, device=DEV, generator=gen) * 3).to(dtype)
    out = torch.empty(tokens, d, dtype=dtype, device=DEV)
    A.ops.silu_and_mul(out, x)
    _ulp_close(out.cpu(), lg.silu_and_mul(x.cpu()), frac=0.02)
    ref = _ref()
    if ref is not None:
        ro = torch.empty_like(out)
        ref.silu_and_mul(ro, x)
        assert torch.equal(out, ro)


def test_silu_and_mul_more_tokens_than_one_grid():
    x = torch.randn(70000, 2 * 64, device=DEV).to(torch.bfloat16)
    out = torch.empty(70000, 64, dtype=torch.bfloat16, device=DEV)
    A.ops.silu_and_mul(out, x)
    _ulp_close(out[-3000:].cpu(), lg.silu_and_mul(x[-3000:].cpu()), frac=0.02)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("neox", [True, False])
@pytest.mark.parametrize("tokens,nh,nkv,d,rot", [(64, 32, 8, 128, 128), (1, 8, 1, 128, 128), (11, 12, 12, 64, 32), (7, 6, 2, 96, 96),
                                                 (300, 64, 8, 128, 64), (5, 4, 4, 256, 256), (3, 4, 2, 80, 20), (9, 8, 8, 64, 8)])
def test_rotary_embedding(tokens, nh, nkv, d, rot, neox, dtype):
    gen = torch.Generator(device=DEV).manual_seed(tokens + nh + d)
    max_pos = 8192
    inv = 1.0 / (10000 ** (torch.arange(0, rot, 2, dtype=torch.float) / rot))
    fr = torch.einsum("i,j->ij", torch.arange(max_pos, dtype=torch.float), inv)
    cache = torch.cat((fr.cos(), fr.sin()), -1).to(dtype).to(DEV)
    pos = torch.randint(0, max_pos, (tokens,), device=DEV, generator=gen)
    # q and k as slices of one qkv projection output: token stride larger than the row
    qkv = torch.randn(tokens, (nh + 2 * nkv) * d, device=DEV, generator=gen).to(dtype)
    q, k, v = qkv.split([nh * d, nkv * d, nkv * d], dim=-1)
    v0 = v.clone()
    wq, wk = lg.rotary_embedding(pos.cpu(), q.cpu().contiguous(), k.cpu().contiguous(), d, cache.cpu(), neox)
    ref = _ref()
    if ref is not None:
        # copies (contiguous() would alias q when there is one token and the reference kernel works in place)
        rq, rk = q.clone(memory_format=torch.contiguous_format), k.clone(memory_format=torch.contiguous_format)
        ref.rotary_embedding(pos, rq, rk, d, cache, neox)
    A.ops.rotary_embedding(pos, q, k, d, cache, neox)
    assert torch.equal(v, v0)
    if dtype == torch.float32:
        torch.testing.assert_close(q.cpu(), wq, atol=1e-6, rtol=1e-6)
        torch.testing.assert_close(k.cpu(), wk, atol=1e-6, rtol=1e-6)
    else:
        assert torch.equal(q.cpu(), wq) and torch.equal(k.cpu(), wk)
    if ref is not None:
        if dtype == torch.float32:
            torch.testing.assert_close(q, rq, atol=1e-6, rtol=1e-6)
            torch.testing.assert_close(k, rk, atol=1e-6, rtol=1e-6)
        else:
            assert torch.equal(q, rq) and torch.equal(k, rk)


def test_rotary_embedding_batched_positions_shape():
    """positions [batch, seq] with query [batch, seq, heads * head_size] (pos_encoding_kernels.cu:149-156)."""
    b, s, nh, d = 3, 5, 4, 64
    cache = torch.randn(128, d, device=DEV).to(torch.bfloat16)
    pos = torch.randint(0, 128, (b, s), device=DEV)
    q = torch.randn(b, s, nh * d, device=DEV).to(torch.bfloat16)
    k = torch.randn(b, s, nh * d, device=DEV).to(torch.bfloat16)
    wq, wk = lg.rotary_embedding(pos.cpu(), q.cpu().reshape(b * s, -1), k.cpu().reshape(b * s, -1), d, cache.cpu(), True)
    A.ops.rotary_embedding(pos, q, k, d, cache, True)
    assert torch.equal(q.cpu().reshape(b * s, -1), wq) and torch.equal(k.cpu().reshape(b * s, -1), wk)


def test_glue_errors_and_empty():
    x = torch.zeros(0, 64, dtype=torch.bfloat16, device=DEV)
    w = torch.ones(64, dtype=torch.bfloat16, device=DEV)
    A.ops.rms_norm(torch.empty_like(x), x, w, 1e-6)
    A.ops.fused_add_rms_norm(x, x.clone(), w, 1e-6)
    A.ops.silu_and_mul(torch.zeros(0, 32, dtype=torch.bfloat16, device=DEV), x)
    y = torch.zeros(4, 64, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="dtype mismatch"):
        A.ops.rms_norm(torch.empty_like(y), y, w.float(), 1e-6)
    with pytest.raises(RuntimeError, match="shape mismatch"):
        A.ops.rms_norm(torch.empty_like(y), y, w[:32], 1e-6)
    with pytest.raises(RuntimeError, match="same number of tokens"):
        A.ops.rotary_embedding(torch.zeros(3, dtype=torch.int64, device=DEV), y, y.clone(), 64, torch.zeros(8, 64, dtype=torch.bfloat16, device=DEV), True)
    with pytest.raises(RuntimeError):  # CPU tensors: the dispatcher has no CPU kernel for the op (no fallback)
        A.ops.silu_and_mul(torch.zeros(4, 32, dtype=torch.bfloat16), y.cpu())


GATED = ["silu_and_mul", "mul_and_silu", "gelu_and_mul", "gelu_tanh_and_mul", "fatrelu_and_mul"]
PLAIN = ["gelu_new", "gelu_fast", "gelu_quick"]


def _call(o, name, out, x):
    if name == "fatrelu_and_mul":
        getattr(o, name)(out, x, 0.3)
    else:
        getattr(o, name)(out, x)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("name", GATED + PLAIN)
@pytest.mark.parametrize("tokens,d", [(7, 512), (64, 14336), (5, 1001)])
def test_activation_family(name, tokens, d, dtype):
    """kernels/activation_kernels.cu:64-225 / tests/kernels/test_activation.py. Bit-exact against the reference CUDA
    kernels for the 16-bit types (same operations, same roundings, same libdevice functions); against the CPU oracle
    within 2 ulp (libm vs libdevice exp / tanh / erf)."""
    gen = torch.Generator(device=DEV).manual_seed(tokens + d)
    gated = name in GATED
    x = (torch.randn(tokens, 2 * d if gated else d, device=DEV, generator=gen) * 2).to(dtype)
    out = torch.empty(tokens, d, dtype=dtype, device=DEV)
    _call(A.ops, name, out, x)
    want = lg.act_and_mul(name, x.cpu(), 0.3) if gated else lg.activation(name, x.cpu())
    if dtype == torch.float32:
        torch.testing.assert_close(out.cpu(), want, atol=2e-6, rtol=2e-5)
    else:
        _ulp_close(out.cpu(), want, frac=0.03)
    ref = _ref()
    if ref is not None and hasattr(ref, name):
        ro = torch.empty_like(out)
        _call(ref, name, ro, x)
        if dtype == torch.float32:
            torch.testing.assert_close(out, ro, atol=1e-6, rtol=1e-5)
        else:
            assert torch.equal(out, ro)


def test_activation_golden_fixtures():
    g = np.load(os.path.join(G, "glue_golden.npz"))
    for ci in range(int(g["num_norm"][0])):
        pre = f"norm{ci}_"
        dt = DT[int(g[pre + "meta"][2])]
        sx = _t(g[pre + "silu_x"], dt)
        for name in GATED + PLAIN:
            gated = name in GATED
            out = torch.empty(sx.shape[0], sx.shape[1] // 2 if gated else sx.shape[1], dtype=dt, device=DEV)
            _call(A.ops, name, out, sx)
            key = pre + ("silu_out" if name == "silu_and_mul" else name + "_out")
            torch.testing.assert_close(out, _t(g[key], dt), atol=1e-2, rtol=1.6e-2)


@pytest.mark.parametrize("neox", [True, False])
def test_batched_rotary_embedding(neox):
    """pos_encoding_kernels.cu:98-121: row = positions + offset of the token's scaling block in one stacked table."""
    tokens, nh, nkv, d = 33, 8, 2, 128
    gen = torch.Generator(device=DEV).manual_seed(5)
    cache = torch.randn(3 * 512, d, device=DEV, generator=gen).to(torch.bfloat16)
    pos = torch.randint(0, 512, (tokens,), device=DEV, generator=gen)
    off = torch.randint(0, 3, (tokens,), device=DEV, generator=gen) * 512
    q = torch.randn(tokens, nh * d, device=DEV, generator=gen).to(torch.bfloat16)
    k = torch.randn(tokens, nkv * d, device=DEV, generator=gen).to(torch.bfloat16)
    wq, wk = lg.rotary_embedding((pos + off).cpu(), q.cpu(), k.cpu(), d, cache.cpu(), neox)
    ref = _ref()
    if ref is not None and hasattr(ref, "batched_rotary_embedding"):
        rq, rk = q.clone(), k.clone()
        ref.batched_rotary_embedding(pos, rq, rk, d, cache, neox, d, off)
    A.ops.batched_rotary_embedding(pos, q, k, d, cache, neox, d, off)
    assert torch.equal(q.cpu(), wq) and torch.equal(k.cpu(), wk)
    if ref is not None and hasattr(ref, "batched_rotary_embedding"):
        assert torch.equal(q, rq) and torch.equal(k, rk)


@pytest.mark.parametrize("num_seqs,num_queries", [(8, 8), (64, 50), (1000, 700), (1, 0)])
def test_advance_step_flashattn(num_seqs, num_queries):
    """Integer bookkeeping: bit-exact against the oracle loop (prepare_inputs/advance_step.cu:12-66)."""
    bs, max_blocks = 16, 40
    gen = torch.Generator().manual_seed(num_seqs)
    tok = torch.randint(0, 32000, (num_seqs,), generator=gen)
    sampled = torch.randint(0, 32000, (num_queries, 1), generator=gen)
    pos = torch.randint(0, 500, (num_seqs,), generator=gen)
    lens = torch.randint(1, bs * max_blocks - 1, (num_seqs,), generator=gen, dtype=torch.int32)
    slots = torch.randint(0, 10000, (num_seqs,), generator=gen)
    bt = torch.randint(0, 5000, (num_seqs, max_blocks), generator=gen, dtype=torch.int32)
    dev = [t.to(DEV) for t in (tok, sampled, pos, lens, slots, bt)]
    ref = oracle.refcuda()
    if ref is not None and hasattr(ref, "advance_step_flashattn") and num_seqs > 0:
        rdev = [t.clone() for t in dev]
        ref.advance_step_flashattn(num_seqs, num_queries, bs, rdev[0], rdev[1], rdev[2], rdev[3], rdev[4], rdev[5])
    else:
        rdev = None
    A.ops.advance_step_flashattn(num_seqs, num_queries, bs, dev[0], dev[1], dev[2], dev[3], dev[4], dev[5])
    if rdev is not None:
        for got, want in zip(dev, rdev):
            assert torch.equal(got, want)
    lg.advance_step_flashattn(num_seqs, num_queries, bs, tok, sampled, pos, lens, slots, bt)
    for got, want in zip(dev, (tok, sampled, pos, lens, slots, bt)):
        assert torch.equal(got.cpu(), want)
    with pytest.raises(RuntimeError, match="is not as expected"):
        A.ops.advance_step_flashattn(num_seqs + 1, num_queries, bs, dev[0], dev[1], dev[2], dev[3], dev[4], dev[5])


@pytest.mark.parametrize("num_seqs,num_queries,max_blocks", [(8, 8, 40), (64, 50, 40), (3000, 2500, 24), (1, 0, 8), (300, 300, 1100)])
def test_advance_step_flashinfer(num_seqs, num_queries, max_blocks):
    """Integer bookkeeping plus the CSR page list: bit-exact against the oracle loops and the reference's three kernels
    (prepare_inputs/advance_step.cu:91-167). 2500 queries take the scan through three 1024-row chunks; 1100 columns take
    the page copy through two column chunks."""
    bs = 16
    gen = torch.Generator().manual_seed(num_seqs)
    tok = torch.randint(0, 32000, (num_seqs,), generator=gen)
    sampled = torch.randint(0, 32000, (num_queries, 1), generator=gen)
    pos = torch.randint(0, 500, (num_seqs,), generator=gen)
    lens = torch.randint(1, bs * max_blocks - 1, (num_seqs,), generator=gen, dtype=torch.int32)
    slots = torch.randint(0, 10000, (num_seqs,), generator=gen)
    bt = torch.randint(0, 5000, (num_seqs, max_blocks), generator=gen, dtype=torch.int32)
    indices = torch.full((num_seqs * max_blocks + 1,), -7, dtype=torch.int32)
    indptr = torch.full((num_seqs + 1,), -7, dtype=torch.int32)
    last = torch.full((num_seqs,), -7, dtype=torch.int32)
    bound = torch.full((num_seqs,), -7, dtype=torch.int32)
    host = (tok, sampled, pos, lens, slots, bt, indices, indptr, last, bound)
    dev = [t.to(DEV) for t in host]
    ref = oracle.refcuda()
    rdev = None
    if ref is not None and hasattr(ref, "advance_step_flashinfer"):
        rdev = [t.clone() for t in dev]
        ref.advance_step_flashinfer(num_seqs, num_queries, bs, *rdev)
    A.ops.advance_step_flashinfer(num_seqs, num_queries, bs, *dev)
    if rdev is not None:
        for got, want in zip(dev, rdev):
            assert torch.equal(got, want)
    lg.advance_step_flashinfer(num_seqs, num_queries, bs, *host)
    for got, want in zip(dev, host):
        assert torch.equal(got.cpu(), want)
    # a second step on the advanced state (what a multi-step runner does), and a too-short page list: no write lands
    # beyond its end (the element after the capacity given to the op keeps its value)
    A.ops.advance_step_flashinfer(num_seqs, num_queries, bs, *dev)
    lg.advance_step_flashinfer(num_seqs, num_queries, bs, *host)
    for got, want in zip(dev, host):
        assert torch.equal(got.cpu(), want)
    if num_queries:
        short = torch.full((int(host[7][num_queries]),), -7, dtype=torch.int32, device=DEV)
        A.ops.advance_step_flashinfer(num_seqs, num_queries, bs, *dev[:6], short[:-1], *dev[7:])
        assert int(short[-1]) == -7
    with pytest.raises(RuntimeError, match="is not as expected"):
        A.ops.advance_step_flashinfer(num_seqs + 1, num_queries, bs, *dev)
