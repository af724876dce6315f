"""GPU parity: paged_attention_v1 / _v2 through the registered torch ops (-> C ABI) against the
CPU oracle. Mirrors /root/reference/tests/kernels/test_attention.py:117-326 (sweep, ragged
sequence lengths, ALiBi, fp8 KV) and test_blocksparse_attention.py:143-330.

Tolerance: |out - truth| <= atol + rtol*|truth| with atol 1e-3 for bf16/f16 KV and
1e-2 for FP8 KV (the reference tests' atol), rtol = torch's default for the dtype
(bf16 1.6e-2, f16 1e-3). For FP8 the truth is attention over the cache dequantised as the
reference reads it (quantisation error excluded, test_attention.py:284-300)."""
import os

import numpy as np
import pytest
import torch

import synth
from helpers import check_close, make_attention_case
from oracle import oracle

pytestmark = pytest.mark.gpu
DT = {0: torch.float32, 1: torch.float16, 2: torch.bfloat16}
FULL = os.environ.get("APH_FAST_BUILD", "0") != "1"


def _A():
    import aphrodite_fork_b200 as A
    return A


def run_case(S, H, KVH, D, bs, seq_lens, dtype, kv, seed, version, alibi=False, k_scale=1.0, v_scale=1.0,
             sparse=None, q_strided=False, via_class=False):
    A = _A()
    q, kc, vc, bt, sl = make_attention_case(S, H, KVH, D, bs, seq_lens, dtype, kv, seed, k_scale, v_scale)
    al = synth.tensor((H,), seed + 9, torch.float32, 1.0) if alibi else None
    sp = sparse or dict(tp_rank=0, blocksparse_local_blocks=0, blocksparse_vert_stride=0, blocksparse_block_size=64,
                        blocksparse_head_sliding_step=0)
    scale = float(D) ** -0.5
    ref64, _ = oracle.paged_attention(q, kc, vc, KVH, scale, bt, sl, bs, alibi_slopes=al, kv_cache_dtype=kv,
                                      k_scale=k_scale, v_scale=v_scale, **sp)
    dev = "cuda"
    if q_strided:  # query as a slice of a fused qkv projection
        qkv = torch.zeros(S, H + 2 * KVH, D, dtype=dtype)
        qkv[:, :H] = q
        q_d = qkv.to(dev)[:, :H]
    else:
        q_d = q.to(dev)
    kc_d, vc_d, bt_d, sl_d = kc.to(dev), vc.to(dev), bt.to(dev), sl.to(dev)
    al_d = al.to(dev) if al is not None else None
    ks = torch.tensor(k_scale, dtype=torch.float32, device=dev)
    vs = torch.tensor(v_scale, dtype=torch.float32, device=dev)
    max_len = int(sl.max())
    if via_class:
        out = A.PagedAttention.forward_decode(q_d, kc_d, vc_d, bt_d, sl_d, max_len, kv, KVH, scale, al_d, ks, vs,
                                              sp["tp_rank"], sp["blocksparse_local_blocks"],
                                              sp["blocksparse_vert_stride"], sp["blocksparse_block_size"],
                                              sp["blocksparse_head_sliding_step"])
    else:
        out = torch.full((S, H, D), float("nan"), dtype=dtype, device=dev)
        if version == 1:
            A.ops.paged_attention_v1(out, q_d, kc_d, vc_d, KVH, scale, bt_d, sl_d, bs, max_len, al_d, kv, ks, vs,
                                     sp["tp_rank"], sp["blocksparse_local_blocks"], sp["blocksparse_vert_stride"],
                                     sp["blocksparse_block_size"], sp["blocksparse_head_sliding_step"])
        else:
            P = (max_len + 511) // 512
            es = torch.empty(S, H, P, dtype=torch.float32, device=dev)
            ml = torch.empty_like(es)
            tmp = torch.empty(S, H, P, D, dtype=dtype, device=dev)
            A.ops.paged_attention_v2(out, es, ml, tmp, q_d, kc_d, vc_d, KVH, scale, bt_d, sl_d, bs, max_len, al_d, kv,
                                     ks, vs, sp["tp_rank"], sp["blocksparse_local_blocks"],
                                     sp["blocksparse_vert_stride"], sp["blocksparse_block_size"],
                                     sp["blocksparse_head_sliding_step"])
    torch.cuda.synchronize()
    atol = 1e-3 if kv == "auto" else 1e-2
    return check_close(out, ref64, dtype, atol, f"S{S} H{H}/{KVH} D{D} bs{bs} {dtype} {kv} v{version}")


def ragged(S, max_len, seed):
    v = (synth.hash_uniform(S, seed, 1, max_len + 0.999)).astype(np.int64).clip(1, max_len)
    v[-1] = max_len
    return [int(t) for t in v]


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("heads", [(16, 16), (64, 8), (32, 8), (12, 12), (8, 1)])
@pytest.mark.parametrize("D", [64, 128])
def test_paged_attention_sweep(version, kv, dtype, heads, D):
    H, KVH = heads
    S = 5
    run_case(S, H, KVH, D, 16, ragged(S, 1100, D + H), dtype, kv, seed=H * 7 + D, version=version)


@pytest.mark.skipif(not FULL, reason="fast build has only head sizes 64/128 and block size 16")
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("D", [32, 80, 96, 112, 192, 256])
@pytest.mark.parametrize("bs", [8, 16, 32])
def test_paged_attention_head_and_block_sizes(kv, D, bs):
    run_case(5, 16, 4, D, bs, ragged(5, 700, D + bs), torch.bfloat16, kv, seed=D + bs, version=2)
    run_case(3, 8, 8, D, bs, ragged(3, 300, D), torch.float16, kv, seed=D + bs + 1, version=1, alibi=True)


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
def test_paged_attention_alibi_scales_strides(kv):
    run_case(4, 32, 8, 128, 16, ragged(4, 900, 3), torch.bfloat16, kv, seed=41, version=2, alibi=True,
             q_strided=True)
    if kv != "auto":
        run_case(4, 32, 8, 128, 16, ragged(4, 900, 4), torch.bfloat16, kv, seed=42, version=1, k_scale=0.05,
                 v_scale=0.02)
        run_case(2, 8, 8, 64, 16, [17, 333], torch.float32, kv, seed=43, version=1)  # f32 query, fp8 cache


def test_paged_attention_edge_lengths():
    # 1-token sequences, exact page / partition boundaries, one very long sequence next to short ones
    for kv in ("auto", "fp8"):
        run_case(6, 8, 2, 128, 16, [1, 2, 15, 16, 17, 1], torch.bfloat16, kv, seed=51, version=1)
        run_case(5, 8, 2, 128, 16, [255, 256, 257, 512, 1024], torch.float16, kv, seed=52, version=2)
        run_case(3, 32, 8, 128, 16, [1, 9000, 37], torch.bfloat16, kv, seed=53, version=2)
    run_case(1, 8, 1, 128, 16, [8192], torch.bfloat16, "fp8_e5m2", seed=54, version=2)  # C5-like: 1 kv head, split-K


def test_paged_attention_through_reference_caller_class():
    # aphrodite/attention/ops/paged_attn.py heuristic picks v1 or v2; both must agree with the oracle
    run_case(2, 32, 8, 128, 16, [300, 512], torch.bfloat16, "fp8", seed=61, version=0, via_class=True)
    run_case(2, 32, 8, 128, 16, [300, 2000], torch.bfloat16, "fp8", seed=62, version=0, via_class=True)
    run_case(20, 32, 8, 128, 16, ragged(20, 1000, 5), torch.bfloat16, "auto", seed=63, version=0, via_class=True)


@pytest.mark.parametrize("step", [0, 1, -1])
def test_paged_attention_blocksparse(step):
    sp = dict(tp_rank=1, blocksparse_local_blocks=2, blocksparse_vert_stride=4, blocksparse_block_size=64,
              blocksparse_head_sliding_step=step)
    for kv in ("auto", "fp8"):
        run_case(4, 16, 4, 128, 16, ragged(4, 1200, 7), torch.bfloat16, kv, seed=71, version=2, sparse=sp)
        run_case(3, 8, 8, 64, 16, ragged(3, 500, 8), torch.float16, kv, seed=72, version=1, sparse=sp)


def test_paged_attention_garbage_tail_is_ignored():
    # tokens beyond seq_len in the last page hold NaN/Inf bit patterns: they must not leak
    A = _A()
    S, H, KVH, D, bs = 2, 8, 2, 128, 16
    for kv, dtype in (("auto", torch.bfloat16), ("fp8", torch.bfloat16), ("fp8_e5m2", torch.float16)):
        q, kc, vc, bt, sl = make_attention_case(S, H, KVH, D, bs, [5, 21], dtype, kv, 81)
        ref64, _ = oracle.paged_attention(q, kc, vc, KVH, D ** -0.5, bt, sl, bs, kv_cache_dtype=kv)
        kc2, vc2 = kc.clone(), vc.clone()
        for s in range(S):
            L = int(sl[s])
            blk = int(bt[s, (L - 1) // bs])
            off = L % bs
            if off == 0:
                continue
            if kv == "auto":
                kc2[blk, :, :, off:, :] = float("nan")
                vc2[blk, :, :, off:] = float("inf")
            else:
                nanb = 0x7F if kv == "fp8" else 0x7E
                kc2[blk, :, :, off:, :] = nanb
                vc2[blk, :, :, off:] = nanb
        out = torch.empty(S, H, D, dtype=dtype, device="cuda")
        one = torch.tensor(1.0, device="cuda")
        A.ops.paged_attention_v1(out, q.cuda(), kc2.cuda(), vc2.cuda(), KVH, D ** -0.5, bt.cuda(), sl.cuda(), bs, 21,
                                 None, kv, one, one)
        check_close(out, ref64, dtype, 1e-3 if kv == "auto" else 1e-2, f"garbage tail {kv}")


def test_paged_attention_golden_reference_cpu():
    """Outputs of the UNMODIFIED reference CPU backend (tests/golden/attention_golden.npz)."""
    A = _A()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "attention_golden.npz"))
    for ci in range(int(g["num_cases"][0])):
        S, H, KVH, D, max_len, dtc, alibi, ver, nb = [int(v) for v in g[f"a{ci}_meta"]]
        dt = DT[dtc]
        if not FULL and D not in (64, 128):
            continue
        x = 16 // torch.tensor([], dtype=dt).element_size()
        sc = float(D) ** -0.5
        q = synth.tensor((S, H, D), 200 + ci, dt, sc).cuda()
        kc = synth.tensor((nb, KVH, D // x, 16, x), 300 + ci, dt, sc).cuda()
        vc = synth.tensor((nb, KVH, D, 16), 400 + ci, dt, sc).cuda()
        sl = torch.from_numpy(g[f"a{ci}_seq_lens"]).cuda()
        bt = torch.from_numpy(g[f"a{ci}_block_tables"]).cuda()
        al = torch.from_numpy(g[f"a{ci}_alibi"]).cuda() if alibi else None
        ref = torch.from_numpy(g[f"a{ci}_out"]).view(dt).double()
        out = torch.empty(S, H, D, dtype=dt, device="cuda")
        one = torch.tensor(1.0, device="cuda")
        A.ops.paged_attention_v1(out, q, kc, vc, KVH, sc, bt, sl, 16, max_len, al, "auto", one, one)
        # both sides are rounded to the scalar dtype: allow one extra ulp of the dtype
        check_close(out, ref, dt, 2e-3, f"golden case {ci}")


def test_paged_attention_errors():
    A = _A()
    dev = "cuda"
    q = torch.zeros(1, 8, 128, dtype=torch.bfloat16, device=dev)
    kc = torch.zeros(4, 8, 16, 16, 8, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros(4, 8, 128, 16, dtype=torch.bfloat16, device=dev)
    bt = torch.zeros(1, 4, dtype=torch.int32, device=dev)
    sl = torch.ones(1, dtype=torch.int32, device=dev)
    one = torch.tensor(1.0, device=dev)
    out = torch.empty_like(q)
    with pytest.raises(RuntimeError):
        A.ops.paged_attention_v1(out, q, kc, vc, 8, 1.0, bt, sl, 16, 1, None, "int4", one, one)
    with pytest.raises(RuntimeError):
        A.ops.paged_attention_v1(out, q, kc, vc, 8, 1.0, bt, sl, 24, 1, None, "auto", one, one)
    q120 = torch.zeros(1, 8, 120, dtype=torch.bfloat16, device=dev)
    kc8 = torch.zeros(4, 8, 120 * 16, dtype=torch.uint8, device=dev)
    with pytest.raises(RuntimeError):  # an FP8 cache stores 16 elements per chunk: head size 120 cannot be laid out
        A.ops.paged_attention_v1(torch.empty_like(q120), q120, kc8, kc8, 8, 1.0, bt, sl, 16, 1, None, "fp8", one, one)
    q100 = torch.zeros(1, 8, 100, dtype=torch.bfloat16, device=dev)
    with pytest.raises(RuntimeError):
        A.ops.paged_attention_v1(torch.empty_like(q100), q100, kc, vc, 8, 1.0, bt, sl, 16, 1, None, "auto", one, one)


def test_paged_attention_cuda_graph_capture():
    """Capturable (no sync, no cudaMalloc), padded sequences (seq_len = 1) as in
    attention/backends/utils.py:298-334."""
    A = _A()
    S, H, KVH, D, bs = 8, 32, 8, 128, 16
    lens = [700, 1, 1, 33, 2048, 1, 512, 90]
    q, kc, vc, bt, sl = make_attention_case(S, H, KVH, D, bs, lens, torch.bfloat16, "fp8", 91)
    ref64, _ = oracle.paged_attention(q, kc, vc, KVH, D ** -0.5, bt, sl, bs, kv_cache_dtype="fp8")
    qd, kcd, vcd, btd, sld = q.cuda(), kc.cuda(), vc.cuda(), bt.cuda(), sl.cuda()
    one = torch.tensor(1.0, device="cuda")
    out = torch.zeros(S, H, D, dtype=torch.bfloat16, device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        A.ops.paged_attention_v1(out, qd, kcd, vcd, KVH, D ** -0.5, btd, sld, bs, 4096, None, "fp8", one, one)
        st.synchronize()
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, stream=st):
            A.ops.paged_attention_v1(out, qd, kcd, vcd, KVH, D ** -0.5, btd, sld, bs, 4096, None, "fp8", one, one)
        out.zero_()
        gph.replay()
        st.synchronize()
    check_close(out, ref64, torch.bfloat16, 1e-2, "cuda graph replay")


def _full_size_case(S, H, KVH, D, ctx, kv, seed, ragged_tail=True):
    """Full-size problems (BASELINE configs): the FP8 cache is drawn directly as bytes (|x| < 1 in either format) with
    a seeded CPU generator - the hash generator's float64 staging would need gigabytes here."""
    bs = 16
    g = torch.Generator().manual_seed(seed)
    nb = S * ctx // bs
    q = ((torch.rand(S, H, D, generator=g) * 2 - 1) * D ** -0.5).to(torch.bfloat16)
    draw = lambda shape: (torch.randint(0, 0x38, shape, dtype=torch.uint8, generator=g)
                          + 128 * torch.randint(0, 2, shape, dtype=torch.uint8, generator=g))
    kc = draw((nb, KVH, D // 16, bs, 16))
    vc = draw((nb, KVH, D, bs))
    bt = torch.randperm(nb, generator=g).to(torch.int32).reshape(S, ctx // bs)
    lens = [ctx] * S
    if ragged_tail and S >= 4:
        lens[1], lens[2], lens[3] = 1, ctx - 17, ctx // 2 + 5
    return q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32)


@pytest.mark.parametrize("version", [1, 2])
def test_paged_attention_c2_full_geometry(version):
    """BASELINE config 2 as benchmarked: 64 sequences x 4096 context, 32 q / 8 kv heads, d 128, FP8-E4M3 KV, block 16."""
    A = _A()
    S, H, KVH, D, ctx = 64, 32, 8, 128, 4096
    q, kc, vc, bt, sl = _full_size_case(S, H, KVH, D, ctx, "fp8", 1001)
    ref64, _ = oracle.paged_attention(q, kc, vc, KVH, D ** -0.5, bt, sl, 16, kv_cache_dtype="fp8")
    dev = "cuda"
    one = torch.tensor(1.0, device=dev)
    out = torch.full((S, H, D), float("nan"), dtype=torch.bfloat16, device=dev)
    args = (q.to(dev), kc.to(dev), vc.to(dev), KVH, D ** -0.5, bt.to(dev), sl.to(dev), 16, ctx, None, "fp8", one, one)
    if version == 1:
        A.ops.paged_attention_v1(out, *args)
    else:
        P = ctx // 512
        es = torch.empty(S, H, P, dtype=torch.float32, device=dev)
        A.ops.paged_attention_v2(out, es, torch.empty_like(es), torch.empty(S, H, P, D, dtype=torch.bfloat16, device=dev), *args)
    check_close(out, ref64, torch.bfloat16, 1e-2, f"C2 full geometry v{version}")


@pytest.mark.parametrize("S", [1, 8])
def test_paged_attention_c5_long_context(S):
    """BASELINE config 5 per GPU: 8 q / 1 kv heads (Qwen2-72B at TP 8), 32 768 context, FP8-E5M2 KV."""
    A = _A()
    H, KVH, D, ctx = 8, 1, 128, 32768
    q, kc, vc, bt, sl = _full_size_case(S, H, KVH, D, ctx, "fp8_e5m2", 1005)
    ref64, _ = oracle.paged_attention(q, kc, vc, KVH, D ** -0.5, bt, sl, 16, kv_cache_dtype="fp8_e5m2")
    dev = "cuda"
    one = torch.tensor(1.0, device=dev)
    out = A.PagedAttention.forward_decode(q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), sl.to(dev), ctx, "fp8_e5m2", KVH,
                                          D ** -0.5, None, one, one)
    check_close(out, ref64, torch.bfloat16, 1e-2, f"C5 x{S}")


def test_paged_attention_tp8_shard_of_c2_and_c4():
    """One rank's shard at TP 8: 4 q / 1 kv heads (C2 sharded, bench.py --gpus 8) and 8 q / 1 kv heads (C4), 64 x 4096."""
    A = _A()
    dev = "cuda"
    one = torch.tensor(1.0, device=dev)
    for H in (4, 8):
        q, kc, vc, bt, sl = _full_size_case(64, H, 1, 128, 4096, "fp8", 1010 + H)
        ref64, _ = oracle.paged_attention(q, kc, vc, 1, 128 ** -0.5, bt, sl, 16, kv_cache_dtype="fp8")
        out = A.PagedAttention.forward_decode(q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), sl.to(dev), 4096, "fp8", 1, 128 ** -0.5,
                                              None, one, one)
        check_close(out, ref64, torch.bfloat16, 1e-2, f"TP8 shard H{H}")


@pytest.mark.skipif(not FULL, reason="fast build has only head sizes 64/128")
@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_paged_attention_head_size_120(version, dtype):
    # paged_attention_v1.cu:119-120 lists head size 120 (15 chunks of 8 16-bit elements / 30 chunks of 4 floats)
    run_case(5, 16, 4, 120, 16, ragged(5, 900, 120), dtype, "auto", seed=120, version=version)
    run_case(3, 8, 8, 120, 8, [1, 40, 333], dtype, "auto", seed=121, version=version, alibi=True)


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("D,bs", [(64, 16), (128, 16), (128, 32), (80, 8), (256, 16)])
def test_paged_attention_float32_cache(version, D, bs):
    # test_attention.py:117-326 sweeps dtype = float with kv_cache_dtype "auto": query, cache and output in float32
    if not FULL and (D not in (64, 128) or bs != 16):
        pytest.skip("fast build")
    run_case(4, 16, 4, D, bs, ragged(4, 700, D + bs), torch.float32, "auto", seed=200 + D, version=version)
    run_case(2, 8, 1, D, bs, [2100, 17], torch.float32, "auto", seed=201 + D, version=version, q_strided=True)
