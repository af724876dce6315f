"""gptq_marlin_repack / awq_marlin_repack on the GPU: bit-exact against the reference-Python fixtures, the oracle at
the BASELINE weight shapes, and (when built) the reference's own CUDA kernels. Mirrors
/root/reference/tests/kernels/test_marlin_gemm.py:66-176."""
import os

import numpy as np
import pytest
import torch

import aphrodite_fork_b200 as A
from oracle import marlin_format as mf
from oracle import oracle

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEV = "cuda"


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def test_repack_golden_fixtures():
    g = np.load(os.path.join(G, "marlin_repack_golden.npz"))
    empty = torch.empty(0, dtype=torch.int32, device=DEV)
    for ci in range(int(g["num_cases"][0])):
        k, n, bits, act = (int(v) for v in g[f"c{ci}_meta"])
        want = torch.from_numpy(g[f"c{ci}_marlin"])
        got = A.ops.gptq_marlin_repack(_dev(g[f"c{ci}_gptq"]), empty, k, n, bits)
        assert got.shape == want.shape and got.dtype == torch.int32
        assert torch.equal(got.cpu(), want)
        assert torch.equal(A.ops.awq_marlin_repack(_dev(g[f"c{ci}_awq"]), k, n, bits).cpu(), want)
        if act:
            got = A.ops.gptq_marlin_repack(_dev(g[f"c{ci}_gptq"]), _dev(g[f"c{ci}_sort_indices"]), k, n, bits)
            assert torch.equal(got.cpu(), torch.from_numpy(g[f"c{ci}_marlin_sorted"]))


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("act_order", [False, True])
@pytest.mark.parametrize("k,n", [(4096, 6144), (14336, 4096), (4096, 28672), (16, 64)])
def test_repack_baseline_shapes_vs_oracle_and_reference_cuda(k, n, bits, act_order):
    rng = np.random.default_rng(k + n + bits)
    gptq = rng.integers(-2**31, 2**31, (k // (32 // bits), n), dtype=np.int64).astype(np.int32)
    awq = rng.integers(-2**31, 2**31, (k, n // (32 // bits)), dtype=np.int64).astype(np.int32)
    perm = rng.permutation(k).astype(np.int32) if act_order else np.zeros(0, dtype=np.int32)
    gd, ad, pd = _dev(gptq), _dev(awq), _dev(perm)
    got_g = A.ops.gptq_marlin_repack(gd, pd, k, n, bits)
    got_a = A.ops.awq_marlin_repack(ad, k, n, bits)
    assert torch.equal(got_g.cpu(), torch.from_numpy(mf.gptq_marlin_repack(gptq, perm, k, n, bits)))
    assert torch.equal(got_a.cpu(), torch.from_numpy(mf.awq_marlin_repack(awq, k, n, bits)))
    ref = oracle.refcuda()
    if ref is not None and hasattr(ref, "gptq_marlin_repack"):
        assert torch.equal(got_g, ref.gptq_marlin_repack(gd, pd, k, n, bits))
        assert torch.equal(got_a, ref.awq_marlin_repack(ad, k, n, bits))


def test_repack_then_gemm_round_trip():
    """Checkpoint -> repack -> gptq_marlin_gemm equals the GEMM on weights packed by the oracle."""
    k, n, m = 1024, 512, 8
    g = torch.Generator().manual_seed(3)
    w = torch.randn(k, n, generator=g).to(torch.bfloat16)
    w_ref, bq, s = mf.marlin_quantize(w, 4, 128)[:3]
    q = mf.unpack_weights(bq.numpy(), k, n, 4)
    empty = torch.empty(0, dtype=torch.int32, device=DEV)
    b1 = A.ops.gptq_marlin_repack(_dev(mf.gptq_pack(q, 4)), empty, k, n, 4)
    b2 = A.ops.awq_marlin_repack(_dev(mf.awq_pack(q, 4)), k, n, 4)
    assert torch.equal(b1.cpu(), bq) and torch.equal(b2.cpu(), bq)
    a = torch.randn(m, k, generator=g).to(torch.bfloat16).to(DEV)
    ws = torch.zeros(148 * 64, dtype=torch.int32, device=DEV)
    c = A.ops.gptq_marlin_gemm(a, b1, s.to(DEV), empty, empty, empty, ws, mf.UINT4B8, m, n, k, True, False, False, True, False)
    want = a.float().cpu() @ w_ref.float()
    assert float((c.float().cpu() - want).abs().max()) <= 2e-2 * float(want.abs().max())


def test_repack_errors():
    empty = torch.empty(0, dtype=torch.int32, device=DEV)
    w = torch.zeros(16, 64, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError, match="num_bits must be 4 or 8"):
        A.ops.gptq_marlin_repack(w, empty, 128, 64, 3)
    with pytest.raises(RuntimeError, match="not divisible by tile_k_size"):
        A.ops.gptq_marlin_repack(w, empty, 120, 64, 4)
    with pytest.raises(RuntimeError, match="not divisible by tile_n_size"):
        A.ops.awq_marlin_repack(w, 16, 96, 4)
    with pytest.raises(RuntimeError, match="Shape mismatch"):
        A.ops.gptq_marlin_repack(w, empty, 256, 64, 4)
    with pytest.raises(RuntimeError, match="not on GPU"):
        A.ops.gptq_marlin_repack(w.cpu(), empty, 128, 64, 4)


@pytest.mark.parametrize("m", [1, 16, 64, 300])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_act_order_gemm_full_k(m, dtype):
    """desc_act checkpoints (tests/kernels/test_marlin_gemm.py:189-260 with act_order=True, is_k_full=True): rows of the
    weight carry a shuffled group index; repack with sort_indices, GEMM with (g_idx, sort_indices)."""
    k, n, gs = 2048, 512, 128
    g = torch.Generator().manual_seed(7 + m)
    w = torch.randn(k, n, generator=g).to(dtype)
    w_ref, bq, s = mf.marlin_quantize(w, 4, gs)[:3]
    q = mf.unpack_weights(bq.numpy(), k, n, 4)
    rand_perm = torch.randperm(k, generator=g)
    g_idx = (torch.arange(k) // gs)[rand_perm].to(torch.int32)       # quant_utils.permute_rows
    w_shuffled = w_ref[rand_perm]
    checkpoint = mf.gptq_pack(q[rand_perm.numpy()], 4)
    sort_indices = torch.argsort(g_idx, stable=True).to(torch.int32)   # quant_utils.sort_weights
    g_idx_sorted = g_idx[sort_indices.long()].contiguous()
    b = A.ops.gptq_marlin_repack(_dev(checkpoint), sort_indices.to(DEV), k, n, 4)
    a = torch.randn(m, k, generator=g).to(dtype).to(DEV)
    ws = torch.zeros(148 * 64, dtype=torch.int32, device=DEV)
    empty = torch.empty(0, dtype=torch.int32, device=DEV)
    args = (a, b, s.to(DEV), empty, g_idx_sorted.to(DEV), sort_indices.to(DEV), ws, mf.UINT4B8, m, n, k, True, False, False, True, False)
    c = A.ops.gptq_marlin_gemm(*args)
    want = a.float().cpu().double() @ w_shuffled.double()
    assert float((c.cpu().double() - want).abs().max()) <= 2e-2 * float(want.abs().max())
    assert int(ws.abs().max()) == 0
    ref = oracle.refcuda()
    if ref is not None:
        rc = ref.gptq_marlin_gemm(*args)
        assert float((c.float() - rc.float()).abs().max()) <= 1e-2 * float(rc.float().abs().max())
    # the same call with is_k_full = False (what a caller passes for any row-parallel shard) takes the act-order kernel
    c2 = A.ops.gptq_marlin_gemm(*args[:11], False, *args[12:])
    assert float((c2.cpu().double() - want).abs().max()) <= 2e-2 * float(want.abs().max())


@pytest.mark.parametrize("m", [1, 16, 33, 300])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("tp,rank", [(2, 1), (4, 2)])
def test_act_order_gemm_row_parallel_shard(m, dtype, tp, rank):
    """The row-parallel shard of a desc_act checkpoint (is_k_full = False; marlin_utils.py:176-185, gptq_marlin.py): the
    rank holds K / tp shuffled rows with THEIR slice of g_idx and the scale table of the whole weight; rows are sorted
    by group inside the shard (marlin_sort_g_idx), groups start and end anywhere. Truth = a_shard @ (those rows of w_ref);
    also compared with the reference CUDA kernel when it is built (gptq_marlin.cu:921-985,1280-1300)."""
    k, n, gs = 2048, 512, 128
    g = torch.Generator().manual_seed(100 + m + tp)
    w = torch.randn(k, n, generator=g).to(dtype)
    w_ref, bq, s = mf.marlin_quantize(w, 4, gs)[:3]       # s: Marlin-permuted scales of the whole weight
    q = mf.unpack_weights(bq.numpy(), k, n, 4)
    rand_perm = torch.randperm(k, generator=g)
    g_idx = (torch.arange(k) // gs)[rand_perm].to(torch.int32)
    kl = k // tp
    rows = slice(rank * kl, (rank + 1) * kl)
    g_idx_l = g_idx[rows]
    w_l = w_ref[rand_perm][rows]
    checkpoint_l = mf.gptq_pack(q[rand_perm.numpy()][rows], 4)
    sort_indices = torch.argsort(g_idx_l, stable=True).to(torch.int32)
    g_idx_sorted = g_idx_l[sort_indices.long()].contiguous()
    b = A.ops.gptq_marlin_repack(_dev(checkpoint_l), sort_indices.to(DEV), kl, n, 4)
    a = torch.randn(m, kl, generator=g).to(dtype).to(DEV)
    ws = torch.zeros(148 * 64, dtype=torch.int32, device=DEV)
    empty = torch.empty(0, dtype=torch.int32, device=DEV)
    args = (a, b, s.to(DEV), empty, g_idx_sorted.to(DEV), sort_indices.to(DEV), ws, mf.UINT4B8, m, n, kl, False, False, False, True, False)
    c = A.ops.gptq_marlin_gemm(*args)
    want = a.float().cpu().double() @ w_l.double()
    err = (c.cpu().double() - want).abs()
    assert float(err.mean() / want.abs().mean()) < 5e-3
    assert float(err.max()) <= (1e-2 if dtype == torch.bfloat16 else 2e-3) * float(want.abs().max())
    assert int(ws.abs().max()) == 0
    ref = oracle.refcuda()
    if ref is not None:
        rc = ref.gptq_marlin_gemm(*args)
        assert float((c.float() - rc.float()).abs().max()) <= 1e-2 * float(rc.float().abs().max())


@pytest.mark.parametrize("fmt,bits,gs,act_order", [("gptq", 4, 128, False), ("gptq", 8, -1, False), ("gptq", 4, 64, True),
                                                  ("awq", 4, 128, False), ("awq", 8, 64, False)])
def test_checkpoint_to_linear_through_the_host_helpers(fmt, bits, gs, act_order):
    """What the reference's GPTQMarlinLinearMethod / AWQMarlinLinearMethod do at load time
    (quantization/gptq_marlin.py process_weights_after_loading, awq_marlin.py), with this library's ops and its
    marlin_utils on DEVICE tensors: checkpoint layout -> repack, permute scales, lay out zero points, sort g_idx ->
    apply_*_marlin_linear. Checked against x @ w_ref of the oracle's quantiser (test_marlin_gemm.py:254: max-rel 4e-2)."""
    from aphrodite_fork_b200 import marlin_utils as mu
    k, n = 1024, 768
    g = torch.Generator().manual_seed(bits + k + (gs if gs > 0 else 7))
    w = (torch.randn(k, n, generator=g) * 0.05).to(torch.bfloat16)
    x = torch.randn(3, 5, k, generator=g).to(torch.bfloat16).to(DEV)  # any leading shape
    groups = 1 if gs == -1 else k // gs
    if fmt == "gptq":
        if act_order:  # random group membership: rows are quantised with the scale of g_idx[row]
            g_idx = torch.randint(0, groups, (k,), generator=g, dtype=torch.int32)
            order = torch.argsort(g_idx, stable=True)
            w_sorted_ref, q_sorted, s, _ = mf.quantize(w[order], bits, gs, False)
            inv = torch.argsort(order)
            w_ref, q = w_sorted_ref[inv], q_sorted[inv]
        else:
            w_ref, q, s, _ = mf.quantize(w, bits, gs, False)
        qweight = _dev(mf.gptq_pack(q.numpy().astype(np.uint32), bits))       # checkpoint: packed along K
        scales = s.to(DEV)
        qtype = mu.scalar_types.uint4b8 if bits == 4 else mu.scalar_types.uint8b128
        mu.verify_marlin_supported(qtype, gs, has_zp=False)
        mu.verify_marlin_supports_shape(n, k, k, gs)
        if act_order:
            g_sorted, sort_idx = mu.marlin_sort_g_idx(g_idx.to(DEV))
        else:
            g_sorted, sort_idx = mu.marlin_make_empty_g_idx(DEV), mu.marlin_make_empty_g_idx(DEV)
        b = A.ops.gptq_marlin_repack(qweight, sort_idx, k, n, bits)
        sm = mu.marlin_permute_scales(scales, k, n, gs)
        out = mu.apply_gptq_marlin_linear(x, b, sm, mu.marlin_make_empty_zp(DEV), g_sorted, sort_idx, mu.marlin_make_workspace(n, DEV),
                                          qtype, n, k, has_zp=False, is_k_full=mu.marlin_is_k_full(act_order, False))
    else:
        w_ref, q, s, zp = mf.quantize(w, bits, gs, True)
        qweight = _dev(mf.awq_pack(q.numpy().astype(np.uint32), bits))        # checkpoint: packed along N, AWQ's interleave
        inter = [0, 2, 4, 6, 1, 3, 5, 7] if bits == 4 else [0, 2, 1, 3]
        qzeros = mu.pack_cols(zp.reshape(-1, len(inter))[:, inter].reshape(groups, n).contiguous().to(DEV), bits, groups, n)
        qtype = mu.scalar_types.uint4 if bits == 4 else mu.scalar_types.uint8
        mu.verify_marlin_supported(qtype, gs, has_zp=True)
        b = A.ops.awq_marlin_repack(qweight, k, n, bits)
        sm = mu.marlin_permute_scales(s.to(DEV), k, n, gs)
        zm = mu.awq_to_marlin_zero_points(qzeros, groups, n, bits)
        assert zm.is_cuda and sm.is_cuda
        e = mu.marlin_make_empty_g_idx(DEV)
        out = mu.apply_awq_marlin_linear(x, b, sm, zm, e, e, mu.marlin_make_workspace(n, DEV), qtype, n, k,
                                         bias=torch.ones(n, dtype=torch.bfloat16, device=DEV))
        out = out - 1
    assert out.shape == (3, 5, n)
    want = x.cpu().float() @ w_ref.float()
    err = (out.cpu().float() - want).abs().mean() / want.abs().mean()
    assert float(err) < 4e-2, float(err)
