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
