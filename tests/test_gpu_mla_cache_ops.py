"""MLA cache ops on the GPU (concat_and_cache_mla, copy_blocks_mla, gather_cache): bit-exact against the oracle loops
(oracle/oracle.py) and - when oracle/_ref/refcuda.so is built - the reference's own CUDA kernels. Shapes follow
/root/reference/tests/kernels/test_cache.py (kv_lora_rank 512, qk_rope_head_dim 64)."""
import pytest
import torch

import aphrodite_fork_b200 as A
from oracle import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref(name):
    r = oracle.refcuda()
    return r if r is not None and hasattr(r, name) else None


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("tokens,rank,pe,bs,nb", [(42, 512, 64, 16, 8), (1, 512, 64, 64, 2), (300, 512, 64, 16, 40), (17, 100, 13, 8, 5)])
def test_concat_and_cache_mla(tokens, rank, pe, bs, nb, dtype, kv):
    gen = torch.Generator(device=DEV).manual_seed(tokens + rank)
    kv_c = torch.randn(tokens, rank, device=DEV, generator=gen).to(dtype)
    # k_pe as a slice of a wider tensor: strided rows
    k_pe = torch.randn(tokens, pe + 3, device=DEV, generator=gen).to(dtype)[:, :pe]
    slots = torch.randperm(nb * bs, device=DEV, generator=gen)[:tokens].to(torch.int64)
    slots[tokens // 2] = -1  # padded token
    cdt = dtype if kv == "auto" else torch.uint8
    cache = torch.zeros(nb, bs, rank + pe, dtype=cdt, device=DEV)
    scale = torch.tensor(0.5 if kv != "auto" else 1.0, dtype=torch.float32, device=DEV)
    want = cache.cpu().clone()
    oracle.concat_and_cache_mla(kv_c.cpu(), k_pe.cpu().contiguous(), want, slots.cpu(), kv, scale.cpu())
    ref = _ref("concat_and_cache_mla")
    if ref is not None:
        rc = cache.clone()
        ref.concat_and_cache_mla(kv_c, k_pe, rc, slots, kv, scale)
    A.ops.concat_and_cache_mla(kv_c, k_pe, cache, slots, kv, scale)
    assert torch.equal(cache.cpu().view(torch.uint8), want.view(torch.uint8))
    if ref is not None:
        assert torch.equal(cache.view(torch.uint8), rc.view(torch.uint8))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.uint8])
def test_copy_blocks_mla(dtype):
    layers, nb, bs, entry = 5, 64, 16, 576
    gen = torch.Generator(device=DEV).manual_seed(1)
    caches = [torch.randint(0, 250, (nb, bs, entry), device=DEV, generator=gen).to(dtype) for _ in range(layers)]
    src = torch.randperm(nb // 2, generator=torch.Generator().manual_seed(2))[:7]
    mapping = torch.stack((src, src + nb // 2), dim=1).to(torch.int64)
    want = [c.cpu().clone() for c in caches]
    oracle.copy_blocks_mla(want, mapping)
    ref = _ref("copy_blocks_mla")
    if ref is not None:
        rcs = [c.clone() for c in caches]
        ref.copy_blocks_mla(rcs, mapping.to(DEV))
    A.ops.copy_blocks_mla(caches, mapping.to(DEV))
    for c, w in zip(caches, want):
        assert torch.equal(c.cpu(), w)
    if ref is not None:
        for c, r in zip(caches, rcs):
            assert torch.equal(c, r)
    A.ops.copy_blocks_mla([], mapping.to(DEV))


@pytest.mark.parametrize("with_starts", [False, True])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.uint8, torch.float32])
@pytest.mark.parametrize("batch,bs,entry", [(5, 16, 576), (130, 16, 576), (3, 64, 100), (1, 16, 7)])
def test_gather_cache(batch, bs, entry, dtype, with_starts):
    gen = torch.Generator().manual_seed(batch + entry)
    lens = torch.randint(1, 200, (batch,), generator=gen)
    lens[0] = 2 * bs  # an exact multiple of the block size
    starts = torch.randint(0, 4, (batch,), generator=gen, dtype=torch.int32) * bs if with_starts else None
    max_blocks = int((lens.max() + bs - 1) // bs) + 4
    nb = batch * max_blocks
    table = torch.randperm(nb, generator=gen).to(torch.int32).reshape(batch, max_blocks)
    cu = torch.zeros(batch + 1, dtype=torch.int32)
    cu[1:] = torch.cumsum(lens, 0)
    cache = torch.randint(0, 250, (nb, bs, entry), generator=gen).to(dtype)
    want = torch.zeros(int(cu[-1]), entry, dtype=dtype)
    oracle.gather_cache(cache, want, table, cu, batch, starts)
    dc, dt_, dcu = cache.to(DEV), table.to(DEV), cu.to(DEV)
    ds = starts.to(DEV) if starts is not None else None
    got = torch.zeros(int(cu[-1]), entry, dtype=dtype, device=DEV)
    A.ops.gather_cache(dc, got, dt_, dcu, batch, ds)
    assert torch.equal(got.cpu(), want)
    ref = _ref("gather_cache")
    if ref is not None:
        rg = torch.zeros_like(got)
        ref.gather_cache(dc, rg, dt_, dcu, batch, ds)
        assert torch.equal(got, rg)


def test_mla_errors():
    c = torch.zeros(4, 16, 576, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError):
        A.ops.concat_and_cache_mla(torch.zeros(2, 512, dtype=torch.bfloat16, device=DEV), torch.zeros(2, 32, dtype=torch.bfloat16, device=DEV),
                                   c, torch.zeros(2, dtype=torch.int64, device=DEV), "auto", torch.ones((), device=DEV))
    with pytest.raises(RuntimeError, match="block_table must be int32"):
        A.ops.gather_cache(c, torch.zeros(4, 576, dtype=torch.bfloat16, device=DEV), torch.zeros(1, 4, dtype=torch.int64, device=DEV),
                           torch.zeros(2, dtype=torch.int32, device=DEV), 1, None)
