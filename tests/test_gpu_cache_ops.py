"""GPU parity of the block-manager side of kernels/cache_kernels.cu: reshape_and_cache_flash,
copy_blocks, swap_blocks through the registered torch ops (-> C ABI) against the CPU oracle and, when
oracle/_ref/refcuda.so is present, against the UNMODIFIED reference CUDA kernels. Bit-exact.
Mirrors /root/reference/tests/kernels/test_cache.py:32-109 (copy_blocks), :209-318 (reshape_and_cache_flash),
:321-397 (swap_blocks)."""
import pytest
import torch

import synth
from oracle import oracle

pytestmark = pytest.mark.gpu


def _ops():
    import aphrodite_fork_b200 as A
    return A.ops


def _bits(t):
    return t.contiguous().view(torch.uint8)


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(64, 8, 128, 16), (37, 12, 64, 16), (19, 4, 80, 32), (23, 2, 256, 8), (500, 8, 120, 16)])
def test_reshape_and_cache_flash_bit_exact(kv, dtype, shape):
    ops = _ops()
    T, H, D, bs = shape
    dev = "cuda"
    nb = (T + bs - 1) // bs + 3
    cdt = torch.uint8 if kv != "auto" else dtype
    qkv = synth.tensor((T, 3, H, D), T + D, dtype, 3.0)  # strided slices of a fused qkv
    key, value = qkv[:, 1], qkv[:, 2]
    slots = torch.from_numpy(synth.permutation(nb * bs, T + 1)[:T].copy()).to(torch.int64)
    slots[::13] = -1
    ks, vs = (0.37, 0.011) if kv != "auto" else (1.0, 1.0)
    kc = torch.full((nb, bs, H, D), 7, dtype=cdt)
    vc = torch.full((nb, bs, H, D), 9, dtype=cdt)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    oracle.reshape_and_cache_flash(key, value, kc_ref, vc_ref, slots, kv, ks, vs)
    qd = qkv.to(dev)
    kcd, vcd = kc.to(dev), vc.to(dev)
    ksd, vsd = torch.tensor(ks, device=dev), torch.tensor(vs, device=dev)
    ops.reshape_and_cache_flash(qd[:, 1], qd[:, 2], kcd, vcd, slots.to(dev), kv, ksd, vsd)
    torch.cuda.synchronize()
    assert torch.equal(_bits(kcd.cpu()), _bits(kc_ref)), "key cache differs from the oracle"
    assert torch.equal(_bits(vcd.cpu()), _bits(vc_ref)), "value cache differs from the oracle"
    ref = oracle.refcuda()
    if ref is not None:
        kcr, vcr = kc.to(dev), vc.to(dev)
        ref.reshape_and_cache_flash(qd[:, 1], qd[:, 2], kcr, vcr, slots.to(dev), kv, ksd, vsd)
        assert torch.equal(_bits(kcr), _bits(kcd)) and torch.equal(_bits(vcr), _bits(vcd)), "differs from the reference CUDA kernel"


def test_reshape_and_cache_flash_padded_cache_and_empty():
    # a cache whose blocks are padded (stride(0) > block elements) and an empty batch
    ops = _ops()
    dev = "cuda"
    T, H, D, bs, nb = 21, 4, 64, 16, 5
    big = torch.zeros(nb, bs * H * D + 64, dtype=torch.bfloat16, device=dev)
    kc = big[:, : bs * H * D].view(nb, bs, H, D)
    vbig = torch.zeros_like(big)
    vc = vbig[:, : bs * H * D].view(nb, bs, H, D)
    key = synth.tensor((T, H, D), 3, torch.bfloat16, 1.0)
    value = synth.tensor((T, H, D), 4, torch.bfloat16, 1.0)
    slots = torch.arange(7, 7 + T, dtype=torch.int64)
    one = torch.tensor(1.0, device=dev)
    ops.reshape_and_cache_flash(key.to(dev), value.to(dev), kc, vc, slots.to(dev), "auto", one, one)
    kc_ref, vc_ref = torch.zeros(nb, bs, H, D, dtype=torch.bfloat16), torch.zeros(nb, bs, H, D, dtype=torch.bfloat16)
    oracle.reshape_and_cache_flash(key, value, kc_ref, vc_ref, slots, "auto", 1.0, 1.0)
    assert torch.equal(_bits(kc.cpu()), _bits(kc_ref)) and torch.equal(_bits(vc.cpu()), _bits(vc_ref))
    assert int(big[:, bs * H * D:].abs().sum()) == 0, "padding was written"
    e = torch.empty(0, dtype=torch.int64, device=dev)
    ops.reshape_and_cache_flash(key.to(dev)[:0], value.to(dev)[:0], kc, vc, e, "auto", one, one)


@pytest.mark.parametrize("cdt", [torch.bfloat16, torch.float16, torch.float32, torch.uint8])
@pytest.mark.parametrize("layers,pairs", [(1, 1), (3, 7), (32, 40)])
def test_copy_blocks_bit_exact(cdt, layers, pairs):
    ops = _ops()
    dev = "cuda"
    nb, H, D, bs = 96, 4, 64, 16
    x = 16 // torch.tensor([], dtype=cdt).element_size()

    def rnd(shape, seed):
        if cdt == torch.uint8:
            return torch.from_numpy(synth.hash_uniform(int(torch.tensor(shape).prod()), seed, 0, 255).astype("uint8")).reshape(shape)
        return synth.tensor(shape, seed, cdt, 1.0)

    kcs = [rnd((nb, H, D // x, bs, x), 10 + l) for l in range(layers)]
    vcs = [rnd((nb, H, D, bs), 50 + l) for l in range(layers)]
    perm = synth.permutation(nb, layers + pairs)
    src = perm[: nb // 2]
    dst = perm[nb // 2:]
    mapping = torch.tensor([[int(src[i % len(src)]), int(dst[i])] for i in range(pairs)], dtype=torch.int64)
    kr, vr = [t.clone() for t in kcs], [t.clone() for t in vcs]
    oracle.copy_blocks(kr, vr, mapping)
    kd, vd = [t.to(dev) for t in kcs], [t.to(dev) for t in vcs]
    ops.copy_blocks(kd, vd, mapping.to(dev))
    torch.cuda.synchronize()
    for l in range(layers):
        assert torch.equal(_bits(kd[l].cpu()), _bits(kr[l])) and torch.equal(_bits(vd[l].cpu()), _bits(vr[l])), f"layer {l}"
    ref = oracle.refcuda()
    if ref is not None:
        k2, v2 = [t.to(dev) for t in kcs], [t.to(dev) for t in vcs]
        ref.copy_blocks(k2, v2, mapping.to(dev))
        for l in range(layers):
            assert torch.equal(_bits(k2[l]), _bits(kd[l])) and torch.equal(_bits(v2[l]), _bits(vd[l]))
    ops.copy_blocks([], [], mapping.to(dev))  # no layers: no-op


@pytest.mark.parametrize("direction", ["d2d", "d2h", "h2d"])
@pytest.mark.parametrize("cdt", [torch.bfloat16, torch.uint8])
def test_swap_blocks_bit_exact(direction, cdt):
    ops = _ops()
    nb, H, D, bs = 40, 8, 128, 16
    shape = (nb, H, D, bs)
    if cdt == torch.uint8:
        src = torch.from_numpy(synth.hash_uniform(nb * H * D * bs, 5, 0, 255).astype("uint8")).reshape(shape)
        dst = torch.from_numpy(synth.hash_uniform(nb * H * D * bs, 6, 0, 255).astype("uint8")).reshape(shape)
    else:
        src, dst = synth.tensor(shape, 5, cdt, 1.0), synth.tensor(shape, 6, cdt, 1.0)
    # scattered pairs plus a run of consecutive blocks on both sides (merged into one copy by the C ABI)
    pairs = [[3, 17], [9, 2], [20, 30], [21, 31], [22, 32], [23, 33], [0, 39]]
    mapping = torch.tensor(pairs, dtype=torch.int64)
    want = dst.clone()
    oracle.swap_blocks(src, want, mapping)
    sd = "cuda" if direction in ("d2d", "d2h") else "cpu"
    dd = "cuda" if direction in ("d2d", "h2d") else "cpu"
    s_t = src.to(sd) if sd == "cuda" else src.clone().pin_memory()
    d_t = dst.to(dd) if dd == "cuda" else dst.clone().pin_memory()
    ops.swap_blocks(s_t, d_t, mapping)
    torch.cuda.synchronize()
    assert torch.equal(_bits(d_t.cpu()), _bits(want))
    with pytest.raises(RuntimeError):
        ops.swap_blocks(s_t, d_t, mapping.cuda())  # block_mapping must be on CPU (cache_kernels.cu:46)


def test_swap_blocks_inside_one_tensor_with_overlapping_runs():
    """src and dst are the SAME cache and consecutive pairs shift blocks by one (0 -> 1, 1 -> 2, 2 -> 3): the reference
    copies block by block in list order (cache_kernels.cu:55-66), so block 0 cascades; a merged run must not change that."""
    ops = _ops()
    cache = synth.tensor((12, 4, 64, 16), 9, torch.bfloat16, 1.0)
    mapping = torch.tensor([[0, 1], [1, 2], [2, 3], [8, 5], [9, 6]], dtype=torch.int64)
    want = cache.clone()
    for s_, d_ in mapping.tolist():
        want[d_] = want[s_]
    dev = cache.cuda()
    ops.swap_blocks(dev, dev, mapping)
    torch.cuda.synchronize()
    assert torch.equal(_bits(dev.cpu()), _bits(want))


def test_paged_attention_class_swap_and_copy_blocks():
    """The reference's PagedAttention.swap_blocks / copy_blocks (attention/ops/paged_attn.py:233-254) take whole
    [2, num_blocks, ...] caches: K plane, then V plane, through the same ops."""
    import aphrodite_fork_b200 as A
    nb, kvh, d, bs = 24, 4, 64, 16
    shape = A.PagedAttention.get_kv_cache_shape(nb, bs, kvh, d)
    assert shape == (2, nb, bs * kvh * d)
    src = synth.tensor(shape, 21, torch.bfloat16, 1.0)
    dst = synth.tensor(shape, 22, torch.bfloat16, 1.0)
    mapping = torch.tensor([[1, 5], [2, 6], [9, 0]], dtype=torch.int64)
    want = dst.clone()
    for s_, d_ in mapping.tolist():
        want[:, d_] = src[:, s_]
    d_dev = dst.cuda()
    A.PagedAttention.swap_blocks(src.cuda(), d_dev, mapping)
    torch.cuda.synchronize()
    assert torch.equal(_bits(d_dev.cpu()), _bits(want))
    # copy_blocks: the same (src, dst) pairs applied inside every layer's cache
    layers = [synth.tensor(shape, 30 + i, torch.bfloat16, 1.0) for i in range(3)]
    wants = [l.clone() for l in layers]
    pairs = torch.tensor([[0, 7], [0, 8], [3, 4]], dtype=torch.int64)
    for w in wants:
        for s_, d_ in pairs.tolist():
            w[:, d_] = w[:, s_]
    devs = [l.cuda() for l in layers]
    A.PagedAttention.copy_blocks(devs, pairs.cuda())
    torch.cuda.synchronize()
    for got, w in zip(devs, wants):
        assert torch.equal(_bits(got.cpu()), _bits(w))
    k, v = A.PagedAttention.split_kv_cache(devs[0], kvh, d)
    assert k.shape == (nb, kvh, d // 8, bs, 8) and v.shape == (nb, kvh, d, bs)
    with pytest.raises(NotImplementedError):
        A.PagedAttention.forward_prefix()
    assert 120 in A.PagedAttention.get_supported_head_sizes()
