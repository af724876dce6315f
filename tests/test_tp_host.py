"""CPU tests of the tensor-parallel host logic: shape rules, head split, slicing of Marlin-format
weights, and the all_reduce exchange over a world_size-2 gloo group. The oracle (CPU restatement of the format and of
the GEMM) is the checker; the product code under test is aphrodite_fork_b200.tp only - no CUDA involved."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import aphrodite_fork_b200.tp as tp
from oracle import marlin_format as mf
import synth


def test_verify_marlin_supports_shape_messages():
    # marlin_utils.py:96-122: same three failure classes, same wording of the leading clause
    tp.verify_marlin_supports_shape(1280, 8192, 8192, 128)          # C4 TP8 qkv
    tp.verify_marlin_supports_shape(8192, 1024, 8192, 128)          # C4 TP8 o
    with pytest.raises(ValueError, match="output_size_per_partition = 96 is not divisible by  min_thread_n = 64"):
        tp.verify_marlin_supports_shape(96, 1024, 1024, 128)
    with pytest.raises(ValueError, match="input_size_per_partition = 192 is not divisible by min_thread_k = 128"):
        tp.verify_marlin_supports_shape(128, 192, 384, 64)
    with pytest.raises(ValueError, match="input_size_per_partition = 1408 is not divisible by group_size = 256"):
        tp.verify_marlin_supports_shape(128, 1408, 29568, 256)
    ok, msg = tp.check_marlin_supports_shape(128, 3696, 29568, 128)   # Qwen2-72B inter / 8
    assert not ok and "3696" in msg


def test_flags_and_workspace():
    assert tp.marlin_is_k_full(False, True) and tp.marlin_is_k_full(True, False) and not tp.marlin_is_k_full(True, True)
    assert tp.marlin_repeat_scales_on_all_ranks(False, -1, True)
    assert not tp.marlin_repeat_scales_on_all_ranks(False, 128, True)
    assert not tp.marlin_repeat_scales_on_all_ranks(False, -1, False)
    assert tp.marlin_repeat_scales_on_all_ranks(True, 128, False)
    ws = tp.marlin_make_workspace(4096, "cpu")
    assert ws.shape == (4096 // 64 * 16,) and ws.dtype == torch.int32 and int(ws.abs().sum()) == 0


def test_shard_heads():
    assert tp.shard_heads(64, 8, 8) == (8, 1)      # C4 TP8
    assert tp.shard_heads(64, 8, 2) == (32, 4)
    assert tp.shard_heads(32, 8, 1) == (32, 8)
    assert tp.shard_heads(64, 8, 16) == (4, 1)     # kv heads replicated (llama.py:118-127)
    with pytest.raises(ValueError):
        tp.shard_heads(30, 8, 4)
    with pytest.raises(ValueError):
        tp.shard_heads(64, 8, 3)


def _quantized(K, N, gs, zp, seed):
    w = synth.tensor((K, N), seed, torch.bfloat16, 1.0)
    return mf.marlin_quantize(w, 4, gs, zp)   # (w_ref, q_weight, scales, zeros)


@pytest.mark.parametrize("gs,zp", [(128, False), (64, True), (-1, False), (32, False)])
@pytest.mark.parametrize("tp_size", [2, 4])
def test_shard_marlin_weight_matches_requantised_slices(gs, zp, tp_size):
    """A slice of the packed tensors decodes to the same weights as the corresponding slice of the dense matrix."""
    K, N = 1024, 512
    w_ref, qw, sc, bz = _quantized(K, N, gs, zp, 3)
    bz_t = bz if zp else None
    for rank in range(tp_size):
        col = tp.shard_marlin_weight(qw, sc, bz_t, K, N, gs, tp_size, rank, "column")
        n_l = N // tp_size
        got = mf.dequantize(col.q_weight, col.scales, col.zeros, K, n_l, 4, gs, zp)
        assert torch.equal(got, w_ref[:, rank * n_l:(rank + 1) * n_l])
        row = tp.shard_marlin_weight(qw, sc, bz_t, K, N, gs, tp_size, rank, "row")
        k_l = K // tp_size
        got = mf.dequantize(row.q_weight, row.scales, row.zeros, k_l, N, 4, gs, zp)
        assert torch.equal(got, w_ref[rank * k_l:(rank + 1) * k_l])
        assert row.is_row_parallel and not col.is_row_parallel
        assert row.scales.shape[0] == (1 if gs == -1 else k_l // gs)


def test_shard_marlin_weight_rejects_bad_splits():
    K, N = 512, 192
    w_ref, qw, sc, bz = _quantized(K, N, 128, False, 5)
    with pytest.raises(ValueError, match="min_thread_n"):
        tp.shard_marlin_weight(qw, sc, None, K, N, 128, 2, 0, "column")   # 96 columns per rank
    with pytest.raises(ValueError, match="min_thread_k|group_size"):
        tp.shard_marlin_weight(qw, sc, None, K, N, 128, 8, 0, "row")      # 64 rows per rank
    with pytest.raises(ValueError):
        tp.shard_marlin_weight(qw, sc, None, K, N, 128, 2, 2, "row")
    with pytest.raises(ValueError):
        tp.shard_marlin_weight(qw, sc, None, K, N, 128, 2, 0, "diagonal")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tp_worker(rank, world, port, result_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        K, N, M, gs = 1024, 256, 5, 128
        w_ref, qw, sc, bz = _quantized(K, N, gs, False, 9)
        a = synth.tensor((M, K), 10, torch.bfloat16, 1.0)
        grp = tp.TPGroup()
        assert grp.world_size == world and grp.rank == rank
        # column parallel: each rank owns N / world output columns, no exchange; gathered here only to check
        col = tp.shard_marlin_weight(qw, sc, None, K, N, gs, world, rank, "column")
        w_col = mf.dequantize(col.q_weight, col.scales, None, K, N // world, 4, gs, False)
        y_col = a.float() @ w_col.float()
        parts = [torch.empty_like(y_col) for _ in range(world)]
        dist.all_gather(parts, y_col)
        # row parallel: each rank owns K / world input rows and produces a full-width partial -> all_reduce
        row = tp.shard_marlin_weight(qw, sc, None, K, N, gs, world, rank, "row")
        k_l = K // world
        w_row = mf.dequantize(row.q_weight, row.scales, None, k_l, N, 4, gs, False)
        y_row = grp.all_reduce(a[:, rank * k_l:(rank + 1) * k_l].float() @ w_row.float())
        full = a.float() @ w_ref.float()
        ok_col = torch.allclose(torch.cat(parts, dim=1), full, rtol=1e-5, atol=1e-4)
        ok_row = torch.allclose(y_row, full, rtol=1e-4, atol=1e-3)
        # the reference-style communicator on a CPU group of 2: stays disabled (callers then use the group's all_reduce),
        # exactly like the reference without its custom all-reduce library
        from aphrodite_fork_b200.custom_all_reduce import CustomAllreduce
        ca = CustomAllreduce(dist.group.WORLD, "cpu")
        ok_row = ok_row and ca.disabled and ca.disabled_reason == "not a CUDA device" and ca.custom_all_reduce(y_row) is None
        result_q.put((rank, ok_col, ok_row))
    finally:
        dist.destroy_process_group()


def test_row_and_column_parallel_over_gloo_world_size_2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_tp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    got = sorted(q.get(timeout=10) for _ in range(2))
    assert got == [(0, True, True), (1, True, True)]


def test_tp_group_is_a_noop_without_a_process_group():
    g = tp.TPGroup()
    x = torch.ones(3)
    assert g.world_size == 1 and g.rank == 0 and g.all_reduce(x) is x and torch.equal(x, torch.ones(3))


def test_one_shot_all_reduce_eligibility_rules():
    """Host logic of custom_all_reduce.OneShotAllReduce.should_use / should_fuse_norm (no GPU, no peers: the object is
    built without its buffers): size limits per world size, alignment, contiguity, dtypes."""
    from aphrodite_fork_b200.custom_all_reduce import OneShotAllReduce

    class FakeCuda(torch.Tensor):
        """A CPU tensor that reports is_cuda (the rules only look at metadata)."""
        @property
        def is_cuda(self):
            return True

    def fake(*shape, dtype=torch.bfloat16):
        return torch.zeros(*shape, dtype=dtype).as_subclass(FakeCuda)

    os.environ.pop("APH_AR_LIMIT_BYTES", None)
    ar = object.__new__(OneShotAllReduce)
    ar.max_size = 8 * 1024 * 1024
    os.environ.pop("APH_AR_TWO_SHOT_BYTES", None)
    for world in (2, 4, 8):  # rows of 8192 bf16 = 16 KiB each: the peer-memory path holds up to max_size at every world size
        ar.world_size = world
        assert ar.should_use(fake(512, 8192)), world
        assert not ar.should_use(fake(513, 8192)), world
    # one-shot below, two-shot from a size on (larger groups pay (world - 1) x the message for a one-shot push)
    for world, one_shot_rows, two_shot_rows in ((2, 512, None), (4, 31, 32), (8, 31, 32)):
        ar.world_size = world
        assert not ar.two_shot(one_shot_rows * 16384), world
        if two_shot_rows is not None:
            assert ar.two_shot(two_shot_rows * 16384), world
    ar.world_size = 2
    assert not ar.should_use(fake(3, 4))                      # 24 bytes: not a multiple of 16
    assert not ar.should_use(fake(64, 128).t())               # not contiguous
    assert not ar.should_use(fake(64, 128, dtype=torch.int32))
    assert not ar.should_use(torch.zeros(64, 128, dtype=torch.bfloat16))  # a CPU tensor
    x, w = fake(64, 8192), fake(8192)
    assert ar.should_fuse_norm(x, fake(64, 8192), w)
    assert not ar.should_fuse_norm(fake(64, 8192, dtype=torch.float32), fake(64, 8192, dtype=torch.float32), fake(8192, dtype=torch.float32))
    assert not ar.should_fuse_norm(fake(4, 32768), fake(4, 32768), fake(32768))   # hidden size above 16384
    assert not ar.should_fuse_norm(x, fake(63, 8192), w)
    os.environ["APH_AR_LIMIT_BYTES"] = "0"
    try:
        assert not ar.should_use(fake(1, 4096))
    finally:
        del os.environ["APH_AR_LIMIT_BYTES"]


def test_all_reduce_mode_policy_and_the_reference_style_communicator():
    """Host policy of custom_all_reduce.py: LL lines up to 1 MiB (512 KiB at two ranks), two-shot above, environment
    overrides; and CustomAllreduce, the adapter with the reference's interface (custom_all_reduce.py:45-304): `disabled`
    instead of exceptions, custom_all_reduce() -> None when the message is not served."""
    from aphrodite_fork_b200.custom_all_reduce import CustomAllreduce, OneShotAllReduce, is_weak_contiguous
    for k in ("APH_AR_MODE", "APH_AR_TWO_SHOT_BYTES", "APH_AR_LIMIT_BYTES"):
        os.environ.pop(k, None)
    ar = object.__new__(OneShotAllReduce)
    ar.max_size, ar.multicast_ptr, ar.ll_max_bytes, ar.ll_max_passes = 8 << 20, 0, 2 << 20, 512
    ar.world_size = 8
    assert ar.mode(16 << 10) == "ll" and ar.mode(1 << 20) == "ll" and ar.mode((1 << 20) + 16) == "two_shot"
    assert ar.mode(1 << 20, rows=64, hidden=8192) == "ll"            # 64 rows x 2 passes of 512 vectors
    assert ar.mode(1 << 20, rows=4096, hidden=128) == "two_shot"     # 4096 passes do not fit the 512 line rows
    ar.world_size = 2
    assert ar.mode(512 << 10) == "ll" and ar.mode(1 << 20) == "one_shot"
    os.environ["APH_AR_MODE"] = "multimem"
    assert ar.mode(1 << 20) == "one_shot"                            # no multicast mapping: the request is ignored
    ar.multicast_ptr = 0x1000
    assert ar.mode(1 << 20) == "multimem"
    os.environ["APH_AR_MODE"] = "two_shot"
    assert ar.mode(16) == "two_shot"
    os.environ.pop("APH_AR_MODE")
    os.environ["APH_AR_TWO_SHOT_BYTES"] = "0"
    assert ar.mode(4 << 20) == "one_shot"
    os.environ.pop("APH_AR_TWO_SHOT_BYTES")

    t = torch.arange(24.0).reshape(4, 6)
    assert is_weak_contiguous(t) and is_weak_contiguous(t.t()) and not is_weak_contiguous(t.t()[:2])
    import torch.distributed as dist
    created = not dist.is_initialized()
    if created:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29600 + os.getpid() % 200))
        dist.init_process_group("gloo", rank=0, world_size=1)
    try:
        ca = CustomAllreduce(dist.group.WORLD, "cpu")
        assert ca.disabled and ca.disabled_reason == "single GPU"
        assert ca.custom_all_reduce(torch.ones(16)) is None and not ca.should_custom_ar(torch.ones(16))
        with ca.capture():
            assert ca._IS_CAPTURING
        assert not ca._IS_CAPTURING
        ca.close()
    finally:
        if created:
            dist.destroy_process_group()
    # the eligibility rule of an enabled communicator (metadata only)
    ca = object.__new__(CustomAllreduce)
    ca.disabled, ca.world_size, ca.fully_connected, ca.max_size, ca._impl = False, 8, True, 8192 * 1024, None
    assert ca.should_custom_ar(torch.zeros(64, 8192, dtype=torch.bfloat16))
    assert not ca.should_custom_ar(torch.zeros(512, 8192, dtype=torch.bfloat16))   # == max_size: the reference's rule is "<"
    assert not ca.should_custom_ar(torch.zeros(3, 4, dtype=torch.bfloat16))        # 24 bytes
    assert not ca.should_custom_ar(torch.zeros(64, 8, dtype=torch.int32))          # not a floating type this path sums
    assert ca.should_custom_ar(torch.zeros(64, 128, dtype=torch.bfloat16).t())     # dense with permuted strides
    ca.disabled = True
    assert not ca.should_custom_ar(torch.zeros(64, 8192, dtype=torch.bfloat16))


def test_decode_layer_shards_partition_the_unsharded_layer():
    """DecodeLayer.shard_of (decode.py): the ranks' shards of an unsharded layer - q / k / v and gate / up split
    separately (linear.py:474-484), row-parallel K splits (:1279-1283), kv heads split or replicated
    (llama.py:118-127) - tile the full tensors exactly. Pure tensor slicing: runs on the CPU."""
    from aphrodite_fork_b200.decode import DecodeConfig, DecodeLayer
    cfg = DecodeConfig(hidden=1024, intermediate=2048, num_heads=16, num_kv_heads=4, head_size=128, num_layers=1)
    gen = torch.Generator().manual_seed(0)
    full = DecodeLayer(cfg, tp.TPGroup.single(), 6, "cpu", gen)
    d, H, KVH = cfg.head_size, cfg.num_heads, cfg.num_kv_heads
    for t in (2, 4, 8):
        shards = []
        for r in range(t):
            g = tp.TPGroup.single()
            g.world_size, g.rank = t, r
            shards.append(DecodeLayer.shard_of(full, g))
        hl, kvl = H // t, max(1, KVH // t)
        wq = full.qkv.w
        for r, sh in enumerate(shards):
            assert (sh.h, sh.kvh) == (hl, kvl)
            kv_first = (r * min(t, KVH) // t) * kvl  # first kv head held by rank r (replicated once t > KVH)
            cols = lambda c0, n: wq.scales[:, c0:c0 + n]
            want_s = torch.cat([cols(r * hl * d, hl * d), cols(H * d + kv_first * d, kvl * d), cols(H * d + KVH * d + kv_first * d, kvl * d)], 1)
            assert torch.equal(sh.qkv.w.scales, want_s)
            assert sh.qkv.w.q_weight.shape == (cfg.hidden // 16, 2 * (hl + 2 * kvl) * d)
            assert torch.equal(sh.key_cache, full.key_cache[:, kv_first:kv_first + kvl])
            assert torch.equal(sh.value_cache, full.value_cache[:, kv_first:kv_first + kvl])
            assert torch.equal(sh.o.w.q_weight, full.o.w.q_weight[r * hl * d // 16:(r + 1) * hl * d // 16])
            il = cfg.intermediate // t
            assert torch.equal(sh.gate_up.w.scales, torch.cat([full.gate_up.w.scales[:, r * il:(r + 1) * il],
                                                               full.gate_up.w.scales[:, cfg.intermediate + r * il:cfg.intermediate + (r + 1) * il]], 1))
            assert torch.equal(sh.down.w.q_weight, full.down.w.q_weight[r * il // 16:(r + 1) * il // 16])
        # the row-parallel shards together hold every k16-row of o and down exactly once
        assert torch.equal(torch.cat([sh.o.w.q_weight for sh in shards]), full.o.w.q_weight)
