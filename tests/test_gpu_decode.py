"""GPU parity of one decode step (GEMM -> cache write -> paged attention -> GEMM -> ... ) against the oracle, and
the same step over 2 GPUs with NCCL when the box has them. Tolerance: the step chains four bf16 GEMM
outputs and one attention per layer, so the bound is 3e-2 of the output's max magnitude (each stage is covered at its
own, tighter, tolerance in test_gpu_marlin.py / test_gpu_attention.py)."""
import os
import socket

import pytest
import torch

import aphrodite_fork_b200 as A
from aphrodite_fork_b200.decode import DecodeConfig, DecodeModel
from oracle import oracle
from oracle import layer_glue as lg
from oracle import marlin_format as mf

pytestmark = pytest.mark.gpu


def _dq(lin):
    w = lin.w
    return mf.dequantize(w.q_weight.cpu(), w.scales.cpu(), None, w.size_k, w.size_n, 4, w.group_size, False)


def _bf16(x64):
    return x64.to(torch.bfloat16)


def _reference_step(model: DecodeModel, caches_before):
    """The same step on the CPU: fp64 GEMMs on the dequantised weights rounded to bf16 where the kernels round, the
    glue from oracle/layer_glue.py."""
    cfg = model.cfg
    x, residual = model.x.cpu().clone(), None
    bt, sl, slots = model.block_tables.cpu(), model.seq_lens.cpu(), model.slot_mapping.cpu()
    pos, cache = model.positions.cpu(), model.cos_sin_cache.cpu()
    for layer, (kc, vc) in zip(model.layers, caches_before):
        d = cfg.head_size
        if residual is None:
            residual, h = x, lg.rms_norm(x, layer.input_layernorm.cpu(), cfg.rms_norm_eps)
        else:
            h, residual = lg.fused_add_rms_norm(x, residual, layer.input_layernorm.cpu(), cfg.rms_norm_eps)
        qkv = _bf16(oracle.gemm(h, _dq(layer.qkv)))
        q, k, v = qkv.split([layer.q_size, layer.kv_size, layer.kv_size], dim=-1)
        q, k = lg.rotary_embedding(pos, q.contiguous(), k.contiguous(), d, cache, True)
        k = k.reshape(-1, layer.kvh, d).contiguous()
        v = v.reshape(-1, layer.kvh, d).contiguous()
        oracle.reshape_and_cache(k, v, kc, vc, slots, cfg.kv_cache_dtype, 1.0, 1.0)
        _, a = oracle.paged_attention(q.reshape(-1, layer.h, d).contiguous(), kc, vc, layer.kvh, layer.scale, bt, sl,
                                      cfg.block_size, None, cfg.kv_cache_dtype, 1.0, 1.0)
        o = _bf16(oracle.gemm(a.reshape(-1, layer.q_size), _dq(layer.o)))
        h, residual = lg.fused_add_rms_norm(o, residual, layer.post_attention_layernorm.cpu(), cfg.rms_norm_eps)
        gu = _bf16(oracle.gemm(h, _dq(layer.gate_up)))
        x = _bf16(oracle.gemm(lg.silu_and_mul(gu), _dq(layer.down)))
    return lg.fused_add_rms_norm(x, residual, model.norm.cpu(), cfg.rms_norm_eps)[0]


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("batch", [3, 20])
def test_decode_step_matches_oracle(kv, batch):
    cfg = DecodeConfig(hidden=512, intermediate=1024, num_heads=8, num_kv_heads=2, head_size=64, num_layers=2,
                       group_size=128, kv_cache_dtype=kv)
    model = DecodeModel(cfg, batch=batch, ctx=77, seed=1)
    before = [(l.key_cache.cpu().clone(), l.value_cache.cpu().clone()) for l in model.layers]
    out = model.step().cpu().double()
    ref = _reference_step(model, [(k.clone(), v.clone()) for k, v in before]).double()
    assert torch.isfinite(out).all()
    err = (out - ref).abs().max() / ref.abs().max()
    assert float(err) < 3e-2, f"decode step differs from the oracle: {float(err):.3e}"
    # positions other than the new token's slot are untouched by the step
    kc_gpu, (kc_before, _) = model.layers[0].key_cache.cpu(), before[0]
    changed = (kc_gpu.view(torch.uint8) != kc_before.view(torch.uint8)).flatten(2).any(dim=2).any(dim=1)
    written_blocks = set((model.slot_mapping.cpu() // cfg.block_size).tolist())
    assert set(changed.nonzero().flatten().tolist()) <= written_blocks


def test_decode_step_under_cuda_graph_is_repeatable():
    cfg = DecodeConfig(hidden=1024, intermediate=2048, num_heads=16, num_kv_heads=4, head_size=128, num_layers=3)
    model = DecodeModel(cfg, batch=16, ctx=300, seed=2)
    eager = model.step().clone()
    model.capture()
    g1 = model.step().clone()
    g2 = model.step().clone()
    assert torch.equal(g1, g2)
    assert torch.allclose(g1.float(), eager.float(), rtol=2e-2, atol=1e-3)
    assert model.launches_per_step() == 34


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, q):
    import faulthandler
    import torch.distributed as dist
    faulthandler.dump_traceback_later(300, exit=True)  # a hang must end with a stack trace, not with the suite's timeout
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from aphrodite_fork_b200.tp import TPGroup
        cfg = DecodeConfig(hidden=1024, intermediate=2048, num_heads=16, num_kv_heads=4, head_size=128, num_layers=2)
        model = DecodeModel(cfg, batch=8, ctx=200, tp=TPGroup(), device=f"cuda:{rank}", seed=3)
        assert (model.layers[0].h, model.layers[0].kvh) == (16 // world, 4 // world)
        out = model.step().float()
        # every rank must hold the same hidden state after the all_reduces
        ref = out.clone()
        dist.broadcast(ref, src=0)
        same = bool(torch.equal(out, ref))
        model.capture()
        g = model.step().float()
        # against the UNSHARDED model: every rank builds the same full model (same seed), steps it alone, then steps its
        # shard of that model with the other ranks; a sharding mistake that is consistent across ranks shows up here
        full = DecodeModel(cfg, batch=8, ctx=200, tp=TPGroup.single(), device=f"cuda:{rank}", seed=5)
        sharded = DecodeModel.shard_of(full, TPGroup())
        want = full.step().float()
        got = sharded.step().float()
        rel = float((got - want).abs().max() / want.abs().max())
        assert sharded.layers[0].key_cache.shape[1] == sharded.layers[0].kvh == 4 // world
        kv_ok = all(torch.equal(ls.key_cache, lf.key_cache[:, rank * ls.kvh:(rank + 1) * ls.kvh])
                    and torch.equal(ls.value_cache, lf.value_cache[:, rank * ls.kvh:(rank + 1) * ls.kvh])
                    for ls, lf in zip(sharded.layers[:1], full.layers[:1]))  # layer 0 sees identical inputs: cache writes bit-equal
        res = (rank, same, bool(torch.isfinite(out).all()), bool(torch.allclose(g, out, rtol=2e-2, atol=1e-3)), rel, kv_ok)
        del full, sharded
        # the captured graph holds the communicator: drop it before the group is torn down
        model.graph = None
        del model
        torch.cuda.synchronize()
        q.put(res)
    except Exception:  # the parent must see the reason, not a queue timeout
        import traceback
        q.put((rank, "EXCEPTION " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tensor_parallel_decode_step_over_nccl():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=360) for _ in range(2))  # drain the queue before joining
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r, res in enumerate(got):
        assert res[:4] == (r, True, True, True), res
        # sharded vs unsharded step: same weights, same KV, only the all-reduce's summation order differs (bf16 hidden states)
        assert res[4] < 3e-2, f"rank {r}: tensor-parallel step differs from the unsharded model by {res[4]:.3e} of max|ref|"
        assert res[5], f"rank {r}: the shard's layer-0 cache write differs from the unsharded model's"


def _ar_worker(rank, world, port, q):
    import faulthandler
    import torch.distributed as dist
    faulthandler.dump_traceback_later(300, exit=True)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from aphrodite_fork_b200.tp import TPGroup
        tp = TPGroup()
        ok = tp.enable_custom_all_reduce(dev)
        if not ok:
            q.put((rank, "setup failed: " + str(tp.custom_error)))
            return
        bad = []
        gen = torch.Generator(device=dev).manual_seed(100 + rank)
        # every message once through the one-shot and once through the two-shot kernel (the threshold between them is a
        # host-side policy, APH_AR_TWO_SHOT_BYTES forces it), then with the default policy: the calls mix freely
        # ... and, where the box has multicast (NVLS) support, through the multimem kernel: the switch adds in its own
        # order, so that result is checked to one ulp of the rank-order sum and to be identical on every rank
        has_mc = bool(tp.custom.multicast_ptr)
        for mode in ("0", "16", "ll", None) + (("multimem",) if has_mc else ()):
          os.environ.pop("APH_AR_TWO_SHOT_BYTES", None)
          os.environ.pop("APH_AR_MODE", None)
          if mode in ("multimem", "ll"):
              os.environ["APH_AR_MODE"] = mode
          elif mode is not None:
              os.environ["APH_AR_TWO_SHOT_BYTES"] = mode
          exact = mode != "multimem"  # the default policy only picks the rank-order kernels
          for dtype in (torch.bfloat16, torch.float16, torch.float32):
            for shape in ((1, 4096), (64, 8192), (256, 8192), (3, 8), (64, 8200)):
                for rep in range(3):
                    x = torch.randn(shape, device=dev, generator=gen).to(dtype)
                    if not tp.custom.should_use(x):  # larger groups: only the small messages are one-shot
                        continue
                    parts = [torch.empty_like(x) for _ in range(world)]
                    dist.all_gather(parts, x)
                    want = parts[0].float()
                    for p in parts[1:]:
                        want = want + p.float()  # fp32, rank order, one rounding: the kernel's definition
                    want = want.to(dtype)
                    assert tp.custom.should_use(x)
                    got = tp.all_reduce(x.clone())
                    if exact:
                        if not torch.equal(got, want):
                            bad.append((mode, str(dtype), shape, rep, float((got.float() - want.float()).abs().max())))
                        continue
                    # one ulp of the result for the final rounding, plus the fp32 slop of adding in another order
                    mag = parts[0].float().abs()
                    for p in parts[1:]:
                        mag = mag + p.float().abs()
                    ulp = (0 if dtype == torch.float32 else torch.finfo(dtype).eps) * want.float().abs() + 4 * 1.2e-7 * mag
                    if not bool(((got.float() - want.float()).abs() <= ulp).all()):
                        bad.append((mode, "beyond one ulp", str(dtype), shape, rep, float((got.float() - want.float()).abs().max())))
                    gots = [torch.empty_like(got) for _ in range(world)]
                    dist.all_gather(gots, got)
                    if not all(torch.equal(g, gots[0]) for g in gots):
                        bad.append((mode, "ranks differ", str(dtype), shape, rep))
        os.environ.pop("APH_AR_MODE", None)
        # all-reduce fused with the residual add + RMSNorm that follows it: against the unfused pair of kernels
        from aphrodite_fork_b200 import _custom_ops as ops
        for mode in ("0", "16", "ll") + (("multimem",) if has_mc else ()):
          os.environ.pop("APH_AR_TWO_SHOT_BYTES", None)
          os.environ.pop("APH_AR_MODE", None)
          if mode in ("multimem", "ll"):
              os.environ["APH_AR_MODE"] = mode  # the unfused call below goes through the same kind of kernel
          else:
              os.environ["APH_AR_TWO_SHOT_BYTES"] = mode
          for dtype in (torch.bfloat16, torch.float16):
            for m, h in ((1, 4096), (64, 8192), (7, 1024), (130, 5120)):
                x = torch.randn(m, h, device=dev, generator=gen).to(dtype)
                res = torch.randn(m, h, device=dev, generator=gen).to(dtype)
                w = (1 + 0.1 * torch.randn(h, device=dev, generator=gen)).to(dtype)
                dist.broadcast(res, src=0)  # the residual stream and the weights are replicated
                dist.broadcast(w, src=0)
                if not tp.custom.should_fuse_norm(x, res, w):
                    continue
                x2, res2 = x.clone(), res.clone()
                tp.custom.all_reduce(x2)
                ops.fused_add_rms_norm(x2, res2, w, 1e-5)
                tp.all_reduce_add_rms_norm(x, res, w, 1e-5)
                # multimem: a vector's owner rank differs between the two kernels, and with it the order in which the
                # switch adds the copies - the sums may differ in the last bit
                dres = (res.view(torch.int16).int() - res2.view(torch.int16).int()).abs()
                if (mode != "multimem" and not torch.equal(res, res2)) or int(dres.max()) > 1:
                    bad.append(("fused residual", mode, str(dtype), m, h, int(dres.max())))
                d = (x.view(torch.int16).int() - x2.view(torch.int16).int()).abs()
                tiny = (x.float() - x2.float()).abs() <= 1e-5
                lim = 4 if mode == "multimem" else 2
                if not bool(((d <= lim) | tiny).all()) or float(((d > 0) & ~tiny).float().mean()) > (0.05 if mode == "multimem" else 0.01):
                    bad.append(("fused norm", mode, str(dtype), m, h, int(d.max())))
        os.environ.pop("APH_AR_TWO_SHOT_BYTES", None)
        os.environ.pop("APH_AR_MODE", None)
        # (the default policy only picks rank-order kernels: the exact checks below run through it - LL for these sizes)
        # too large / misaligned messages take NCCL
        big = torch.ones(5 * 1024 * 1024, device=dev, dtype=torch.bfloat16)
        assert not tp.custom.should_use(big)
        assert float(tp.all_reduce(big)[0]) == world
        # many back-to-back calls under a CUDA graph: the flags must survive replays
        x = torch.randn(8 if world > 2 else 64, 8192, device=dev, generator=gen).to(torch.bfloat16)
        y = torch.empty_like(x)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            tp.custom.all_reduce(x, y)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                tp.custom.all_reduce(x, y)
        first = None
        for _ in range(10):
            g.replay()
            torch.cuda.synchronize()
            first = y.clone() if first is None else first
            if not torch.equal(y, first):
                bad.append(("graph replay differs",))
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        want = parts[0].float()
        for p in parts[1:]:
            want = want + p.float()
        if not torch.equal(first, want.to(torch.bfloat16)):
            bad.append(("graph result wrong",))
        del g
        torch.cuda.synchronize()
        # the adapter with the reference's communicator interface (custom_all_reduce.py:45-304): eager, warm-up inside
        # capture(), and a captured graph; ineligible messages come back as None
        from aphrodite_fork_b200.custom_all_reduce import CustomAllreduce
        ca = CustomAllreduce(dist.group.WORLD, dev)
        if ca.disabled:
            bad.append(("CustomAllreduce disabled", ca.disabled_reason))
        else:
            x = torch.randn(16, 4096, device=dev, generator=gen).to(torch.bfloat16)
            parts = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            want = parts[0].float()
            for p in parts[1:]:
                want = want + p.float()
            want = want.to(torch.bfloat16)
            got = ca.custom_all_reduce(x)
            if got is None or got.data_ptr() == x.data_ptr() or not torch.equal(got, want):
                bad.append(("CustomAllreduce eager",))
            xt = x.t()  # dense, permuted strides: weakly contiguous
            got_t = ca.custom_all_reduce(xt)
            if got_t is None or got_t.shape != xt.shape or not torch.equal(got_t, want.t()):
                bad.append(("CustomAllreduce weakly contiguous",))
            if ca.custom_all_reduce(torch.ones(3, device=dev)) is not None or ca.custom_all_reduce(
                    torch.ones(8192 * 1024 // 2, dtype=torch.bfloat16, device=dev)) is not None:
                bad.append(("CustomAllreduce served an ineligible message",))
            with ca.capture():
                s2 = torch.cuda.Stream()
                s2.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s2):
                    warm = ca.custom_all_reduce(x)
                torch.cuda.current_stream().wait_stream(s2)
                torch.cuda.synchronize()
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2):
                    y2 = ca.custom_all_reduce(x)
            g2.replay()
            torch.cuda.synchronize()
            if not torch.equal(warm, want) or not torch.equal(y2, want):
                bad.append(("CustomAllreduce under capture",))
            del g2
            ca.close()
        torch.cuda.synchronize()
        q.put((rank, bad))
    except Exception:
        import traceback
        q.put((rank, "EXCEPTION " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_one_shot_all_reduce_over_peer_memory():
    """csrc/all_reduce.cu against an fp32 rank-order sum of all_gather'ed inputs: bit-exact on every rank, eager and
    under CUDA-graph replay (the reference's tests/distributed/test_custom_all_reduce.py runs the same two modes)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    world = 8 if torch.cuda.device_count() >= 8 else 4 if torch.cuda.device_count() >= 4 else 2
    procs = [ctx.Process(target=_ar_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=360) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, bad in res:
        assert bad == [], (rank, bad)
