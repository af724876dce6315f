"""Per-kernel time of ONE decoder layer of the decode step on ONE GPU, at the shapes a rank sees at a given tensor-parallel
degree (no collective here: the all-reduce is timed by tools/bench_all_reduce.py on the real ranks). Each kernel alone
under CUDA-graph replay (20 launches, weights / KV rotated through > 256 MB), so the numbers are per-launch floors
including launch latency - what the layer costs when nothing overlaps.
    python tools/layer_breakdown.py 8b:1 70b:2 70b:4 70b:8   ->  one JSON line per configuration"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import aphrodite_fork_b200 as A
from aphrodite_fork_b200 import _custom_ops as ops

UINT4B8 = 1125899907892224
MODELS = {"8b": (4096, 14336, 32, 8, 128), "70b": (8192, 28672, 64, 8, 128)}
BATCH, CTX, BS = 64, 4096, 16
dev = "cuda"


def timed(make_call, n_variants, iters=20):
    calls = [make_call(i) for i in range(n_variants)]
    for c in calls[:2]:
        c()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for i in range(iters):
                calls[i % n_variants]()
        g.replay()
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(2_000_000)  # the host enqueues the launches while the device is busy: the events see kernels only
        e0.record(st)
        g.replay()
        g.replay()
        e1.record(st)
        st.synchronize()
    return round(e0.elapsed_time(e1) / (2 * iters) * 1e3, 2)


def gemm(K, N, M):
    nbuf = max(2, int(300e6 // (K * N // 2)) + 1)
    gen = torch.Generator(device=dev).manual_seed(0)
    bs = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=dev, generator=gen) for _ in range(nbuf)]
    ss = [(torch.rand(K // 128, N, device=dev, generator=gen) * 0.01 + 0.005).to(torch.bfloat16) for _ in range(nbuf)]
    a = torch.randn(M, K, device=dev, generator=gen).to(torch.bfloat16)
    e = torch.empty(0, dtype=torch.int32, device=dev)
    ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device=dev)
    return timed(lambda i: (lambda: ops.gptq_marlin_gemm(a, bs[i], ss[i], e, e, e, ws, UINT4B8, M, N, K, True)), nbuf)


def layer(model, tp):
    hidden, inter, H, KVH, d = MODELS[model]
    h, kvh = H // tp, max(1, KVH // tp)
    q_size, kv_size, il = h * d, kvh * d, inter // tp
    r = {"model": model, "tp": tp, "batch": BATCH, "ctx": CTX}
    r["qkv_gemm"] = gemm(hidden, q_size + 2 * kv_size, BATCH)
    r["o_gemm"] = gemm(q_size, hidden, BATCH)
    r["gate_up_gemm"] = gemm(hidden, 2 * il, BATCH)
    r["down_gemm"] = gemm(il, hidden, BATCH)
    gen = torch.Generator(device=dev).manual_seed(1)
    x = torch.randn(BATCH, hidden, device=dev, generator=gen).to(torch.bfloat16)
    res = torch.randn(BATCH, hidden, device=dev, generator=gen).to(torch.bfloat16)
    w = torch.ones(hidden, device=dev, dtype=torch.bfloat16)
    r["fused_add_rms_norm"] = timed(lambda i: (lambda: ops.fused_add_rms_norm(x, res, w, 1e-5)), 1)
    qkv = torch.randn(BATCH, q_size + 2 * kv_size, device=dev, generator=gen).to(torch.bfloat16)
    q, k, v = qkv.split([q_size, kv_size, kv_size], dim=-1)
    pos = torch.full((BATCH,), CTX - 1, dtype=torch.int64, device=dev)
    cs = torch.randn(CTX, d, device=dev, generator=gen).to(torch.bfloat16)
    r["rotary_embedding"] = timed(lambda i: (lambda: ops.rotary_embedding(pos, q, k, d, cs, True)), 1)
    nblk = BATCH * CTX // BS
    byt = 2 * BATCH * CTX * kvh * d
    npools = max(2, min(16, -(-(300 << 20) // byt)))
    pools = [(torch.randint(0, 120, (nblk, kvh, d // 16, BS, 16), dtype=torch.uint8, device=dev, generator=gen),
              torch.randint(0, 120, (nblk, kvh, d, BS), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(npools)]
    bt = torch.randperm(nblk, generator=gen, device=dev).to(torch.int32).reshape(BATCH, CTX // BS)
    sl = torch.full((BATCH,), CTX, dtype=torch.int32, device=dev)
    slots = bt[:, -1].to(torch.int64) * BS + BS - 1
    one = torch.tensor(1.0, device=dev)
    r["reshape_and_cache"] = timed(lambda i: (lambda: A.PagedAttention.write_to_paged_cache(
        k.reshape(-1, kvh, d), v.reshape(-1, kvh, d), pools[i][0], pools[i][1], slots, "fp8", one, one)), npools)
    qq = q.reshape(-1, h, d)
    r["paged_attention"] = timed(lambda i: (lambda: A.PagedAttention.forward_decode(qq, pools[i][0], pools[i][1], bt, sl, CTX, "fp8", kvh,
                                                                                     d ** -0.5, None, one, one)), npools)
    gu = torch.randn(BATCH, 2 * il, device=dev, generator=gen).to(torch.bfloat16)
    act = torch.empty(BATCH, il, device=dev, dtype=torch.bfloat16)
    r["silu_and_mul"] = timed(lambda i: (lambda: ops.silu_and_mul(act, gu)), 1)
    r["sum_us"] = round(sum(v for kk, v in r.items() if isinstance(v, float)) + r["fused_add_rms_norm"], 2)  # two norms per layer
    wb = (hidden * (q_size + 2 * kv_size) + q_size * hidden + hidden * 2 * il + il * hidden) // 2 * (1 + 4 / 128)
    r["hbm_MB_per_layer"] = round((wb + byt) / 1e6, 1)
    r["hbm_floor_us"] = round((wb + byt) / 6552.3e3, 2)
    del pools
    torch.cuda.empty_cache()
    return r


if __name__ == "__main__":
    A.load()
    for spec in sys.argv[1:] or ["8b:1", "70b:2", "70b:4", "70b:8"]:
        m, tp = spec.split(":")
        print(json.dumps(layer(m, int(tp))), flush=True)
