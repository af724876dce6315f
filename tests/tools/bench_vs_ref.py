"""Same-box comparison: our kernels vs the UNMODIFIED reference CUDA kernels (oracle/_ref/refcuda.so,
`make -C oracle refcuda`) on the BASELINE shapes. Checker/measurement tool only (tests/tools/ is checker tooling, not the
product path). Timing: CUDA graph of `iters` back-to-back calls over rotating buffers larger than
L2, CUDA events on the capturing stream, one warm replay before the timed one.

    python tests/tools/bench_vs_ref.py [attention] [gemm] [cache]  > profiles/rNN_vs_reference_cuda.jsonl
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import aphrodite_fork_b200 as A
from oracle import oracle

UINT4B8 = 1125899907892224
DEV = "cuda"


def timed(fn, iters):
    """fn(i) launches call i; returns us per call under one CUDA graph."""
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for i in range(iters):
                fn(i)
        g.replay()
        st.synchronize()
        e0.record(st)
        g.replay()
        e1.record(st)
        st.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    del g
    return us


def impls():
    ref = oracle.refcuda()
    out = [("ours", A.ops)]
    if ref is not None and not os.environ.get("ONLY_OURS"):
        out.append(("reference_cuda", ref))
    return out


def attention(S, ctx, H, KVH, D, kv, version, label, iters=20):
    bs = 16
    esz = 1 if kv != "auto" else 2
    x = 16 // esz
    nblk = S * ctx // bs
    npool = max(2, int(300e6 // (2 * nblk * KVH * D * bs * esz)) + 1)
    gen = torch.Generator(device=DEV).manual_seed(0)
    sc = D ** -0.5
    pools = []
    for _ in range(npool):
        k = ((torch.rand(nblk, KVH, D // x, bs, x, device=DEV, generator=gen) * 2 - 1) * sc).to(torch.bfloat16)
        v = ((torch.rand(nblk, KVH, D, bs, device=DEV, generator=gen) * 2 - 1) * sc).to(torch.bfloat16)
        if kv != "auto":
            k8, v8 = torch.empty(k.shape, dtype=torch.uint8, device=DEV), torch.empty(v.shape, dtype=torch.uint8, device=DEV)
            A.ops.convert_fp8(k8, k, 1.0, "fp8")
            A.ops.convert_fp8(v8, v, 1.0, "fp8")
            k, v = k8, v8
        pools.append((k, v))
    q = ((torch.rand(S, H, D, device=DEV, generator=gen) * 2 - 1) * sc).to(torch.bfloat16)
    bt = torch.randperm(nblk, device=DEV, generator=gen).to(torch.int32).reshape(S, ctx // bs)
    sl = torch.full((S,), ctx, dtype=torch.int32, device=DEV)
    one = torch.tensor(1.0, device=DEV)
    P = (ctx + 511) // 512
    es = torch.empty(S, H, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    tmp = torch.empty(S, H, P, D, dtype=torch.bfloat16, device=DEV)
    byt = 2 * S * ctx * KVH * D * esz
    outs = {}
    for name, o in impls():
        out = torch.empty_like(q)

        def f(i, o=o, out=out):
            k, v = pools[i % npool]
            if version == 1:
                o.paged_attention_v1(out, q, k, v, KVH, sc, bt, sl, bs, ctx, None, kv, one, one, 0, 0, 0, 64, 0)
            else:
                o.paged_attention_v2(out, es, ml, tmp, q, k, v, KVH, sc, bt, sl, bs, ctx, None, kv, one, one, 0, 0, 0, 64, 0)
        us = timed(f, iters)
        f(0)
        torch.cuda.synchronize()
        outs[name] = out.float().clone()
        print(json.dumps({"op": f"paged_attention_v{version}", "workload": label, "impl": name, "us": round(us, 1),
                          "KV_GBps": round(byt / us / 1e3, 1)}), flush=True)
    if len(outs) == 2:
        d = float((outs["ours"] - outs["reference_cuda"]).abs().max())
        print(json.dumps({"op": f"paged_attention_v{version}", "workload": label, "max_abs_diff_ours_vs_reference": d}), flush=True)


def gemm(name, K, N, M, g=128, iters=40):
    nbuf = max(2, int(300e6 // (K * N // 2)) + 1)
    gen = torch.Generator(device=DEV).manual_seed(0)
    bs = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=DEV, generator=gen) for _ in range(nbuf)]
    ss = [(torch.rand(K // g, N, device=DEV, generator=gen) * 0.01 + 0.005).to(torch.bfloat16) for _ in range(nbuf)]
    a = torch.randn(M, K, device=DEV, generator=gen).to(torch.bfloat16)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    ws = torch.zeros(max(N // 64 * 16, 148 * 16 * 4), dtype=torch.int32, device=DEV)
    byt = K * N // 2 + (K // g) * N * 2 + M * K * 2 + M * N * 2
    res = {}
    for iname, o in impls():
        f = lambda i, o=o: o.gptq_marlin_gemm(a, bs[i % nbuf], ss[i % nbuf], e, e, e, ws, UINT4B8, M, N, K, True, False,
                                              False, True, False)
        try:
            f(0)
        except NotImplementedError:
            continue
        us = timed(f, iters)
        res[iname] = f(0).float()
        print(json.dumps({"op": "gptq_marlin_gemm", "shape": name, "K": K, "N": N, "M": M, "impl": iname, "us": round(us, 2),
                          "GBps": round(byt / us / 1e3, 1), "TFLOPs": round(2 * M * N * K / us / 1e6, 1)}), flush=True)
    if len(res) == 2:
        d = float((res["ours"] - res["reference_cuda"]).abs().max() / res["reference_cuda"].abs().max())
        print(json.dumps({"op": "gptq_marlin_gemm", "shape": name, "M": M, "max_diff_over_max_ref": d}), flush=True)


def cache(T, KVH, D, kv, iters=50, consecutive=False):
    bs = 16
    nb = T // bs + 64
    esz = 1 if kv != "auto" else 2
    cdt = torch.uint8 if kv != "auto" else torch.bfloat16
    x = 16 // esz
    gen = torch.Generator(device=DEV).manual_seed(0)
    qkv = torch.randn(T, 3, KVH, D, device=DEV, generator=gen).to(torch.bfloat16)
    slots = torch.randperm(nb * bs, device=DEV, generator=gen)[:T].to(torch.int64)
    if consecutive:  # prefill: every sequence fills whole blocks in order
        slots = torch.arange(T, device=DEV, dtype=torch.int64)
    kc = torch.zeros(nb, KVH, D // x, bs, x, dtype=cdt, device=DEV)
    vc = torch.zeros(nb, KVH, D, bs, dtype=cdt, device=DEV)
    one = torch.tensor(1.0, device=DEV)
    byt = T * KVH * D * (2 * 2 + 2 * esz) + 8 * T
    for name, o in impls():
        us = timed(lambda i, o=o: o.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots, kv, one, one), iters)
        print(json.dumps({"op": "reshape_and_cache", "T": T, "kv": kv, "slots": "consecutive" if consecutive else "random", "impl": name, "us": round(us, 2),
                          "GBps": round(byt / us / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    A.load()
    what = sys.argv[1:] or ["attention", "gemm", "cache"]
    if oracle.refcuda() is None:
        print(json.dumps({"note": "oracle/_ref/refcuda.so not built: only our kernels are timed"}), flush=True)
    if "attention" in what:
        attention(64, 4096, 32, 8, 128, "fp8", 2, "C2: 32q/8kv d128 fp8_e4m3 block16 batch64 x ctx4096")
        attention(64, 4096, 32, 8, 128, "fp8", 1, "C2 through v1")
        attention(64, 4096, 32, 8, 128, "auto", 2, "C2 shape with bf16 KV")
        attention(8, 32768, 8, 1, 128, "fp8_e5m2", 2, "C5 per GPU: 8q/1kv d128 fp8_e5m2 batch8 x ctx32768")
        attention(8, 512, 12, 12, 64, "auto", 1, "C1 on GPU: 12 heads d64 bf16 8 x 512")
    if "gemm" in what:
        shapes = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096),
                  "70b_gate_up": (8192, 57344), "70b_down": (28672, 8192)}
        for name, (K, N) in shapes.items():
            for M in ([1, 16, 64, 256] if not name.startswith("70b") else [1, 16, 256]):
                gemm(name, K, N, M)
    if "cache" in what:
        for T in (64, 8192):
            for kv in ("auto", "fp8"):
                cache(T, 8, 128, kv)
        for kv in ("auto", "fp8"):
            cache(8192, 8, 128, kv, consecutive=True)
