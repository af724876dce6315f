"""W4A16 GEMM microbench on the C3 shapes (Llama-3.1-8B, g128): us per call, HBM GB/s on the
algorithmic bytes (K*N/2 + K/g*N*2 + M*K*2 + M*N*2) and TFLOP/s, with rotating weight buffers > L2."""
import json
import os
import sys

os.environ.setdefault("APH_DEV_ENV", "1")  # development overrides are re-read on every call

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import aphrodite_fork_b200 as A

UINT4B8 = 1125899907892224
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}


def bench(name, K, N, M, g=128, iters=50):
    dev = "cuda"
    nbuf = max(2, int(300e6 // (K * N // 2)) + 1)
    gen = torch.Generator(device=dev).manual_seed(0)
    bs = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=dev, generator=gen) for _ in range(nbuf)]
    ss = [(torch.rand(K // g, N, device=dev, generator=gen) * 0.01 + 0.005).to(torch.bfloat16) for _ in range(nbuf)]
    a = torch.randn(M, K, device=dev, generator=gen).to(torch.bfloat16)
    e = torch.empty(0, dtype=torch.int32, device=dev)
    ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device=dev)
    f = lambda i: A.ops.gptq_marlin_gemm(a, bs[i % nbuf], ss[i % nbuf], e, e, e, ws, UINT4B8, M, N, K, True)
    for i in range(5):
        f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if os.environ.get("NOGRAPH"):
        e0.record()
        for i in range(iters):
            f(i)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / iters * 1e3
    else:
        # CUDA graph of `iters` calls over rotating weight buffers: device time without host launch gaps
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph, stream=st):
                for i in range(iters):
                    f(i)
            gph.replay()
            st.synchronize()
            torch.cuda._sleep(2_000_000)  # the host enqueues the launch while the device is busy: the events see kernels only
            e0.record(st)
            gph.replay()
            e1.record(st)
            st.synchronize()
        us = e0.elapsed_time(e1) / iters * 1e3
    byt = K * N // 2 + (K // g) * N * 2 + M * K * 2 + M * N * 2
    return {"shape": name, "K": K, "N": N, "M": M, "us": round(us, 2), "GBps": round(byt / us / 1e3, 1),
            "TFLOPs": round(2 * M * N * K / us / 1e6, 1)}


if __name__ == "__main__":
    A.load()
    shapes = dict(SHAPES)
    shapes.update({"70b_tp8_qkv": (8192, 1280), "70b_tp8_o": (1024, 8192), "70b_tp8_gate_up": (8192, 7168), "70b_tp8_down": (3584, 8192)})
    ms = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 16, 64, 256]
    only = sys.argv[2].split(",") if len(sys.argv) > 2 else list(shapes)
    for name in only:
        K, N = shapes[name]
        print(json.dumps({"shape": name, **{f"m{M}": bench(name, K, N, M)["us"] for M in ms}}), flush=True)
