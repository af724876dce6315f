"""Same-box comparison (test infrastructure, not product): gptq_marlin_gemm of this library against the UNMODIFIED
reference CUDA kernel (oracle/_ref/refcuda.so) on the C3 / TP8 shapes. CUDA-graph replay of 30 calls over weight
buffers rotated through > 300 MB. One JSON line per shape: us per call for ours / the reference at each batch size.
usage: python tests/tools/bench_gemm_vs_ref.py [1,16,64,256] [shape,shape]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import aphrodite_fork_b200 as A
from oracle import oracle

UINT4B8 = 1125899907892224
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096),
          "70b_tp8_qkv": (8192, 1280), "70b_tp8_o": (1024, 8192), "70b_tp8_gate_up": (8192, 7168), "70b_tp8_down": (3584, 8192)}


def bench(fn, K, N, M, g=128, iters=30):
    dev = "cuda"
    nbuf = max(2, int(300e6 // (K * N // 2)) + 1)
    gen = torch.Generator(device=dev).manual_seed(0)
    bs = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=dev, generator=gen) for _ in range(nbuf)]
    ss = [(torch.rand(K // g, N, device=dev, generator=gen) * 0.01 + 0.005).to(torch.bfloat16) for _ in range(nbuf)]
    a = torch.randn(M, K, device=dev, generator=gen).to(torch.bfloat16)
    e = torch.empty(0, dtype=torch.int32, device=dev)
    ws = torch.zeros(max(N // 64 * 16, 148 * 16), dtype=torch.int32, device=dev)
    f = lambda i: fn(a, bs[i % nbuf], ss[i % nbuf], e, e, e, ws, UINT4B8, M, N, K, True, False, False, True, False)
    for i in range(3):
        f(i)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
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
    return round(e0.elapsed_time(e1) / iters * 1e3, 2)


if __name__ == "__main__":
    A.load()
    ref = oracle.refcuda()
    assert ref is not None, "oracle/_ref/refcuda.so is not built"
    ms = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 16, 64, 256]
    only = sys.argv[2].split(",") if len(sys.argv) > 2 else list(SHAPES)
    for name in only:
        K, N = SHAPES[name]
        row = {"shape": name, "K": K, "N": N}
        for M in ms:
            row[f"m{M}"] = {"ours_us": bench(A.ops.gptq_marlin_gemm, K, N, M), "ref_us": bench(ref.gptq_marlin_gemm, K, N, M)}
        print(json.dumps(row), flush=True)
