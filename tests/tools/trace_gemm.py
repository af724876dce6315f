"""APH_GEMM_PATH=tc APH_GEMM_TRACE=gpurun_out/trace.txt python tests/tools/trace_gemm.py M K N  (one warm call, then the traced call)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import aphrodite_fork_b200 as A
A.load()
M, K, N = [int(x) for x in sys.argv[1:4]]
dev = "cuda"
b = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=dev)
s = (torch.rand(K // 128, N, device=dev) * 0.01 + 0.005).to(torch.bfloat16)
a = torch.randn(M, K, device=dev).to(torch.bfloat16)
e = torch.empty(0, dtype=torch.int32, device=dev)
ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device=dev)
for i in range(3):
    out = A.ops.gptq_marlin_gemm(a, b, s, e, e, e, ws, 1125899907892224, M, N, K, True)
torch.cuda.synchronize()
print(open(os.environ["APH_GEMM_TRACE"]).read())
