import sys, os, faulthandler
faulthandler.enable()
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests", "golden"))
import torch
import aphrodite_fork_b200 as A
from oracle import marlin_format as mf
ops = A.ops
dev = "cuda"
K, N, M = 256, 128, 4
b = torch.zeros(K // 16, N * 2, dtype=torch.int32, device=dev)
s = torch.ones(K // 128, N, dtype=torch.bfloat16, device=dev)
a = torch.zeros(M, K, dtype=torch.bfloat16, device=dev)
e = torch.empty(0, dtype=torch.int32, device=dev)
ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device=dev)
order = sys.argv[1] if len(sys.argv) > 1 else "123"
for ch in order:
    print("case", ch, flush=True)
    try:
        if ch == "1":
            ops.gptq_marlin_gemm(a, b, s, e, e, e, ws, mf.UINT4, M, N, K, True, False, False, True, False)
        elif ch == "2":
            ops.gptq_marlin_gemm(a, b, s, e, e, e, ws, mf.UINT4B8, M + 1, N, K, True, False, False, True, False)
        elif ch == "3":
            ops.gptq_marlin_gemm(a, b, s, e, e, e, ws[:4], mf.UINT4B8, M, N, K, True, False, False, True, False)
        elif ch == "4":
            o = ops.gptq_marlin_gemm(a[:0], b, s, e, e, e, ws, mf.UINT4B8, 0, N, K, True, False, False, True, False)
            print(o.shape)
        elif ch == "5":
            o = ops.gptq_marlin_gemm(a, b, s, e, e, e, ws, mf.UINT4B8, M, N, K, True, False, False, True, False)
            print(o.shape, float(o.float().abs().sum()))
    except RuntimeError as ex:
        print("  RuntimeError:", str(ex)[:120], flush=True)
print("done")
