"""Development stress: repeat one W4A16 GEMM shape eagerly, check run-to-run determinism and catch sporadic faults."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import aphrodite_fork_b200 as A

A.load()
M, K, N = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 200
dev = "cuda"
gen = torch.Generator(device=dev).manual_seed(0)
nbuf = 3
bs = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=dev, generator=gen) for _ in range(nbuf)]
ss = [(torch.rand(K // 128, N, device=dev, generator=gen) * 0.01 + 0.005).to(torch.bfloat16) for _ in range(nbuf)]
a = torch.randn(M, K, device=dev, generator=gen).to(torch.bfloat16)
e = torch.empty(0, dtype=torch.int32, device=dev)
ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device=dev)
ref = [None] * nbuf
bad = 0
for r in range(reps):
    outs = []
    for i in range(5):
        outs.append((i % nbuf, A.ops.gptq_marlin_gemm(a, bs[i % nbuf], ss[i % nbuf], e, e, e, ws, 1125899907892224, M, N, K, True)))
    torch.cuda.synchronize()
    for j, o in outs:
        if ref[j] is None:
            ref[j] = o.clone()
        elif not torch.equal(ref[j], o):
            bad += 1
            if bad <= 6:
                d = (ref[j].float() - o.float())
                nz = d.nonzero()
                rows, cols = nz[:, 0], nz[:, 1]
                print(f"rep {r} buf {j}: {nz.shape[0]} elements differ, rows {int(rows.min())}..{int(rows.max())} "
                      f"cols {int(cols.min())}..{int(cols.max())}, max |d| {float(d.abs().max()):.4g}, "
                      f"ref max {float(ref[j].float().abs().max()):.4g}, distinct rows {rows.unique().numel()} "
                      f"distinct cols {cols.unique().numel()}, nan {int(torch.isnan(o.float()).sum())}", flush=True)
    assert int(ws.abs().sum()) == 0, "workspace not zero"
print(f"M={M} K={K} N={N}: {reps} x 5 calls, {bad} mismatching outputs")
