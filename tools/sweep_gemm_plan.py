"""Development sweep: register-path W4A16 GEMM time as a function of the slice width NT and the grid
(stream-K ranges; grid = n_slices * split with split | k_stages reproduces plain split-K)."""
import json
import os
import sys

os.environ["APH_DEV_ENV"] = "1"  # the library re-reads its development overrides on every call

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_gemm
import aphrodite_fork_b200 as A

A.load()
shapes = dict(bench_gemm.SHAPES)
shapes.update({"70b_qkv": (8192, 10240), "70b_gate_up": (8192, 57344), "70b_down": (28672, 8192)})
only = sys.argv[2].split(",") if len(sys.argv) > 2 else list(shapes)
for M in [int(x) for x in sys.argv[1].split(",")]:
    for name in only:
        K, N = shapes[name]
        for nt in (64, 128, 256):
            if N % nt:
                continue
            ks = K // (128 * (4 // (nt // 64)))
            n_slices = N // nt
            grids = {148, 222, 296}
            for split in range(1, 9):
                if ks % split == 0 and n_slices * split <= 1200:
                    grids.add(n_slices * split)
            for g in sorted(grids):
                os.environ["APH_GEMM_NT"] = str(nt)
                os.environ["APH_GEMM_GRID"] = str(g)
                r = bench_gemm.bench(name, K, N, M, iters=30)
                print(json.dumps({"shape": name, "M": M, "nt": nt, "grid": g, "aligned": g % n_slices == 0 and ks % max(g // n_slices, 1) == 0,
                                  "us": r["us"], "GBps": r["GBps"]}), flush=True)
