import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_gemm import bench, SHAPES
import aphrodite_fork_b200 as A
A.load()
M = int(sys.argv[1]) if len(sys.argv) > 1 else 16
name = sys.argv[2] if len(sys.argv) > 2 else "gate_up"
K, N = SHAPES[name]
print(bench(name, K, N, M, iters=5))
