"""Development sweep: split-K partitioning of the paged-attention kernel (waves of CTAs aimed at, smallest
partition) on the C2 and C5 shapes."""
import os
import sys

os.environ["APH_DEV_ENV"] = "1"  # the library re-reads its development overrides on every call

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
os.environ["ONLY_OURS"] = "1"
import bench_vs_ref as B

B.A.load()
for waves in (1, 4):
    for min_part in (256, 1024, 2048, 4096):
        os.environ["APH_PA_WAVES"] = str(waves)
        os.environ["APH_PA_MIN_PART"] = str(min_part)
        B.attention(64, 4096, 32, 8, 128, "fp8", 2, f"C2 w{waves} m{min_part}")
        B.attention(8, 32768, 8, 1, 128, "fp8_e5m2", 2, f"C5 w{waves} m{min_part}")
        B.attention(32, 32768, 8, 1, 128, "fp8_e5m2", 2, f"C5x32 w{waves} m{min_part}")
        B.attention(1, 2048, 32, 8, 128, "fp8", 2, f"1x2048 w{waves} m{min_part}", iters=50)
        B.attention(4, 8192, 32, 8, 128, "auto", 2, f"4x8192bf16 w{waves} m{min_part}", iters=50)
        B.attention(256, 1024, 32, 8, 128, "fp8", 2, f"256x1024 w{waves} m{min_part}")
