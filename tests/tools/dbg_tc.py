"""Debug driver for the tcgen05 GEMM path: APH_GEMM_PATH=tc python tools/dbg_tc.py [M K N gs zp dtype]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import test_gpu_marlin as T
from oracle import oracle

def one(M, K, N, gs, zp, dtype, seed=1):
    w_ref, b, s, bz = T.quantized(K, N, 4, gs, zp, dtype, seed)
    a = T.synth.tensor((M, K), seed + 100, dtype, 1.0)
    ref = oracle.gemm(a.contiguous(), w_ref)
    out = T.run_gemm(a.cuda(), b, s, bz, M, N, K, 4, zp)
    torch.cuda.synchronize()
    o = out.cpu().double()
    err = (o - ref).abs()
    print(f"M{M} K{K} N{N} g{gs} zp{zp} {dtype}: finite={bool(torch.isfinite(o).all())} mean_rel={float(err.mean()/ref.abs().mean()):.3e} "
          f"max_rel={float(err.max()/ref.abs().max()):.3e}", flush=True)
    if float(err.max() / ref.abs().max()) > 2e-2:
        bad = (err > 2e-2 * ref.abs().max())
        print("  bad rows:", sorted(set(bad.nonzero()[:, 0].tolist()))[:20], " bad cols(sample):", sorted(set(bad.nonzero()[:, 1].tolist()))[:40])
        print("  out[0,:8]", o[0, :8].tolist()); print("  ref[0,:8]", ref[0, :8].tolist())

if __name__ == "__main__":
    if len(sys.argv) > 1:
        M, K, N, gs, zp = [int(x) for x in sys.argv[1:6]]
        one(M, K, N, gs, bool(zp), torch.bfloat16 if len(sys.argv) < 7 or sys.argv[6] == "bf16" else torch.float16)
    else:
        for args in [(16, 128, 128, 128, False), (16, 256, 128, 128, False), (16, 512, 256, 128, False), (16, 1024, 512, -1, False),
                     (40, 512, 256, 32, True), (100, 1024, 512, 64, False), (256, 4096, 6144, 128, False), (300, 1024, 512, 128, False)]:
            one(*args, torch.bfloat16)
        one(16, 512, 256, 128, False, torch.float16)
        one(33, 512, 256, -1, True, torch.float16)
