"""Proof for the two-group ring fix (ptx.cuh RingCursor): W4A16 GEMMs through every kernel instantiation whose ring
depth is ODD, checked against the oracle's fp64 product, with one of three builds of the library:

  product            lib/                     per-group sub-rings
  fault              lib_ringfault/           same + -DAPH_RING_FAULT: producers land the odd items late
  fault + legacy     lib_ringfault_legacy/    -DAPH_RING_FAULT -DAPH_RING_LEGACY: round 1's slot = it % depth mapping

Expected: the first two are clean, the third returns wrong results on every odd-depth instantiation (the group that
runs ahead passes a full-barrier wait on parity alone and consumes a slot that still holds an older stage).
Oracle use: test infrastructure only (this script lives under tests/tools/).

  python tests/tools/ring_fault_proof.py build        # on the CPU box: the two variant libraries
  python tests/tools/ring_fault_proof.py run <out>    # on the GPU box: runs the three builds, writes JSON lines
  python tests/tools/ring_fault_proof.py child        # one build (selected by APH_B200_LIB_DIR), prints JSON lines
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
PKG = os.path.join(ROOT, "aphrodite-fork_b200")
VARIANTS = {"product": None, "fault": ["-DAPH_RING_FAULT"], "fault_legacy": ["-DAPH_RING_FAULT", "-DAPH_RING_LEGACY"]}
# (path, M, K, N, forced slice width or 0, forced grid or 0). Every CTA must run several times round its ring, so the
# register kernel gets a small forced grid (16 items per CTA): NB=1/NT=64 (depth 3), NB=2/NT=256 (3), NB=4/NT=64 (3),
# NB=4/NT=256 (5) through APH_GEMM_PATH=mma; tensor-memory kernel with >= 10 stages per CTA at 256 batch rows (weight
# ring 3 + activation ring 3, one and three batch tiles) and at 16 rows (activation ring 7).
CASES = [("mma", 4, 4096, 1024, 64, 8), ("mma", 16, 4096, 1024, 256, 8), ("mma", 32, 4096, 1024, 64, 8),
         ("mma", 32, 4096, 1024, 256, 8), ("tc", 256, 8192, 4096, 0, 0), ("tc", 700, 4096, 2048, 0, 0), ("tc", 16, 8192, 4096, 0, 0)]


def lib_dir(name):
    return os.path.join(PKG, "lib" if name == "product" else "lib_ring" + name)


def build():
    import importlib.util
    spec = importlib.util.spec_from_file_location("aph_b200_build", os.path.join(PKG, "build.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    for name, flags in VARIANTS.items():
        if flags:
            print(m.build_variant("ring" + name, flags))


def child():
    import torch
    import aphrodite_fork_b200 as A
    import synth
    from oracle import marlin_format as mf
    from oracle import oracle
    A.load()
    path, M, K, N = os.environ["APH_GEMM_PATH"], int(os.environ["CASE_M"]), int(os.environ["CASE_K"]), int(os.environ["CASE_N"])
    reps = int(os.environ.get("CASE_REPS", "100"))
    w = synth.tensor((K, N), 3, torch.bfloat16, 1.0)
    w_ref, b, s, _ = mf.marlin_quantize(w, 4, 128, False)
    a = synth.tensor((M, K), 103, torch.bfloat16, 1.0)
    ref = oracle.gemm(a, w_ref).cuda()
    a, b, s = a.cuda(), b.cuda(), s.cuda()
    e = torch.empty(0, dtype=torch.int32, device="cuda")
    ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device="cuda")
    bad, worst = 0, 0.0
    scale = float(ref.abs().max())
    for r in range(reps):
        out = A.ops.gptq_marlin_gemm(a, b, s, e, e, e, ws, mf.UINT4B8, M, N, K, True)
        err = float((out.double() - ref).abs().max()) / scale
        worst = max(worst, err)
        bad += err > 1e-2  # the parity tests' bound for bf16: max |err| <= 1e-2 max |ref|
    print(json.dumps({"lib": os.path.basename(os.environ.get("APH_B200_LIB_DIR") or "lib"), "path": path, "M": M, "K": K, "N": N,
                      "forced_nt": int(os.environ.get("APH_GEMM_NT", "0")), "forced_grid": int(os.environ.get("APH_GEMM_GRID", "0")),
                      "calls": reps, "wrong_calls": int(bad),
                      "worst_max_err_over_max_ref": round(worst, 5), "lock_buffer_zero": int(ws.abs().sum()) == 0}), flush=True)


def run(out_path):
    with open(out_path, "w") as f:
        for name in VARIANTS:
            for path, M, K, N, nt, grid in CASES:
                env = dict(os.environ, APH_B200_LIB_DIR=lib_dir(name), APH_GEMM_PATH=path, CASE_M=str(M), CASE_K=str(K), CASE_N=str(N))
                if nt:
                    env["APH_GEMM_NT"] = str(nt)
                if grid:
                    env["APH_GEMM_GRID"] = str(grid)
                r = subprocess.run([sys.executable, os.path.abspath(__file__), "child"], env=env, capture_output=True, text=True,
                                   timeout=600)
                line = r.stdout.strip().splitlines()[-1] if r.returncode == 0 and r.stdout.strip() else json.dumps(
                    {"lib": name, "path": path, "M": M, "K": K, "N": N, "forced_nt": nt, "failed": (r.stderr or "")[-300:]})
                f.write(line + "\n")
                f.flush()
                print(line, flush=True)


if __name__ == "__main__":
    {"build": build, "child": child}.get(sys.argv[1], lambda: run(sys.argv[2]))()
