"""Same-box timing of reshape_and_cache_flash and copy_blocks: ours vs the reference CUDA kernels (oracle/_ref)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import bench_vs_ref as B

DEV = "cuda"
B.A.load()
for T in (64, 8192):
    for kv in ("auto", "fp8"):
        H, D, bs = 8, 128, 16
        nb = T // bs + 64
        cdt = torch.uint8 if kv != "auto" else torch.bfloat16
        qkv = torch.randn(T, 3, H, D, device=DEV).to(torch.bfloat16)
        slots = torch.randperm(nb * bs, device=DEV)[:T].to(torch.int64)
        kc = torch.zeros(nb, bs, H, D, dtype=cdt, device=DEV)
        vc = torch.zeros_like(kc)
        one = torch.tensor(1.0, device=DEV)
        byt = T * H * D * (4 + 2 * kc.element_size()) + 8 * T
        for name, o in B.impls():
            us = B.timed(lambda i, o=o: o.reshape_and_cache_flash(qkv[:, 1], qkv[:, 2], kc, vc, slots, kv, one, one), 50)
            print(json.dumps({"op": "reshape_and_cache_flash", "T": T, "kv": kv, "impl": name, "us": round(us, 2),
                              "GBps": round(byt / us / 1e3, 1)}), flush=True)
for layers, pairs in ((32, 8), (32, 256)):
    nb, H, D, bs = 1024, 8, 128, 16
    kcs = [torch.zeros(nb, H, D // 16, bs, 16, dtype=torch.uint8, device=DEV) for _ in range(layers)]
    vcs = [torch.zeros(nb, H, D, bs, dtype=torch.uint8, device=DEV) for _ in range(layers)]
    perm = torch.randperm(nb)
    mapping = torch.stack([perm[:pairs], perm[pairs:2 * pairs]], dim=1).to(torch.int64).to(DEV)
    byt = layers * pairs * 2 * 2 * H * D * bs
    for name, o in B.impls():
        # eager timing: the op uploads its pointer table synchronously (reference too), so no graph
        for _ in range(3):
            o.copy_blocks(kcs, vcs, mapping)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            o.copy_blocks(kcs, vcs, mapping)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        print(json.dumps({"op": "copy_blocks", "layers": layers, "pairs": pairs, "impl": name, "us_per_call_eager": round(us, 1),
                          "GBps": round(byt / us / 1e3, 1)}), flush=True)
