"""Development microbench: paged attention (forward_decode -> our kernels) on the BASELINE shapes under the merge
policies of paged_attention.cu (APH_PA_CLUSTER = 0 never / 1 always / 2 one-wave grids only). CUDA-graph replay over KV
pools rotated through > 256 MB. Prints one JSON line per (case, policy)."""
import json
import os
import sys

os.environ["APH_DEV_ENV"] = "1"  # the library re-reads its development overrides on every call
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import aphrodite_fork_b200 as A

BS = 16
CASES = [("C2", 32, 8, 128, 64, 4096, "fp8"), ("C2_bf16", 32, 8, 128, 64, 4096, "auto"), ("C2_tp2", 16, 4, 128, 64, 4096, "fp8"),
         ("C2_tp4", 8, 2, 128, 64, 4096, "fp8"), ("C2_tp8", 4, 1, 128, 64, 4096, "fp8"), ("C4_tp8", 8, 1, 128, 64, 4096, "fp8"),
         ("C4_tp2", 32, 4, 128, 64, 4096, "fp8"), ("C5_b1", 8, 1, 128, 1, 32768, "fp8_e5m2"), ("C5_b8", 8, 1, 128, 8, 32768, "fp8_e5m2"),
         ("C5_b32", 8, 1, 128, 32, 32768, "fp8_e5m2"), ("b1_2k", 32, 8, 128, 1, 2048, "fp8"), ("b16_1k", 32, 8, 128, 16, 1024, "fp8"),
         ("b256_1k", 32, 8, 128, 256, 1024, "fp8"), ("C1_gpu", 12, 12, 64, 8, 512, "auto")]


def bench(name, h, kvh, d, batch, ctx, kv, iters=20):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    one = torch.tensor(1.0, device=dev)
    esz = 2 if kv == "auto" else 1
    x = 16 // esz
    nblk = batch * ctx // BS
    byt = 2 * batch * ctx * kvh * d * esz
    npools = max(2, min(64, -(-(300 << 20) // byt)))
    pools = []
    for _ in range(npools):
        if kv == "auto":
            pools.append(((torch.rand(nblk, kvh, d // x, BS, x, generator=g, device=dev) - 0.5).to(torch.bfloat16),
                          (torch.rand(nblk, kvh, d, BS, generator=g, device=dev) - 0.5).to(torch.bfloat16)))
        else:
            pools.append((torch.randint(0, 120, (nblk, kvh, d // x, BS, x), dtype=torch.uint8, device=dev, generator=g),
                          torch.randint(0, 120, (nblk, kvh, d, BS), dtype=torch.uint8, device=dev, generator=g)))
    q = ((torch.rand(batch, h, d, generator=g, device=dev) * 2 - 1) * d ** -0.5).to(torch.bfloat16)
    bt = torch.randperm(nblk, generator=g, device=dev).to(torch.int32).reshape(batch, ctx // BS)
    sl = torch.full((batch,), ctx, dtype=torch.int32, device=dev)
    f = lambda i: A.PagedAttention.forward_decode(q, pools[i % npools][0], pools[i % npools][1], bt, sl, ctx, kv, kvh, d ** -0.5, None, one, one)
    out = {}
    for policy in (os.environ.get("POLICIES", "0,1,2").split(",")):
        os.environ["APH_PA_CLUSTER"] = policy
        for i in range(3):
            f(i)
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph, stream=st):
                for i in range(iters):
                    f(i)
            gph.replay()
            st.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(2_000_000)  # the host enqueues the launches while the device is busy: the events see kernels only
            e0.record(st)
            gph.replay()
            gph.replay()
            e1.record(st)
            st.synchronize()
        us = e0.elapsed_time(e1) / (2 * iters) * 1e3
        out[policy] = round(us, 2)
        del gph
    print(json.dumps({"case": name, "heads": f"{h}q/{kvh}kv d{d}", "batch": batch, "ctx": ctx, "kv": kv, "kv_MB": round(byt / 1e6, 1),
                      "us_by_policy": out, "GBps_best": round(byt / min(out.values()) / 1e3, 1)}), flush=True)
    del pools
    torch.cuda.empty_cache()


if __name__ == "__main__":
    A.load()
    only = sys.argv[1].split(",") if len(sys.argv) > 1 else None
    for c in CASES:
        if only is None or c[0] in only:
            bench(*c)
