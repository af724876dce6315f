"""Development microbench (torchrun, one rank per GPU): the row-parallel outputs' all-reduce of a decode step in
isolation - NCCL, the one-shot, two-shot and (where the box has multicast support) multimem kernels (csrc/all_reduce.cu), alone and fused with the
residual add + RMSNorm - for messages of rows x 8192 bf16. CUDA-graph replay of 50 calls, max over ranks; rank 0 prints
one JSON line per size with the microseconds per call and the NVLink egress per rank each variant implies.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/bench_all_reduce.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import aphrodite_fork_b200 as A
from aphrodite_fork_b200 import _custom_ops as ops
from aphrodite_fork_b200.tp import TPGroup


def timed(fn, iters=50):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(2_000_000)  # the host enqueues the launches while the device is busy: the events see kernels only
    e0.record()
    g.replay()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (2 * iters) * 1e3], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del g
    return round(float(t.item()), 2)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    A.load()
    tp = TPGroup()
    assert tp.enable_custom_all_reduce(dev), tp.custom_error
    hidden = 8192
    w = torch.ones(hidden, dtype=torch.bfloat16, device=dev)
    for rows in (1, 4, 16, 64, 128, 256):
        x = torch.randn(rows, hidden, device=dev).to(torch.bfloat16) * 0.01
        res = torch.zeros_like(x)
        nbytes = x.numel() * 2
        out = {"rows": rows, "KiB": nbytes // 1024, "world": world}
        out["nccl"] = timed(lambda: dist.all_reduce(x))
        out["nccl_plus_norm"] = timed(lambda: (dist.all_reduce(x), ops.fused_add_rms_norm(x, res, w, 1e-5)))
        modes = ("one_shot", "two_shot", "ll") + (("multimem",) if tp.custom.multicast_ptr and os.environ.get("WITH_MULTIMEM") else ())
        for name in modes:
            os.environ["APH_AR_MODE"] = name.split("_32")[0]
            tp.custom.max_blocks = 32 if name.endswith("32cta") else 0
            x.normal_(0, 0.01)
            out[name] = timed(lambda: tp.custom.all_reduce(x))
            x.normal_(0, 0.01)
            res.zero_()
            out[name + "_fused_norm"] = timed(lambda: tp.custom.all_reduce_add_rms_norm(x, res, w, 1e-5))
        os.environ.pop("APH_AR_MODE", None)
        tp.custom.max_blocks = 0
        out["default_mode"] = tp.custom.mode(nbytes)
        out["egress_KiB_per_rank"] = {"one_shot": nbytes * (world - 1) // 1024, "two_shot": 2 * nbytes * (world - 1) // world // 1024,
                                      "ll": 4 * nbytes * (world - 1) // world // 1024, "multimem": nbytes // 1024}
        if rank == 0:
            print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
