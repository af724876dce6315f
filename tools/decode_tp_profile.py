"""Development (torchrun, one rank per GPU): the tensor-parallel decode step (Llama-3.1-70B shapes, LAYERS layers, batch
64 x ctx 4096, one CUDA graph per step) under each all-reduce policy - NCCL, one-shot and two-shot peer-memory kernels,
and the default size-based choice. Rank 0 prints one JSON line per policy with the microseconds per layer.
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 tools/decode_tp_profile.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import aphrodite_fork_b200 as A
from aphrodite_fork_b200.decode import DecodeConfig, DecodeModel
from aphrodite_fork_b200.tp import TPGroup


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    A.load()
    layers = int(os.environ.get("LAYERS", "16"))
    cfg = DecodeConfig.llama3_70b(num_layers=layers)
    base = None
    only = os.environ.get("POLICIES")
    for name, env in (("default", {}), ("one_shot", {"APH_AR_MODE": "one_shot"}), ("two_shot", {"APH_AR_MODE": "two_shot"}),
                      ("multimem", {"APH_AR_MODE": "multimem"}), ("ll", {"APH_AR_MODE": "ll"}), ("unfused_norm", {"APH_FUSED_AR_NORM": "0"}), ("nccl", {"APH_CUSTOM_AR": "0"})):
        if only and name not in only.split(","):
            continue
        for k in ("APH_AR_TWO_SHOT_BYTES", "APH_AR_MODE", "APH_FUSED_AR_NORM", "APH_CUSTOM_AR"):
            os.environ.pop(k, None)
        os.environ.update(env)
        model = DecodeModel(cfg, batch=64, ctx=4096, tp=TPGroup(), device=dev, seed=0, weights_from=base)
        base = base or model
        model.capture()
        for _ in range(3):
            model.step()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            model.step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 20], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"policy": name, "tp": world, "layers": layers, "ms_per_step": round(float(t), 3),
                              "us_per_layer": round(float(t) * 1e3 / layers, 1), "ms_per_80_layers": round(float(t) * 80 / layers, 2)}), flush=True)
        if model is not base:
            model.graph = None
            del model
        torch.cuda.synchronize()
    base.graph = None
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
