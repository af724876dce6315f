"""Two Llama-3.1-8B-shaped layers of the decode step (batch 64 x ctx 4096), eager, for an ncu launch list:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/decode_step_profile.py
Also prints CUDA-graph time per layer without a profiler."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import aphrodite_fork_b200 as A
from aphrodite_fork_b200.decode import DecodeConfig, DecodeModel

A.load()
layers = int(os.environ.get("LAYERS", "2"))
model = DecodeModel(DecodeConfig.llama3_8b(num_layers=layers), batch=64, ctx=4096, seed=0)
for _ in range(3):
    model.step()
torch.cuda.synchronize()
if not os.environ.get("NO_GRAPH"):
    model.capture()
    for _ in range(3):
        model.step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        model.step()
    e1.record()
    torch.cuda.synchronize()
    print(f"graph: {e0.elapsed_time(e1) / 50 / layers * 1e3:.1f} us per layer ({layers} layers)")
