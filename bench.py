#!/usr/bin/env python
"""Headline benchmark: paged-attention decode KV GB/s on BASELINE.json configs[1] (C2).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the decode attention hot path over one batch: `paged_attention_v2`
(our kernels through the reference's op schema) over 64 sequences x 4096 context, 32 q / 8 kv
heads, head 128, FP8-E4M3 KV, block 16. `value` = algorithmic KV bytes (K and V, every kv head
once) / device time with everything resident in HBM; `e2e` = the same metric through the public
API (PagedAttention.write_to_paged_cache + forward_decode) with the step's inputs (query, new K/V
token, slot mapping, block tables, sequence lengths) starting in pinned HOST memory and the
output copied back, copies inside the timed region. The KV cache itself is device-resident state
(it never lives on the host in the reference either: worker/cache_engine.py).

Multi-GPU (torchrun, one rank per GPU): tensor parallelism as the reference shards this path
(modeling/models/llama.py:118-127): the GLOBAL C2 batch is fixed and rank r of N serves 32/N query heads and
max(1, 8/N) kv heads of every sequence from its own KV cache; attention has no exchange step, so
there is no collective in the timed region. value = global KV bytes (every kv head once) / max-over-ranks
time: STRONG scaling. The tensor-parallel decode step (GEMMs + all-reduces) is reported in `extras.decode`.

After the timed loop the last output of the timed kernel is checked against the CPU oracle on 4 sampled
sequences of the full problem (outside the timed region); a miss fails the run.

--impl reference: the reference's own CPU implementation of the path (oracle/_ref = unmodified
kernels/cpu/attention.cpp, bf16 KV because the CPU backend has no FP8) on the host cores, each step
a bounded sample of the workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, KVH, D, BS = 32, 8, 128, 16
BATCH, CTX = 64, 4096
METRIC = "paged-attn decode KV GB/s"
# dram__bytes_read.sum + dram__bytes_write.sum of one launch of this workload at N = 1, from an `ncu --set full` capture of
# THIS command line; not measured in the run, so the line names the file it comes from
NCU_TRAFFIC_BYTES = 546_121_472
NCU_TRAFFIC_SOURCE = ("profiles/r02_ncu_attention_c2_c4tp8.txt, prof_attn_c2_dense (ncu --set full of the split-K kernel at C2, the 4-CTA-per-SM "
                      "build the planner picks: 538.90 MB read + 7.22 MB written; the reduce kernel that follows re-reads the 7 MB of partials)")
WORKLOAD = "C2: paged_attention_v2 Llama-3-8B shape 32q/8kv d128, fp8_e4m3 KV, block 16, batch 64 x ctx 4096"


def kv_bytes(batch, ctx, kvh, d, esz):
    return 2 * batch * ctx * kvh * d * esz


def keep_device_busy(torch):
    """Queue about a millisecond of spinning on the current stream BEFORE the opening event of a device-timed region: the
    host then enqueues the event, the graph launch and the closing event while the GPU is still busy, and the events bracket
    the kernels of the K steps only. Without it the region also holds the host's graph-launch latency (tens of
    microseconds, during which the GPU idles) - 2.5 us per step when a driver asks for K = 20 steps of a 20 us kernel."""
    torch.cuda._sleep(2_000_000)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_reference_run(seconds_target, batch, ctx, threads=None):
    """Times the reference CPU backend (oracle/_ref) on a bounded sample: `batch` sequences of the C2
    shape with bf16 KV. Returns (GB/s, cores, kind, sample description, ms per call, calls)."""
    import torch
    from oracle import oracle
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = oracle.refcpu()
    kind = "reference" if ref is not None else "port"
    dt = torch.bfloat16
    nblk = batch * ctx // BS
    g = torch.Generator().manual_seed(0)
    sc = D ** -0.5
    q = ((torch.rand(batch, H, D, generator=g) * 2 - 1) * sc).to(dt)
    kc = ((torch.rand(nblk, KVH, D // 8, BS, 8, generator=g) * 2 - 1) * sc).to(dt)
    vc = ((torch.rand(nblk, KVH, D, BS, generator=g) * 2 - 1) * sc).to(dt)
    bt = torch.randperm(nblk, generator=g).to(torch.int32).reshape(batch, ctx // BS)
    sl = torch.full((batch,), ctx, dtype=torch.int32)
    out = torch.empty_like(q)
    one = torch.tensor(1.0)

    def call():
        if ref is not None:
            ref.paged_attention_v1(out, q, kc, vc, KVH, sc, bt, sl, BS, ctx, None, "auto", one, one, 0, 0, 0, 64, 0)
        else:
            oracle.paged_attention(q, kc, vc, KVH, sc, bt, sl, BS)

    call()
    t0 = time.perf_counter()
    n = 0
    while True:
        call()
        n += 1
        el = time.perf_counter() - t0
        if el >= seconds_target or n >= 2000:
            break
    gbs = kv_bytes(batch, ctx, KVH, D, 2) * n / el / 1e9
    sample = (f"{n} calls of {'kernels/cpu/attention.cpp paged_attention_v1 (oracle/_ref)' if ref is not None else 'oracle port'}"
              f" on {batch} seqs x {ctx} ctx, 32q/8kv d128, bf16 KV (CPU backend has no FP8), {cores} threads")
    return gbs, cores, kind, sample, el / n * 1e3, n


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warm = args.steps, args.warmup
    # bounded sample per step: 4 sequences of the C2 shape (67 MB of bf16 KV)
    import torch
    from oracle import oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sb = 4
    ref = oracle.refcpu()
    kind = "reference" if ref is not None else "port"
    dt = torch.bfloat16
    nblk = sb * CTX // BS
    g = torch.Generator().manual_seed(0)
    sc = D ** -0.5
    q = ((torch.rand(sb, H, D, generator=g) * 2 - 1) * sc).to(dt)
    kc = ((torch.rand(nblk, KVH, D // 8, BS, 8, generator=g) * 2 - 1) * sc).to(dt)
    vc = ((torch.rand(nblk, KVH, D, BS, generator=g) * 2 - 1) * sc).to(dt)
    bt = torch.randperm(nblk, generator=g).to(torch.int32).reshape(sb, CTX // BS)
    sl = torch.full((sb,), CTX, dtype=torch.int32)
    out = torch.empty_like(q)
    one = torch.tensor(1.0)

    def call():
        if ref is not None:
            ref.paged_attention_v1(out, q, kc, vc, KVH, sc, bt, sl, BS, CTX, None, "auto", one, one, 0, 0, 0, 64, 0)
        else:
            oracle.paged_attention(q, kc, vc, KVH, sc, bt, sl, BS)

    for _ in range(warm):
        call()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    el = time.perf_counter() - t0
    val = kv_bytes(sb, CTX, KVH, D, 2) * steps / el / 1e9
    sample = (f"each step = {sb} seqs x {CTX} ctx of the C2 shape with bf16 KV (the CPU backend has no FP8), "
              f"{'kernels/cpu/attention.cpp via oracle/_ref' if ref is not None else 'oracle port'}, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": round(el / steps * 1e3, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16 (f32 accumulate)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": round(val, 3), "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(val, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)



def gemm_extras(torch, A, dev, iters=30):
    """W4A16 GEMM on the C3 shapes (Llama-3.1-8B, g128, bf16): us/call under a CUDA graph with weights rotated through
    > 256 MB of distinct buffers, HBM GB/s on the algorithmic bytes and TFLOP/s. Secondary numbers of BASELINE's metric."""
    UINT4B8 = 1125899907892224
    shapes = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096),
              # C4 (Llama-3.1-70B, TP1 shapes of tests/benchmarks/kernels/benchmark_shapes.py:57-62): matrices large
              # enough (235 MB) that one launch can approach the HBM roofline at all
              "70b_gate_up": (8192, 57344), "70b_down": (28672, 8192)}
    hbm_peak, _ = peaks()
    tensor_peak = 1657.3
    try:
        tensor_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
    except Exception:
        pass
    res = {}
    gen = torch.Generator(device=dev).manual_seed(0)
    e = torch.empty(0, dtype=torch.int32, device=dev)
    for name, (K, N) in shapes.items():
        nbuf = max(2, int(300e6 // (K * N // 2)) + 1)
        bs = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), dtype=torch.int32, device=dev, generator=gen) for _ in range(nbuf)]
        ss = [(torch.rand(K // 128, N, device=dev, generator=gen) * 0.01 + 0.005).to(torch.bfloat16) for _ in range(nbuf)]
        ws = torch.zeros(N // 64 * 16, dtype=torch.int32, device=dev)
        for M in ((1, 16, 64, 256, 1024) if not name.startswith("70b") else (1, 16, 256)):
            a = torch.randn(M, K, device=dev, generator=gen).to(torch.bfloat16)
            f = lambda i: A.ops.gptq_marlin_gemm(a, bs[i % nbuf], ss[i % nbuf], e, e, e, ws, UINT4B8, M, N, K, True)
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
                keep_device_busy(torch)
                e0.record(st)
                gph.replay()
                e1.record(st)
                st.synchronize()
            us = e0.elapsed_time(e1) / iters * 1e3
            byt = K * N // 2 + (K // 128) * N * 2 + M * K * 2 + M * N * 2
            res[f"{name}_m{M}"] = {"us": round(us, 2), "GBps": round(byt / us / 1e3, 1), "TFLOPs": round(2 * M * N * K / us / 1e6, 1),
                                   "frac_hbm": round(byt / us / 1e3 / hbm_peak, 3),
                                   "frac_tensor": round(2 * M * N * K / us / 1e6 / tensor_peak, 3)}
        del bs, ss
    return res


def attention_extras(torch, A, dev, iters=20):
    """The other attention configs of BASELINE.json on one GPU, CUDA-graph replay over KV pools rotated through > 256 MB:
    C5 (Qwen2-72B shape at TP 8: 8 q / 1 kv heads per GPU, 32 768 ctx, FP8-E5M2 KV, batch 1 / 8 / 32) and C1's shape on
    the GPU (12 heads x 64, bf16 KV, 8 x 512; latency-bound)."""
    hbm_peak, _ = peaks()
    res = {}
    cases = [("C5_b1", 8, 1, 128, 1, 32768, "fp8_e5m2"), ("C5_b8", 8, 1, 128, 8, 32768, "fp8_e5m2"),
             ("C5_b32", 8, 1, 128, 32, 32768, "fp8_e5m2"), ("C4_tp8_b64", 8, 1, 128, 64, 4096, "fp8"),
             ("C1_gpu", 12, 12, 64, 8, 512, "auto")]
    g = torch.Generator(device=dev).manual_seed(0)
    one = torch.tensor(1.0, device=dev)
    for name, h, kvh, d, batch, ctx, kv in cases:
        esz = 2 if kv == "auto" else 1
        x = 16 // esz
        nblk = batch * ctx // BS
        byt = kv_bytes(batch, ctx, kvh, d, esz)
        npools = max(2, min(64, -(-(300 << 20) // byt)))
        pools = []
        for _ in range(npools):
            if kv == "auto":
                kc = (torch.rand(nblk, kvh, d // x, BS, x, generator=g, device=dev) - 0.5).to(torch.bfloat16)
                vc = (torch.rand(nblk, kvh, d, BS, generator=g, device=dev) - 0.5).to(torch.bfloat16)
            else:  # any byte below 0x78 is a finite value in both FP8 formats
                kc = torch.randint(0, 120, (nblk, kvh, d // x, BS, x), dtype=torch.uint8, device=dev, generator=g)
                vc = torch.randint(0, 120, (nblk, kvh, d, BS), dtype=torch.uint8, device=dev, generator=g)
            pools.append((kc, vc))
        q = ((torch.rand(batch, h, d, generator=g, device=dev) * 2 - 1) * d ** -0.5).to(torch.bfloat16)
        bt = torch.randperm(nblk, generator=g, device=dev).to(torch.int32).reshape(batch, ctx // BS)
        sl = torch.full((batch,), ctx, dtype=torch.int32, device=dev)
        f = lambda i: A.PagedAttention.forward_decode(q, pools[i % npools][0], pools[i % npools][1], bt, sl, ctx, kv, kvh,
                                                      d ** -0.5, None, one, one)
        for i in range(3):
            f(i)
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph, stream=st):
                for i in range(iters):
                    f(i)
            for _ in range(3):  # these graphs run for well under a millisecond: let the clocks settle after the allocations
                gph.replay()
            st.synchronize()
            times = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                keep_device_busy(torch)
                e0.record(st)
                gph.replay()
                e1.record(st)
                st.synchronize()
                times.append(e0.elapsed_time(e1))
        us = sorted(times)[len(times) // 2] / iters * 1e3  # median of five replays
        res[name] = {"heads": f"{h}q/{kvh}kv d{d}", "batch": batch, "ctx": ctx, "kv": kv, "us": round(us, 2),
                     "kv_bytes": byt, "GBps": round(byt / us / 1e3, 1), "frac_hbm": round(byt / us / 1e3 / hbm_peak, 3)}
        del pools, gph
        torch.cuda.empty_cache()
    return res


def c1_cpu_decode_loop(seconds_target=6.0):
    """BASELINE.md 4.2 (b), config C1 on the host cores: a 12-layer OPT-125m-shaped decode loop with random weights -
    torch CPU linears + the reference CPU backend's reshape_and_cache and paged_attention_v1 (oracle/_ref), bf16,
    block 16, 8 sequences x 512 context, 12 heads x 64."""
    import torch
    from oracle import oracle
    ref = oracle.refcpu()
    if ref is None:
        return {"unavailable": "oracle/_ref (reference CPU backend) not built"}
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    L, hid, ffn, h, d, batch, ctx = 12, 768, 3072, 12, 64, 8, 512
    dt = torch.bfloat16
    g = torch.Generator().manual_seed(0)
    nblk = batch * ctx // BS
    mk = lambda *shape: (torch.randn(*shape, generator=g) * 0.02).to(dt)
    layers = [dict(qkv=mk(3 * hid, hid), o=mk(hid, hid), up=mk(ffn, hid), down=mk(hid, ffn),
                   kc=mk(nblk, h, d // 8, BS, 8), vc=mk(nblk, h, d, BS)) for _ in range(L)]
    bt = torch.randperm(nblk, generator=g).to(torch.int32).reshape(batch, ctx // BS)
    sl = torch.full((batch,), ctx, dtype=torch.int32)
    slots = bt[:, -1].to(torch.int64) * BS + (BS - 1)
    one = torch.tensor(1.0)
    x0 = mk(batch, hid)
    ln = torch.nn.functional.layer_norm

    def step():
        x = x0
        for w in layers:
            hdn = ln(x, (hid,))
            qkv = torch.nn.functional.linear(hdn, w["qkv"])
            q, k, v = qkv.split(hid, dim=-1)
            ref.reshape_and_cache(k.reshape(batch, h, d), v.reshape(batch, h, d), w["kc"], w["vc"], slots, "auto", one, one)
            a = torch.empty(batch, h, d, dtype=dt)
            ref.paged_attention_v1(a, q.reshape(batch, h, d).contiguous(), w["kc"], w["vc"], h, d ** -0.5, bt, sl, BS, ctx, None,
                                   "auto", one, one, 0, 0, 0, 64, 0)
            x = x + torch.nn.functional.linear(a.reshape(batch, hid), w["o"])
            hdn = ln(x, (hid,))
            x = x + torch.nn.functional.linear(torch.relu(torch.nn.functional.linear(hdn, w["up"])), w["down"])
        return x

    step()
    t0, n = time.perf_counter(), 0
    while True:
        step()
        n += 1
        el = time.perf_counter() - t0
        if el >= seconds_target or n >= 500:
            break
    ms = el / n * 1e3
    kvb = kv_bytes(batch, ctx, h, d, 2) * L
    return {"config": "C1: OPT-125m shape, 12 layers, bf16, block 16, 8 seqs x 512 ctx, reference CPU backend (oracle/_ref) + torch CPU linears",
            "cores": cores, "steps": n, "ms_per_step": round(ms, 3), "tok_per_s": round(batch / ms * 1e3, 1),
            "kv_GBps": round(kvb / (ms * 1e-3) / 1e9, 2)}


def decode_extras(torch, A, dev, world, steps=10):
    """Decode tok/s of a Llama layer stack built from this library's kernels only (+ the all-reduce when world > 1),
    one CUDA graph per step. 1 GPU: Llama-3.1-8B shapes (C3) at batch 1 / 16 / 64 / 256; N GPUs: Llama-3.1-70B shapes at
    TP = N (C4), batch 64. RMSNorm, RoPE and SiLU-mul are the library's own kernels, see
    aphrodite-fork_b200/decode.py. `roofline` = mandatory HBM bytes of a step on one GPU (this rank's weights + KV, each
    once) / step time against the measured HBM peak."""
    import torch.distributed as dist
    from aphrodite_fork_b200.decode import DecodeConfig, DecodeModel
    from aphrodite_fork_b200.tp import TPGroup
    cfg = DecodeConfig.llama3_8b() if world == 1 else DecodeConfig.llama3_70b()
    hbm_peak, _ = peaks()
    tp = TPGroup()

    def measure(model):
        model.capture()
        for _ in range(3):
            model.step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            model.step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        byt = model.kv_bytes_per_step() + model.weight_bytes()
        return {"batch": model.batch, "ctx": model.ctx, "ms_per_step": round(ms, 3), "tok_per_s": round(model.batch / (ms * 1e-3), 1),
                "kernel_launches_per_step": model.launches_per_step(), "hbm_bytes_per_step_per_gpu": byt,
                "roofline": {"bound": "hbm", "achieved": round(byt / (ms * 1e-3) / 1e9, 1), "peak": hbm_peak, "unit": "GB/s",
                             "frac": round(byt / (ms * 1e-3) / 1e9 / hbm_peak, 4)}}

    main = DecodeModel(cfg, batch=BATCH, ctx=CTX, tp=tp, device=dev, seed=0)
    res = measure(main)
    res.update({"model": ("llama-3.1-8b" if world == 1 else "llama-3.1-70b") + f" shapes, {cfg.num_layers} layers, W4A16 g128, fp8 KV",
                "tp": world,
                "all_reduce": ("none" if world == 1 else f"NCCL ({main.tp.custom_error})" if main.tp.custom is None
                               else main.tp.describe(main.x) if hasattr(main.tp, "describe")
                               else "one-shot peer-memory kernel (csrc/all_reduce.cu)" if main.tp.custom.should_use(main.x)
                               else "NCCL (message above the one-shot threshold of this world size)")})
    if world == 1:  # C3: "decode loop, batch 1-256"
        sweep = {}
        for b in (1, 16, 256):
            main.graph = None
            m = DecodeModel(cfg, batch=b, ctx=CTX, tp=tp, device=dev, seed=1, weights_from=main)
            r = measure(m)
            sweep[f"batch{b}"] = {k: r[k] for k in ("ms_per_step", "tok_per_s", "hbm_bytes_per_step_per_gpu", "roofline")}
            del m
            torch.cuda.empty_cache()
        sweep[f"batch{BATCH}"] = {k: res[k] for k in ("ms_per_step", "tok_per_s", "hbm_bytes_per_step_per_gpu", "roofline")}
        res["batch_sweep"] = sweep
    return res


def check_parity(torch, out, q, pool, bt, sl, kvh, scale, seqs=(0, 21, 42, 63), atol=1e-2, rtol=1.6e-2):
    """The timed kernel's output for `seqs` against oracle/ (the checker, never the thing measured): the sequences'
    KV blocks are gathered into a compact CPU cache, attention is evaluated in double over the FP8 cache dequantised
    the way the reference reads it. Bound: |out - truth| <= atol + rtol |truth| (reference tests' atol for FP8 KV,
    torch's bf16 rtol) - the bound of tests/test_gpu_attention.py."""
    from oracle import oracle
    kc, vc = pool
    idx = torch.tensor(list(seqs), device=out.device)
    blocks = bt[idx].to(torch.int64)                      # [n, blocks per seq]
    n, per = blocks.shape
    kc_s = kc[blocks.reshape(-1)].cpu()
    vc_s = vc[blocks.reshape(-1)].cpu()
    bt_s = torch.arange(n * per, dtype=torch.int32).reshape(n, per)
    ref64, _ = oracle.paged_attention(q[idx].cpu(), kc_s, vc_s, kvh, scale, bt_s, sl[idx].cpu(), BS, kv_cache_dtype="fp8")
    got = out[idx].cpu().double()
    err = (got - ref64).abs()
    return {"checked_seqs": len(seqs), "checked_values": int(err.numel()), "max_abs": float(err.max()),
            "worst_excess": float((err - (atol + rtol * ref64.abs())).max()), "bound": f"{atol} + {rtol}*|truth|",
            "against": "oracle/ (C restatement, fp64) on the full-size problem, last timed step"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import aphrodite_fork_b200 as A

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    A.load()
    steps, warm = args.steps, max(args.warmup, 3)

    # ---- tensor-parallel shard of C2 (llama.py:118-127): fixed global batch, heads split over the ranks ----
    assert H % world == 0 and (KVH % world == 0 or world % KVH == 0), f"{world} ranks cannot shard {H}q/{KVH}kv heads"
    Hr, KVHr = H // world, max(1, KVH // world)
    kv_replicas = max(1, world // KVH)  # ranks holding the same kv head (none up to 8 ranks)
    if world == 1 and os.environ.get("APH_BENCH_SHARD"):  # development: one rank's shard of an N-GPU run on a single GPU
        Hr, KVHr = H // int(os.environ["APH_BENCH_SHARD"]), max(1, KVH // int(os.environ["APH_BENCH_SHARD"]))
    # ---- synthetic inputs (seed 0): per pool 16384 distinct blocks; as many pools as it takes to rotate through
    #      >= 1 GiB of KV, so no launch finds its KV in the 126 MB L2 (a rank's pool is only 64 MiB at N = 8) ----
    g = torch.Generator(device=dev).manual_seed(0 + rank)
    nblk = BATCH * CTX // BS
    sc = D ** -0.5
    one = torch.tensor(1.0, device=dev)
    pool_bytes = kv_bytes(BATCH, CTX, KVHr, D, 1)
    npools = max(2, -(-(1 << 30) // pool_bytes))
    npools += npools % 2  # even: the end-to-end leg's two lanes take alternate pools
    pools = []
    for _ in range(npools):
        src = ((torch.rand(nblk, KVHr, D // 16, BS, 16, generator=g, device=dev) * 2 - 1) * sc).to(torch.bfloat16)
        kc = torch.empty(src.shape, dtype=torch.uint8, device=dev)
        A.ops.convert_fp8(kc, src, 1.0, "fp8")
        src = ((torch.rand(nblk, KVHr, D, BS, generator=g, device=dev) * 2 - 1) * sc).to(torch.bfloat16)
        vc = torch.empty(src.shape, dtype=torch.uint8, device=dev)
        A.ops.convert_fp8(vc, src, 1.0, "fp8")
        pools.append((kc, vc))
    del src
    q = ((torch.rand(BATCH, Hr, D, generator=g, device=dev) * 2 - 1) * sc).to(torch.bfloat16)
    bt = torch.randperm(nblk, generator=g, device=dev).to(torch.int32).reshape(BATCH, CTX // BS)
    sl = torch.full((BATCH,), CTX, dtype=torch.int32, device=dev)
    out = torch.empty_like(q)
    P = (CTX + 511) // 512
    exp_sums = torch.empty(BATCH, Hr, P, dtype=torch.float32, device=dev)
    max_logits = torch.empty_like(exp_sums)
    tmp_out = torch.empty(BATCH, Hr, P, D, dtype=torch.bfloat16, device=dev)

    def step(i):
        kc, vc = pools[i % npools]
        A.ops.paged_attention_v2(out, exp_sums, max_logits, tmp_out, q, kc, vc, KVHr, sc, bt, sl, BS, CTX, None, "fp8",
                                 one, one)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(warm):
        step(i)
    # the K timed steps are replayed from one CUDA graph (the reference replays captured decode steps the same way):
    # a rank's kernel is only ~20 us at 8 GPUs, eager launches from Python would measure the host, not the device
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        timed_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(timed_graph, stream=side):
            for i in range(steps):
                step(i)
    torch.cuda.current_stream().wait_stream(side)
    timed_graph.replay()  # untimed: K more warm-up steps
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    keep_device_busy(torch)  # not a step: the host gets ahead of the device, so that ev0 .. ev1 holds the K steps' kernels only
    ev0.record()
    timed_graph.replay()  # exactly K steps
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # algorithmic bytes: K and V of every kv head of the GLOBAL problem once (a kv head replicated over several ranks
    # still counts once); what one rank's launch streams is pool_bytes
    # kernels of this repo per step: the attention kernel, plus the split-K reduce kernel unless the partitions are merged
    # on chip by thread-block clusters (no workspace needed then)
    capi = A.c_api()
    launches_per_step = 2 if capi.aph_paged_attention_workspace_bytes(BATCH, Hr, KVHr, D, CTX) > 0 else 1
    per_step_bytes = kv_bytes(BATCH, CTX, KVH, D, 1)
    value = per_step_bytes * steps / (ms * 1e-3) / 1e9
    per_gpu = pool_bytes * steps / (ms * 1e-3) / 1e9

    # ---- parity of the TIMED kernel on the FULL problem (outside the timed region): the output of the last timed
    #      step against the CPU oracle on 4 sampled sequences; tolerance of tests/test_gpu_attention.py for FP8 KV ----
    parity = check_parity(torch, out, q, pools[(steps - 1) % npools], bt, sl, KVHr, sc)
    if world > 1:
        t = torch.tensor([parity["max_abs"], parity["worst_excess"]], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        parity["max_abs"], parity["worst_excess"] = float(t[0]), float(t[1])
        parity["ranks_checked"] = world
    parity["ok"] = parity["worst_excess"] <= 0
    if not parity["ok"]:
        raise RuntimeError(f"bench.py: the timed kernel's output misses the oracle: {parity}")

    # ---- e2e: public API, host-resident step inputs, copies inside the timed region ----
    # Every step copies ITS inputs (query, new K/V token, block tables, sequence lengths, slot mapping) from pinned host
    # memory and copies ITS output back. The way a serving loop drives the API: the step's inputs are packed in one
    # pinned staging buffer (one H2D copy), and three streams (copy-in / compute / copy-out) with two device staging
    # buffers let step i+1's upload and step i-1's download run under step i's kernels.
    def seg(nbytes):
        return (nbytes + 255) // 256 * 256

    parts = [("q", (BATCH, Hr, D), torch.bfloat16), ("newkv", (BATCH, 2, KVHr, D), torch.bfloat16),
             ("bt", (BATCH, CTX // BS), torch.int32), ("sl", (BATCH,), torch.int32), ("slots", (BATCH,), torch.int64)]
    offs, total_in = {}, 0
    for name, shape, dt in parts:
        n = int(torch.tensor(shape).prod()) * torch.tensor([], dtype=dt).element_size()
        offs[name] = (total_in, n, shape, dt)
        total_in += seg(n)

    def views(buf):
        return {name: buf[o:o + n].view(dt).view(shape) for name, (o, n, shape, dt) in offs.items()}

    in_h = torch.empty(total_in, dtype=torch.uint8, pin_memory=True)
    hv = views(in_h)
    hv["q"].copy_(q.cpu())
    hv["newkv"].copy_(((torch.rand(BATCH, 2, KVHr, D) * 2 - 1) * sc).to(torch.bfloat16))
    hv["bt"].copy_(bt.cpu())
    hv["sl"].copy_(sl.cpu())
    hv["slots"].copy_(hv["bt"][:, -1].to(torch.int64) * BS + (BS - 1))  # slot of token ctx-1 of every sequence
    NBUF = 2
    in_d = [torch.empty(total_in, dtype=torch.uint8, device=dev) for _ in range(NBUF)]
    dv = [views(b) for b in in_d]
    out_h = [torch.empty(BATCH, Hr, D, dtype=torch.bfloat16, pin_memory=True) for _ in range(NBUF)]
    h2d = sum(n for _, (o, n, shape, dt) in offs.items())
    d2h = out_h[0].numel() * out_h[0].element_size()
    # One step = upload of ITS inputs, write_to_paged_cache + forward_decode, download of ITS output. The way the
    # reference serves decode (worker/model_runner.py:1487-1646 replays one captured CUDA graph per step), every step is
    # one graph replay; two lanes (stream + staging buffers each) take alternate steps so that a step's copies run under
    # its neighbour's kernels. Consecutive steps use different KV pools, so the lanes never touch the same cache.
    lanes = [torch.cuda.Stream(device=dev) for _ in range(NBUF)]
    s_main = torch.cuda.current_stream()

    def e2e_body(b, pool):
        kc, vc = pools[pool]
        in_d[b].copy_(in_h, non_blocking=True)
        v = dv[b]
        A.PagedAttention.write_to_paged_cache(v["newkv"][:, 0], v["newkv"][:, 1], kc, vc, v["slots"], "fp8", one, one)
        o = A.PagedAttention.forward_decode(v["q"], kc, vc, v["bt"], v["sl"], CTX, "fp8", KVHr, sc, None, one, one)
        out_h[b].copy_(o, non_blocking=True)

    graphs = {}
    for pool in range(npools):  # npools is even: pool parity = lane
        b = pool % NBUF
        lanes[b].wait_stream(s_main)
        with torch.cuda.stream(lanes[b]):
            e2e_body(b, pool)  # warm-up outside capture
            lanes[b].synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=lanes[b]):
                e2e_body(b, pool)
        graphs[pool] = g

    def e2e_run(n):
        ev = torch.cuda.Event()
        ev.record(s_main)
        for ln in lanes:
            ln.wait_event(ev)
        for i in range(n):
            with torch.cuda.stream(lanes[i % NBUF]):
                graphs[i % npools].replay()
        for ln in lanes:
            s_main.wait_stream(ln)  # the last downloads are inside the timed region

    e2e_run(warm)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_run(steps)
    e1.record()
    barrier()
    ems = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ems], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ems = float(t.item())
    e2e_val = per_step_bytes * steps / (ems * 1e-3) / 1e9

    # ---- secondary numbers of BASELINE's metric: W4A16 GEMM TFLOP/s (rank 0) and decode tok/s (all ranks) ----
    extras = {}
    if not args.no_extras:
        del pools
        torch.cuda.empty_cache()
        try:
            if rank == 0:
                extras["gemm_w4a16"] = gemm_extras(torch, A, dev)
                if world == 1:
                    extras["attention_configs"] = attention_extras(torch, A, dev)
            barrier()
            torch.cuda.empty_cache()
            extras["decode"] = decode_extras(torch, A, dev, world)
        except Exception as ex:  # the headline line must still be printed
            extras["error"] = f"{type(ex).__name__}: {ex}"[:300]

    if rank == 0:
        peak, peak_src = peaks()
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": steps,
            "warmup": warm, "ms_per_step": round(ms / steps, 5), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16 mma / f32 accumulate over fp8_e4m3 KV, bf16 q/out", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "parallelism": f"tp{world}: fixed global batch {BATCH} x {CTX}, {Hr} q / {KVHr} kv heads per rank"
                                      + (f" (each kv head on {kv_replicas} ranks)" if kv_replicas > 1 else "")
                                      + ", no collective in attention (llama.py:118-127)",
                       "l2": f"{npools} FP8 KV pools of {pool_bytes >> 20} MiB per rank in rotation (>= 1 GiB), block tables = random "
                             "permutation of distinct blocks: inputs larger than the 126 MB L2",
                       "timing": "CUDA events on the launching stream around ONE graph replay of the K steps, max over ranks; "
                                 "about 1 ms of device-side spinning is queued before the opening event so that the host's "
                                 "graph-launch latency is not inside the timed region",
                       "kv_bytes_per_step": per_step_bytes, "kv_bytes_per_step_per_gpu": pool_bytes},
            "roofline": {"bound": "hbm", "achieved": round(per_gpu, 2), "peak": peak, "unit": "GB/s",
                         "frac": round(per_gpu / peak, 4), "traffic": NCU_TRAFFIC_BYTES if world == 1 else None,
                         "traffic_source": NCU_TRAFFIC_SOURCE if world == 1 else None,
                         "note": f"per GPU: algorithmic KV bytes of one rank's launch / CUDA-event time of the whole op per "
                                 f"launch (max over ranks); peak = {peak_src}"},
            "parity": parity,
            "e2e": {"value": round(e2e_val, 2), "unit": "GB/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                    "ms_per_step": round(ems / steps, 5),
                    "api": "PagedAttention.write_to_paged_cache + forward_decode; every step uploads its q / new K,V token / "
                           "block tables / seq lens / slot mapping from ONE packed pinned host buffer and downloads its "
                           "output; a step is one CUDA-graph replay (copies included), two lanes take alternate steps"},
            "gpu_launches": launches_per_step * steps,
            "clocks": clocks,
            "extras": extras,
        }
        if world == 1 and not args.no_cpu:
            gbs, cores, kind, sample, cms, n = cpu_reference_run(12.0, 4, CTX)
            line["cpu_baseline"] = {"value": round(gbs, 3), "unit": "GB/s", "cores": cores, "kind": kind,
                                    "sample": sample, "ms_per_call": round(cms, 3)}
            if not args.no_extras:
                try:
                    extras["c1_cpu_decode_loop"] = c1_cpu_decode_loop()
                except Exception as ex:
                    extras["c1_cpu_decode_loop"] = {"error": f"{type(ex).__name__}: {ex}"[:200]}
        if extras:  # the secondary legs also as their own stderr lines (a long stdout line gets truncated in logs)
            for k, v in extras.items():
                print(json.dumps({"extra": k, "n_gpus": world, "result": v}), file=sys.stderr, flush=True)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the GEMM / decode-loop secondary measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
