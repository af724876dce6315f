"""Child process of tests/test_overlay.py: plays the part of a live `aphrodite._C` (defines the reference's schemas from
Python and registers stand-in kernels that only record that they were called), then loads this library as an overlay.
Mode "cpu": dispatcher plumbing only (CPU tensors must fall through to the stand-in). Mode "gpu": CUDA calls reach the
sm_100a kernels and match the oracle; a combination this library reports as unsupported reaches the stand-in."""
import json
import os
import re
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")]


def reference_schemas():
    """{namespace: [schema strings]} as torch_binding.cpp defines them (byte-identical to kernels/torch_bindings.cpp)."""
    src = open(os.path.join(ROOT, "aphrodite-fork_b200", "csrc", "torch_binding.cpp")).read()
    out = {}
    for ns, body in re.findall(r"TORCH_LIBRARY_FRAGMENT\((\w+), ops\) \{(.*?)\n\}", src, flags=re.S):
        defs = []
        for call in re.findall(r"ops\.def\((.*?)\);", body, flags=re.S):
            call = re.sub(r",\s*\{[^}]*\}\s*$", "", call.strip())  # drop a trailing tag list
            defs.append(re.sub(r"\s+", " ", "".join(re.findall(r'"((?:[^"\\]|\\.)*)"', call))).strip())
        out[ns] = defs
    return out


def main(mode):
    calls = []
    libs = []
    for ns, defs in reference_schemas().items():
        lib = torch.library.Library(ns, "DEF")
        libs.append(lib)
        for d in defs:
            name = d.split("(")[0]
            lib.define(d)

            def stand_in(*args, _name=name, **kw):
                calls.append(_name)
                if _name == "gptq_marlin_gemm":
                    return torch.zeros(args[8], args[9], dtype=args[0].dtype, device=args[0].device)
                if _name in ("gptq_marlin_repack", "awq_marlin_repack"):
                    return torch.zeros(1, dtype=torch.int32, device=args[0].device)
                return None

            for key in ("CPU", "CUDA"):
                lib.impl(name, stand_in, key)
    import aphrodite_fork_b200 as A
    res = {"mode": A.load()}
    ops = torch.ops._C
    if mode == "cpu":
        x = torch.randn(4, 64)
        ops.silu_and_mul(torch.empty(4, 32), x)
        torch.ops._C_cache_ops.convert_fp8(torch.empty(8, dtype=torch.uint8), torch.randn(8), 1.0, "fp8")
        res["stand_in_calls"] = calls
        print(json.dumps(res))
        return
    from helpers import check_close, make_attention_case
    from oracle import oracle
    dev = "cuda"
    S, H, KVH, D, bs, lens = 3, 16, 4, 128, 16, [700, 33, 1500]
    q, kc, vc, bt, sl = make_attention_case(S, H, KVH, D, bs, lens, torch.bfloat16, "fp8", 5)
    ref64, _ = oracle.paged_attention(q, kc, vc, KVH, D ** -0.5, bt, sl, bs, kv_cache_dtype="fp8")
    one = torch.tensor(1.0, device=dev)
    out = torch.full((S, H, D), float("nan"), dtype=torch.bfloat16, device=dev)
    args = (out, q.to(dev), kc.to(dev), vc.to(dev), KVH, D ** -0.5, bt.to(dev), sl.to(dev), bs, max(lens), None, "fp8", one, one, 0, 0, 0,
            64, 0)
    ops.paged_attention_v1(*args)
    torch.cuda.synchronize()
    res["attention_max_abs_err"] = check_close(out, ref64, torch.bfloat16, 1e-2, "overlay paged_attention_v1")
    res["stand_in_calls_after_supported_call"] = list(calls)
    # silu_and_mul through the overlay against torch
    x = torch.randn(5, 256, device=dev, dtype=torch.bfloat16)
    y = torch.empty(5, 128, device=dev, dtype=torch.bfloat16)
    ops.silu_and_mul(y, x)
    want = (torch.nn.functional.silu(x[:, :128].float()) * x[:, 128:].float())
    res["silu_max_abs_err"] = float((y.float() - want).abs().max())
    # a combination the reference serves and this library reports as unsupported: it must reach the reference's kernel
    n_before = len(calls)
    a = torch.zeros(4, 256, dtype=torch.float16, device=dev)
    e = torch.empty(0, dtype=torch.int32, device=dev)
    ws = torch.zeros(64 // 64 * 16, dtype=torch.int32, device=dev)
    bq = torch.zeros(256 // 16, 64 * 16 // 4, dtype=torch.int32, device=dev)
    sc = torch.ones(2, 64, dtype=torch.float16, device=dev)
    UINT8B128 = 1125899923621888
    # 8-bit weights with 64 output columns: this library's 8-bit path works on 128-column slices (APH_ERR_UNSUPPORTED)
    r = ops.gptq_marlin_gemm(a, bq, sc, e, e, e, ws, UINT8B128, 4, 64, 256, True, False, False, True, False)
    res["unsupported_call_reached"] = calls[n_before:]
    res["unsupported_call_shape"] = list(r.shape)
    # torch.library.opcheck of the hot ops' schemas (tests/kernels/utils.py:950-963 does the same in the reference)
    out2 = torch.empty_like(out)
    torch.library.opcheck(ops.paged_attention_v1.default, (out2,) + args[1:], test_utils=("test_schema",))
    kcd, vcd = args[2].clone(), args[3].clone()
    key = torch.randn(S, KVH, D, device=dev, dtype=torch.bfloat16)
    slots = torch.tensor([5, 77, 130], device=dev)
    torch.library.opcheck(torch.ops._C_cache_ops.reshape_and_cache.default, (key, key.clone(), kcd, vcd, slots, "fp8", one, one),
                          test_utils=("test_schema",))
    res["opcheck"] = "ok"
    print(json.dumps(res))


if __name__ == "__main__":
    main(sys.argv[1])
