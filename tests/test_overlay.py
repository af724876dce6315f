"""The overlay flavour of the binding (lib/_C_b200_overlay.so) loaded over a process that already defines the reference's
`_C` / `_C_cache_ops` schemas - the drop-in seam of SURVEY.md 8b (kernels/torch_bindings.cpp:19,678;
aphrodite/_custom_ops.py:13-17). aphrodite._C itself cannot be imported here, so tests/overlay_child.py stands in for it:
it defines the same schema strings from Python and registers stand-in kernels. Each case runs in its own process (the
test process may already have loaded the standalone flavour, which defines the schemas itself)."""
import json
import os
import subprocess
import sys

import pytest

CHILD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "overlay_child.py")


def run_child(mode):
    r = subprocess.run([sys.executable, CHILD, mode], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def test_overlay_loads_over_existing_schemas_and_passes_non_cuda_calls_through():
    res = run_child("cpu")
    assert res["mode"] == "overlay"
    # CPU tensors are not this library's business: the calls must reach the kernels that were registered before it
    assert res["stand_in_calls"] == ["silu_and_mul", "convert_fp8"]


@pytest.mark.gpu
def test_overlay_serves_cuda_calls_and_falls_back_for_unsupported_arguments():
    res = run_child("gpu")
    assert res["mode"] == "overlay"
    assert res["stand_in_calls_after_supported_call"] == [], "a supported CUDA call must not reach the reference's kernel"
    assert res["attention_max_abs_err"] < 1e-2 and res["silu_max_abs_err"] < 0.07
    assert res["unsupported_call_reached"] == ["gptq_marlin_gemm"] and res["unsupported_call_shape"] == [4, 64]
    assert res["opcheck"] == "ok"
