"""Loads the in-tree native libraries. There is NO fallback: if the CUDA extension is missing
or cannot be loaded this raises, it never routes to PyTorch or the CPU oracle."""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# APH_B200_LIB_DIR: development only - a variant build of the same library (build.py build_variant)
_LIB = os.environ.get("APH_B200_LIB_DIR") or os.path.join(_HERE, "lib")
_loaded = None
_capi = None


def library_paths() -> dict:
    return {
        "c_abi": os.path.join(_LIB, "libaphrodite_b200.so"),
        "standalone": os.path.join(_LIB, "_C_b200.so"),
        "overlay": os.path.join(_LIB, "_C_b200_overlay.so"),
    }


def load() -> str:
    """Register the CUDA kernels with the torch dispatcher; returns the mode used.

    overlay    - aphrodite._C already defined the schemas: attach impl-only registrations that
                 override its CUDA kernels for the four hot-path ops;
    standalone - define the schemas ourselves (same strings as kernels/torch_bindings.cpp).
    """
    global _loaded
    if _loaded is not None:
        return _loaded
    paths = library_paths()
    for k, p in paths.items():
        if not os.path.exists(p):
            raise RuntimeError(
                f"aphrodite_fork_b200: native library missing ({p}); run `python -c 'import __graft_entry__ as g; "
                "g.build()'` - there is no non-CUDA fallback")
    have_ref = False
    try:
        have_ref = hasattr(torch.ops._C, "paged_attention_v1")
    except Exception:
        have_ref = False
    mode = "overlay" if have_ref else "standalone"
    torch.ops.load_library(paths[mode])
    _loaded = mode
    return mode


def c_api() -> ctypes.CDLL:
    """The raw C ABI (include/aphrodite_b200.h) for callers that bypass the torch dispatcher."""
    global _capi
    if _capi is None:
        p = library_paths()["c_abi"]
        if not os.path.exists(p):
            raise RuntimeError(f"aphrodite_fork_b200: {p} missing; build first (no fallback)")
        lib = ctypes.CDLL(p)
        lib.aph_b200_last_error.restype = ctypes.c_char_p
        lib.aph_paged_attention_workspace_bytes.restype = ctypes.c_size_t
        lib.aph_gptq_marlin_gemm_scratch_bytes.restype = ctypes.c_size_t
        _capi = lib
    return _capi
