"""In-tree build of the sm_100a CUDA library and the torch binding.

  lib/libaphrodite_b200.so   nvcc -gencode arch=compute_100a,code=sm_100a, torch-free C ABI
  lib/_C_b200.so             torch binding, standalone flavour (defines the four schemas)
  lib/_C_b200_overlay.so     torch binding, overlay flavour (impl-only, over aphrodite._C)

Called by __graft_entry__.build(); cross-compiles without a GPU. The built files are
git-ignored but travel to the GPU box.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build", "fast" if os.environ.get("APH_FAST_BUILD", "0") == "1" else "full")
LIB = os.path.join(HERE, "lib")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
if os.environ.get("APH_GEMM_TRACE_BUILD"):  # development: per-stage clock stamps in the tcgen05 GEMM (tools/trace_gemm.py)
    NVCC_FLAGS.append("-DAPH_GEMM_TRACE")
HEAD_SIZES = [32, 64, 80, 96, 112, 120, 128, 192, 256]
FAST = os.environ.get("APH_FAST_BUILD", "0") == "1"


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: " + " ".join(cmd) + "\n" + r.stdout)


def _headers() -> list[str]:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    return hs + [os.path.join(INCLUDE, "aphrodite_b200.h"), os.path.abspath(__file__)]


def build(verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIB, exist_ok=True)
    hdrs = _headers()
    extra = ["-DAPH_FAST_BUILD"] if FAST else []
    jobs = []  # (object, command, deps)
    for name in ("runtime", "cache_kernels", "paged_attention", "marlin_gemm", "marlin_repack", "layer_glue", "all_reduce"):
        src = os.path.join(CSRC, name + ".cu")
        obj = os.path.join(OBJ, name + ".o")
        jobs.append((obj, [NVCC, *NVCC_FLAGS, *extra, "-c", src, "-o", obj], [src, *hdrs]))
    inst = os.path.join(CSRC, "paged_attention_inst.cu")
    for d in ([64, 128] if FAST else HEAD_SIZES):
        obj = os.path.join(OBJ, f"paged_attention_d{d}.o")
        jobs.append((obj, [NVCC, *NVCC_FLAGS, *extra, f"-DAPH_PA_HEAD_SIZE={d}", "-c", inst, "-o", obj], [inst, *hdrs]))
    todo = [(o, c) for o, c, deps in jobs if _newer(o, deps)]
    if verbose and todo:
        print(f"[build] compiling {len(todo)} CUDA objects", file=sys.stderr)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
        list(ex.map(lambda oc: _run(oc[1]), todo))
    objs = [o for o, _, _ in jobs]
    libpath = os.path.join(LIB, "libaphrodite_b200.so")
    stamp = os.path.join(LIB, ".mode")
    mode = "fast" if FAST else "full"
    if _newer(libpath, objs) or not os.path.exists(stamp) or open(stamp).read() != mode:
        open(stamp, "w").write(mode)
        _run([NVCC, "-shared", "-o", libpath, *objs, "-gencode", "arch=compute_100a,code=sm_100a"])

    # torch binding (host-only C++)
    import torch
    from torch.utils import cpp_extension as ce
    binding = os.path.join(CSRC, "torch_binding.cpp")
    inc = [f"-I{p}" for p in ce.include_paths(device_type="cuda")] + [f"-I{sysconfig.get_paths()['include']}"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    # NOT $CXX: the image exports CXX=/opt/gcc/bin/g++, a wrapper that links libstdc++ statically; a second,
    # uninitialised copy of the C++ runtime inside the torch process segfaults on the first ostream << int
    # (seen as a crash inside TORCH_CHECK message formatting). Use the system driver.
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    common = [cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-D_GLIBCXX_USE_CXX11_ABI=1", "-w", *inc, binding,
              f"-L{LIB}", "-laphrodite_b200", f"-L{tlib}", "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10",
              "-lc10_cuda", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
    bjobs = []
    for name, flags in (("_C_b200.so", ["-DAPH_DEFINE_SCHEMAS"]), ("_C_b200_overlay.so", [])):
        out = os.path.join(LIB, name)
        if _newer(out, [binding, os.path.join(INCLUDE, "aphrodite_b200.h"), os.path.abspath(__file__)]):
            bjobs.append(common + flags + ["-o", out])
    if verbose and bjobs:
        print(f"[build] compiling {len(bjobs)} torch bindings", file=sys.stderr)
    with ThreadPoolExecutor(max_workers=2) as ex:
        list(ex.map(_run, bjobs))
    return libpath


# translation units a variant build may recompile: (object stem, source, extra flags, flag prefixes that select it)
_VARIANT_UNITS = (
    ("marlin_gemm", "marlin_gemm.cu", [], ("-DAPH_RING", "-DAPH_GEMM")),
    ("paged_attention", "paged_attention.cu", [], ("-DAPH_PA_",)),
    ("paged_attention_d128", "paged_attention_inst.cu", ["-DAPH_PA_HEAD_SIZE=128"], ("-DAPH_PA_",)),
)


def build_variant(name: str, extra_flags: list[str]) -> str:
    """Development: the library with the translation units `extra_flags` address recompiled with them, in
    aphrodite-fork_b200/lib_<name>/ (libaphrodite_b200.so + copies of the torch bindings, which find their sibling
    through an $ORIGIN rpath). Select it with APH_B200_LIB_DIR. Used by tests/tools/ring_fault_proof.py
    (-DAPH_RING_FAULT, -DAPH_RING_LEGACY), the GEMM trace (-DAPH_GEMM_TRACE) and attention experiments (any -DAPH_PA_...
    flag recompiles paged_attention.cu and the head-size-128 kernels)."""
    import shutil
    build()
    out_dir = os.path.join(HERE, "lib_" + name)
    var_obj = os.path.join(OBJ, "variants", name)
    os.makedirs(out_dir, exist_ok=True)
    os.makedirs(var_obj, exist_ok=True)
    rebuilt, replaced = [], set()
    for stem, src_name, more, prefixes in _VARIANT_UNITS:
        if not any(f.startswith(prefixes) for f in extra_flags):
            continue
        src = os.path.join(CSRC, src_name)
        obj = os.path.join(var_obj, stem + ".o")
        if _newer(obj, [src, *_headers()]):
            _run([NVCC, *NVCC_FLAGS, *extra_flags, *more, "-c", src, "-o", obj])
        rebuilt.append(obj)
        replaced.add(stem + ".o")
    if not rebuilt:
        raise SystemExit(f"no translation unit is selected by {extra_flags}")
    objs = [os.path.join(OBJ, f) for f in sorted(os.listdir(OBJ)) if f.endswith(".o") and f not in replaced] + rebuilt
    libpath = os.path.join(out_dir, "libaphrodite_b200.so")
    if _newer(libpath, objs):
        _run([NVCC, "-shared", "-o", libpath, *objs, "-gencode", "arch=compute_100a,code=sm_100a"])
    shutil.copy2(os.path.join(LIB, "_C_b200.so"), os.path.join(out_dir, "_C_b200.so"))
    shutil.copy2(os.path.join(LIB, "_C_b200_overlay.so"), os.path.join(out_dir, "_C_b200_overlay.so"))
    return out_dir


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "variant":  # python build.py variant <name> <flag> ...
        print(build_variant(sys.argv[2], sys.argv[3:]))
    else:
        print(build(verbose=True))
