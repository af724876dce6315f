"""GPU: aph_paged_attention_v1_noalloc (split-K scratch in a static device array inside the library, no allocation in
the call) against aph_paged_attention_v1 with a caller workspace, both called through the C ABI with ctypes, two calls
in a row. The entry point has no counterpart in the reference. This file
is named so that pytest collects it last."""
import ctypes
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests", "golden")):  # also runs as a script (reject-child), without conftest.py
    if _p not in sys.path:
        sys.path.insert(0, _p)
from helpers import RTOL, make_attention_case  # noqa: E402

pytestmark = pytest.mark.gpu
MAX_SCRATCH = 16 << 20  # APH_PA_NOALLOC_MAX_SCRATCH_BYTES, include/aphrodite_b200.h


def _call(lib, name, out, q, kc, vc, bt, sl, KVH, max_len, ks, vs, kv, workspace=None):
    S, H, D = q.shape
    bs = vc.shape[3]
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    i64 = ctypes.c_int64
    args = [p(out), p(q), p(kc), p(vc), S, H, KVH, D, ctypes.c_float(float(D) ** -0.5), p(bt), p(sl), bs, max_len,
            bt.shape[1], None, i64(q.stride(0)), i64(out.stride(0)), i64(kc.stride(0)), i64(kc.stride(1)), 2,
            lib.aph_kv_dtype_from_string(kv.encode()), p(ks), p(vs), 0, 0, 0, 64, 0]
    if workspace is not None:
        args += [p(workspace), ctypes.c_size_t(workspace.numel())]
    args.append(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return getattr(lib, name)(*args)


def test_noalloc_entry_point_matches_the_workspace_entry_point_twice():
    import aphrodite_fork_b200 as A
    lib = A.c_api()
    dev = "cuda"
    ks = torch.tensor(1.0, device=dev)
    vs = torch.tensor(1.0, device=dev)
    # two different problems back to back: the shared scratch keeps the first call's partials. Long contexts at a small
    # batch: splits finer than the 8 partitions a thread-block cluster merges on chip, so the scratch is really used
    cases = [(2, 32, 8, 128, 16, [20000, 9000], "fp8", 11),
             (3, 16, 2, 128, 16, [9000, 4000, 257], "auto", 12)]
    for S, H, KVH, D, bs, lens, kv, seed in cases:
        q, kc, vc, bt, sl = [t.to(dev) for t in make_attention_case(S, H, KVH, D, bs, lens, torch.bfloat16, kv, seed)]
        need = lib.aph_paged_attention_workspace_bytes(S, H, KVH, D, max(lens))
        assert 0 < need <= MAX_SCRATCH, need  # the case must exercise the split-K scratch
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        want = torch.full((S, H, D), float("nan"), dtype=torch.bfloat16, device=dev)
        st = _call(lib, "aph_paged_attention_v1", want, q, kc, vc, bt, sl, KVH, max(lens), ks, vs, kv, workspace=ws)
        assert st == 0, lib.aph_b200_last_error()
        for rep in range(2):
            got = torch.full((S, H, D), float("nan"), dtype=torch.bfloat16, device=dev)
            before = torch.cuda.memory_allocated()
            st = _call(lib, "aph_paged_attention_v1_noalloc", got, q, kc, vc, bt, sl, KVH, max(lens), ks, vs, kv)
            assert st == 0, lib.aph_b200_last_error()
            assert torch.cuda.memory_allocated() == before
            torch.cuda.synchronize()
            atol = 1e-3 if kv == "auto" else 1e-2
            err = (got.double() - want.double()).abs()
            assert torch.isfinite(got.float()).all()
            assert bool((err <= atol + RTOL[torch.bfloat16] * want.double().abs()).all()), (kv, rep, float(err.max()))


def oversized_case(lib):
    """A problem whose split-K partials exceed the built-in scratch (found through the library's own size query)."""
    for S, H, KVH, D, max_len in [(355, 40, 5, 256, 8192), (222, 64, 8, 256, 32768), (1775, 8, 1, 256, 4096),
                                  (300, 48, 6, 256, 16384), (443, 32, 4, 256, 2048)]:
        if lib.aph_paged_attention_workspace_bytes(S, H, KVH, D, max_len) > MAX_SCRATCH:
            return S, H, KVH, D, max_len
    raise AssertionError("no candidate problem exceeds the built-in scratch; update the list")


def reject_child():
    """Runs with APH_PA_CLUSTER=0: with cluster merging no real problem needs more than a few MB of partials, so the
    size check is exercised with the round-1 behaviour (every partition through the scratch)."""
    lib = ctypes.CDLL(os.path.join(ROOT, "aphrodite-fork_b200", "lib", "libaphrodite_b200.so"))
    lib.aph_b200_last_error.restype = ctypes.c_char_p
    lib.aph_paged_attention_workspace_bytes.restype = ctypes.c_size_t
    S, H, KVH, D, max_len = oversized_case(lib)
    one = ctypes.c_void_p(256)  # never dereferenced: the size check comes first
    i64 = ctypes.c_int64
    st = lib.aph_paged_attention_v1_noalloc(one, one, one, one, S, H, KVH, D, ctypes.c_float(0.088), one, one, 16, max_len,
                                            2048, None, i64(H * D), i64(H * D), i64(KVH * D * 16), i64(D * 16), 2, 1, one,
                                            one, 0, 0, 0, 64, 0, None)
    assert st == 3 and b"built-in scratch" in lib.aph_b200_last_error()  # APH_ERR_WORKSPACE
    print("rejected")


def reject_check():
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "reject-child"], env=dict(os.environ, APH_PA_CLUSTER="0"),
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "rejected" in r.stdout, r.stderr[-2000:]


def test_noalloc_entry_point_rejects_problems_beyond_its_scratch():
    reject_check()


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "reject-child":
    reject_child()
