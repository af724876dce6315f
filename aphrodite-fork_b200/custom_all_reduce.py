"""One-shot all-reduce over NVLink peer memory for the decode step's row-parallel outputs.

Host side of csrc/all_reduce.cu, the counterpart of the reference's CustomAllreduce
(distributed/device_communicators/custom_all_reduce.py:47-304): the reference exchanges CUDA-IPC handles through
Python and registers buffers with a C++ object; here the symmetric buffer and its peer mappings come from
`torch.distributed._symmetric_memory` and the kernel is called through the C ABI on the current stream (capturable in a
CUDA graph). Eligibility mirrors the reference (`should_custom_ar`, :216-229): contiguous, a multiple of 16 bytes, at
most `max_size` bytes, world size 2/4/6/8; anything else goes to NCCL.
"""
from __future__ import annotations

import ctypes
import os
from contextlib import contextmanager
from typing import Optional, Union

import torch
import torch.distributed as dist

from ._loader import c_api

_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


class OneShotAllReduce:
    _SUPPORTED_WORLD_SIZES = (2, 4, 6, 8)  # custom_all_reduce.py:49

    def __init__(self, group: dist.ProcessGroup, device: torch.device, max_size: int = 8 * 1024 * 1024, max_blocks: int = 0):
        import torch.distributed._symmetric_memory as symm

        self.group = group
        self.world_size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world_size not in self._SUPPORTED_WORLD_SIZES:
            raise RuntimeError(f"one-shot all-reduce supports world sizes {self._SUPPORTED_WORLD_SIZES}")
        self.lib = c_api()
        self.lib.aph_all_reduce_signal_bytes.restype = ctypes.c_size_t
        self.signal_bytes = int(self.lib.aph_all_reduce_signal_bytes())
        self.max_size = max_size
        self.max_blocks = max_blocks
        self.device = torch.device(device)
        self.lib.aph_all_reduce_buffer_bytes.restype = ctypes.c_size_t
        total = int(self.lib.aph_all_reduce_buffer_bytes(self.world_size, ctypes.c_int64(max_size)))
        self.buffer = symm.empty(total, dtype=torch.uint8, device=self.device)
        self.buffer.zero_()  # the flags and the LL lines start at zero (a line's flag is the call number, from 1 on)
        self.lib.aph_all_reduce_ll_max_bytes.restype = ctypes.c_size_t
        self.ll_max_bytes = int(self.lib.aph_all_reduce_ll_max_bytes())
        self.ll_max_passes = int(self.lib.aph_all_reduce_ll_max_passes())
        self.handle = symm.rendezvous(self.buffer, group.group_name)
        self.peer_ptrs = int(self.handle.buffer_ptrs_dev)
        # NVLS: the multicast mapping of the same buffer, 0 where the device / driver / container has no multicast support
        try:
            self.multicast_ptr = int(self.handle.multicast_ptr or 0)
        except Exception:  # older torch: no such attribute
            self.multicast_ptr = 0
        self.counters = torch.zeros(int(self.lib.aph_all_reduce_max_blocks()), dtype=torch.int32, device=self.device)
        torch.cuda.synchronize(self.device)
        dist.barrier(group)  # every rank's flags are zero before anybody starts

    def two_shot(self, nbytes: int) -> bool:
        """One-shot pushes (world - 1) x the message per rank, two-shot 2 (world - 1) / world x for one more flag barrier
        (csrc/all_reduce.cu; the reference switches at 512 / 256 KiB for 4 / 8 ranks, custom_all_reduce.cuh:563-570).
        APH_AR_TWO_SHOT_BYTES (development) moves the threshold; 0 = never."""
        env = os.environ.get("APH_AR_TWO_SHOT_BYTES")
        if env is not None:
            return int(env) > 0 and nbytes >= int(env)
        if self.world_size <= 2:
            return False
        # measured on 8 B200s (profiles/r02_all_reduce_n8.jsonl, us one-shot / two-shot): 256 KiB 11.6 / 14.9,
        # 1 MiB 25.4 / 16.9, 4 MiB 60 / 27 - the curves cross near 512 KiB
        return nbytes >= 512 * 1024

    def mode(self, nbytes: int, rows: int = 0, hidden: int = 0) -> str:
        """Which kernel serves a message: "ll", "one_shot", "two_shot" or "multimem" (csrc/all_reduce.cu). rows / hidden
        describe the fused add + RMSNorm call (its LL form holds rows x ceil(hidden / 4096) passes of 512 vectors).
        APH_AR_MODE (development) forces one; multimem needs the multicast mapping."""
        ll_ok = nbytes <= self.ll_max_bytes and rows * ((hidden // 8 + 511) // 512) <= self.ll_max_passes
        env = os.environ.get("APH_AR_MODE")
        if env in ("one_shot", "two_shot"):
            return env
        if env == "multimem" and self.multicast_ptr:
            return env
        if env == "ll" and ll_ok:
            return env
        if os.environ.get("APH_AR_TWO_SHOT_BYTES") is not None:  # an explicit choice between the two barrier kernels
            return "two_shot" if self.two_shot(nbytes) else "one_shot"
        # Measured on 8 B200s (profiles/r02_all_reduce_multimem_n8.jsonl, us one-shot / two-shot / multimem): 256 KiB
        # 11.5 / 14.9 / 16.3, 1 MiB 24.8 / 16.8 / 26.1, 4 MiB 60 / 26.9 / 36.0 - at these sizes the switch reduction's
        # round trip costs more than the NVLink bytes it saves, so multimem stays opt-in (APH_AR_MODE=multimem).
        # LL against the best barrier kernel, 8 B200s, fused with the norm (profiles/r02_all_reduce_ll_n8.jsonl): 16 KiB 9.4 vs
        # 9.9 us, 256 KiB 10.2 vs 13.2, 1 MiB 15.6 vs 18.0, 2 MiB 27.0 vs 20.3 (the doubled bytes win). Two ranks: ahead to 256 KiB.
        if ll_ok and env != "no_ll" and nbytes <= (1 << 20 if self.world_size > 2 else 1 << 19):
            return "ll"
        return "two_shot" if self.two_shot(nbytes) else "one_shot"

    def should_use(self, x: torch.Tensor) -> bool:
        """One-shot pushes world x the message per rank, so on larger groups it only pays up to a size. The reference
        stops at 512 KiB (<= 4 ranks) and 256 KiB (<= 8 ranks) for its pull kernel (custom_all_reduce.cuh:563-570); the
        push kernel was measured against NCCL on 4 and 8 B200s at the decode step's 1 MiB messages and is still ahead
        (profiles/r01_bench_n4_one_shot_vs_nccl.txt), so the limits here are 2 MiB and 1 MiB; larger messages take NCCL."""
        n = x.numel() * x.element_size()
        limit = self.max_size  # with the two-shot kernel for the larger messages the peer-memory path holds up to 8 MiB
        limit = int(os.environ.get("APH_AR_LIMIT_BYTES", limit))  # development: sweep the threshold
        return (x.is_cuda and x.is_contiguous() and x.dtype in _DT and n % 16 == 0 and 0 < n <= min(limit, self.max_size)
                and x.data_ptr() % 16 == 0)

    def all_reduce(self, x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out = sum over ranks of x (out defaults to x: in place, as TPGroup.all_reduce)."""
        out = x if out is None else out
        mode = self.mode(x.numel() * x.element_size())
        common = (self.rank, self.world_size, ctypes.c_int64(x.numel()), _DT[x.dtype], ctypes.c_int64(self.max_size), self.max_blocks,
                  ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        ptrs = (ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(self.peer_ptrs))
        cnt = ctypes.c_void_p(self.counters.data_ptr())
        if mode == "multimem":
            st = self.lib.aph_multimem_all_reduce(*ptrs, ctypes.c_void_p(self.multicast_ptr), cnt, *common)
        elif mode == "ll":
            st = self.lib.aph_ll_all_reduce(*ptrs, cnt, *common)
        elif mode == "two_shot":
            st = self.lib.aph_two_shot_all_reduce(*ptrs, cnt, *common)
        else:
            st = self.lib.aph_one_shot_all_reduce(*ptrs, cnt, *common)
        if st != 0:
            raise RuntimeError(self.lib.aph_b200_last_error().decode())
        return out

    def should_fuse_norm(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor) -> bool:
        h = x.shape[-1]
        return (self.should_use(x) and x.dtype in (torch.float16, torch.bfloat16) and x.dim() == 2 and h % 8 == 0 and h <= 16384
                and residual.is_contiguous() and residual.shape == x.shape and residual.dtype == x.dtype
                and weight.is_contiguous() and weight.dtype == x.dtype and weight.numel() == h
                and residual.data_ptr() % 16 == 0 and weight.data_ptr() % 16 == 0)

    def all_reduce_add_rms_norm(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
        """x <- rms_norm(all_reduce(x) + residual) * weight, residual <- all_reduce(x) + residual: the all-reduce and the
        fused_add_rms_norm that follows it in every decoder layer, in one kernel."""
        mode = self.mode(x.numel() * x.element_size(), x.shape[0], x.shape[1])
        head = (ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(residual.data_ptr()), ctypes.c_void_p(weight.data_ptr()),
                ctypes.c_float(epsilon), ctypes.c_void_p(self.peer_ptrs))
        tail = (ctypes.c_void_p(self.counters.data_ptr()), self.rank, self.world_size, ctypes.c_int64(x.shape[0]),
                ctypes.c_int64(x.shape[1]), _DT[x.dtype], ctypes.c_int64(self.max_size),
                ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        if mode == "multimem":
            st = self.lib.aph_multimem_all_reduce_add_rms_norm(*head, ctypes.c_void_p(self.multicast_ptr), *tail)
        elif mode == "ll":
            st = self.lib.aph_ll_all_reduce_add_rms_norm(*head, *tail)
        elif mode == "two_shot":
            st = self.lib.aph_two_shot_all_reduce_add_rms_norm(*head, *tail)
        else:
            st = self.lib.aph_all_reduce_add_rms_norm(*head, *tail)
        if st != 0:
            raise RuntimeError(self.lib.aph_b200_last_error().decode())


def is_weak_contiguous(inp: torch.Tensor) -> bool:
    """Dense in memory, whatever the stride order (the reference's helper, custom_all_reduce.py:37-42)."""
    return inp.is_contiguous() or (inp.untyped_storage().nbytes() - inp.storage_offset() * inp.element_size()
                                   == inp.numel() * inp.element_size())


class CustomAllreduce:
    """The reference's communicator interface (distributed/device_communicators/custom_all_reduce.py:45-304) on top of
    OneShotAllReduce, for code written against `ca_comm.custom_all_reduce(x)` / `with ca_comm.capture():`. Same
    constructor arguments, same `disabled` convention (anything that rules the peer-memory path out leaves the object
    disabled instead of raising; callers then fall back to NCCL), same eligibility rule.

    What differs underneath: the buffers are torch symmetric memory, not hand-exchanged CUDA-IPC handles, and the
    kernels PUSH from the caller's tensor, so nothing has to be registered - `registered=` and `register_graph_buffers()`
    are accepted and do nothing, and a warm-up call inside `capture()` performs the real reduction (the reference
    returns an uninitialised tensor there)."""

    _SUPPORTED_WORLD_SIZES = [2, 4, 6, 8]

    def __init__(self, group: dist.ProcessGroup, device: Union[int, str, torch.device], max_size: int = 8192 * 1024) -> None:
        self._IS_CAPTURING = False
        self.disabled = True
        self.disabled_reason: Optional[str] = None
        self._impl: Optional[OneShotAllReduce] = None
        self.group = group
        self.max_size = max_size
        if isinstance(device, int):
            device = torch.device(f"cuda:{device}")
        elif isinstance(device, str):
            device = torch.device(device)
        self.device = device
        self.rank = dist.get_rank(group=group)
        self.world_size = dist.get_world_size(group=group)
        self.fully_connected = True  # B200 boards: every GPU reaches every peer through NVSwitch
        if self.world_size == 1:
            self.disabled_reason = "single GPU"
            return
        if self.world_size not in self._SUPPORTED_WORLD_SIZES:
            self.disabled_reason = f"unsupported world size {self.world_size} (supported: {self._SUPPORTED_WORLD_SIZES})"
            return
        if device.type != "cuda" or not torch.cuda.is_available():
            self.disabled_reason = "not a CUDA device"
            return
        try:
            self._impl = OneShotAllReduce(group, device, max_size)
        except Exception as e:  # no peer access / no symmetric memory for this group: stay disabled like the reference
            self.disabled_reason = f"{type(e).__name__}: {e}"
            return
        self.disabled = False

    @contextmanager
    def capture(self):
        """Wrap CUDA-graph capture (and its warm-up) of code that calls custom_all_reduce."""
        try:
            self._IS_CAPTURING = True
            yield
        finally:
            self._IS_CAPTURING = False
            if not self.disabled:
                self.register_graph_buffers()

    def register_graph_buffers(self) -> None:
        """Nothing to register: the kernels read the caller's tensor locally and write peers through symmetric memory."""

    def should_custom_ar(self, inp: torch.Tensor) -> bool:
        if self.disabled:
            return False
        inp_size = inp.numel() * inp.element_size()
        if inp_size % 16 != 0 or not is_weak_contiguous(inp):
            return False
        if inp.dtype not in _DT or inp.data_ptr() % 16 != 0 or inp_size == 0:
            return False
        if self.world_size == 2 or self.fully_connected:
            return inp_size < self.max_size
        return False

    @staticmethod
    def _flat(t: torch.Tensor) -> torch.Tensor:
        return t if t.is_contiguous() else t.as_strided((t.numel(),), (1,), t.storage_offset())

    def all_reduce(self, inp: torch.Tensor, *, out: Optional[torch.Tensor] = None, registered: bool = False) -> torch.Tensor:
        """Out-of-place all-reduce (sum). `registered` is accepted for compatibility (see the class comment)."""
        if out is None:
            out = torch.empty_like(inp)  # same strides as inp: element i of the storage is element i of the result
        self._impl.all_reduce(self._flat(inp), self._flat(out))
        return out

    def custom_all_reduce(self, input: torch.Tensor) -> Optional[torch.Tensor]:
        """None when this path does not serve the message (disabled, too large, misaligned): the caller uses NCCL."""
        if self.disabled or not self.should_custom_ar(input):
            return None
        return self.all_reduce(input, registered=self._IS_CAPTURING)

    def close(self) -> None:
        self._impl = None
        self.disabled = True

    def __del__(self):
        self.close()
