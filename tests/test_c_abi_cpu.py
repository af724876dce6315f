"""CPU: the C-ABI library loads without a GPU and exports every entry point include/aphrodite_b200.h declares; the
host-side argument checks of the entry points answer with a status + message (no compute is launched)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "aphrodite_b200.h")
LIB = os.path.join(ROOT, "aphrodite-fork_b200", "lib", "libaphrodite_b200.so")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(aph_[a-z0-9_]+)\s*\(", text)))


def _lib():
    if not os.path.exists(LIB):
        pytest.skip("libaphrodite_b200.so not built (python __graft_entry__.py)")
    return ctypes.CDLL(LIB)


def test_every_declared_symbol_is_exported():
    names = _declared()
    assert len(names) >= 10, names
    lib = _lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_abi_version_and_kv_dtype_strings():
    lib = _lib()
    assert lib.aph_b200_abi_version() >= 1
    f = lib.aph_kv_dtype_from_string
    assert f(b"auto") == 0 and f(b"fp8") == f(b"fp8_e4m3") == 1 and f(b"fp8_e5m2") == 2 and f(b"int8") < 0


def test_argument_checks_return_status_and_message():
    lib = _lib()
    lib.aph_b200_last_error.restype = ctypes.c_char_p
    # bad sizes are rejected on the host before any CUDA call
    st = lib.aph_swap_blocks(None, None, None, ctypes.c_int64(-1), ctypes.c_int64(0), 0, None)
    assert st != 0 and b"swap_blocks" in lib.aph_b200_last_error()
    st = lib.aph_copy_blocks(None, None, None, 1, ctypes.c_int64(70000), ctypes.c_int64(16), None)
    assert st != 0 and b"65535" in lib.aph_b200_last_error()
    # zero work is a successful no-op
    assert lib.aph_copy_blocks(None, None, None, 0, ctypes.c_int64(0), ctypes.c_int64(16), None) == 0


def test_argument_checks_of_the_next_row_entry_points():
    """Host-side validation of the repack / glue / MLA / all-reduce entry points: bad arguments come back as a status
    and a message before any CUDA call, empty work is a successful no-op."""
    lib = _lib()
    lib.aph_b200_last_error.restype = ctypes.c_char_p
    lib.aph_all_reduce_signal_bytes.restype = ctypes.c_size_t
    lib.aph_all_reduce_buffer_bytes.restype = ctypes.c_size_t
    i64, f32 = ctypes.c_int64, ctypes.c_float
    one = ctypes.c_void_p(16)  # a non-null, 16-byte aligned placeholder: only looked at, never dereferenced on these paths

    assert lib.aph_gptq_marlin_repack(one, one, None, i64(128), i64(64), 3, None) != 0
    assert b"num_bits must be 4 or 8" in lib.aph_b200_last_error()
    assert lib.aph_awq_marlin_repack(one, one, i64(120), i64(64), 4, None) != 0
    assert b"tile_k_size" in lib.aph_b200_last_error()
    assert lib.aph_gptq_marlin_repack(one, one, None, i64(128), i64(96), 4, None) != 0
    assert b"tile_n_size" in lib.aph_b200_last_error()
    assert lib.aph_permute_cols(one, one, one, i64(4), i64(100), i64(100), None) != 0

    assert lib.aph_rms_norm(None, None, None, f32(1e-6), i64(0), i64(4096), i64(4096), 2, None) == 0
    assert lib.aph_fused_add_rms_norm(None, None, None, f32(1e-6), i64(4), i64(4096), 2, None) != 0
    assert lib.aph_rms_norm(one, one, one, f32(1e-6), i64(4), i64(4096), i64(4096), 7, None) != 0
    assert b"unknown dtype" in lib.aph_b200_last_error()
    assert lib.aph_silu_and_mul(one, one, i64(70000), i64(128), 2, None) != 0
    assert b"65535" in lib.aph_b200_last_error()
    assert lib.aph_act_and_mul(99, one, one, i64(4), i64(128), f32(0), 2, None) != 0
    assert lib.aph_activation(0, one, one, i64(128), 2, None) != 0  # silu_and_mul is not an element-wise activation
    assert lib.aph_rotary_embedding(one, one, one, one, None, i64(4), 130, i64(1024), i64(256), 8, 2, 128, 1, 2, None) != 0
    assert b"rot_dim" in lib.aph_b200_last_error()
    assert lib.aph_rotary_embedding(one, one, one, one, None, i64(4), 128, i64(1024), i64(256), 8, 3, 128, 1, 2, None) != 0
    assert lib.aph_advance_step_flashattn(4, 5, 16, one, one, one, one, one, one, i64(8), None) != 0
    assert lib.aph_advance_step_flashinfer(4, 5, 16, one, one, one, one, one, one, i64(8), one, i64(64), one, one, one, None) != 0
    assert b"num_queries <= num_seqs" in lib.aph_b200_last_error()
    assert lib.aph_advance_step_flashinfer(4, 4, 16, one, one, one, one, one, one, i64(8), one, i64(64), None, one, one, None) != 0
    assert b"paged_kv_indptr" in lib.aph_b200_last_error()

    assert lib.aph_concat_and_cache_mla(one, one, one, one, i64(4), 512, 64, 16, i64(9216), i64(576), i64(512), i64(64), 2, 1, None,
                                        None) != 0
    assert b"scale" in lib.aph_b200_last_error()
    assert lib.aph_gather_cache(one, one, one, one, None, 0, 16, i64(1152), i64(8), i64(18432), i64(1152), i64(1152), None) == 0
    assert lib.aph_copy_blocks_mla(None, None, 0, i64(0), i64(16), None) == 0

    assert lib.aph_all_reduce_signal_bytes() == 4096
    # flags | 2 halves x world slots | LL lines: world x 512 passes x ceil(512 / world) receive lines + 512 x 512 result lines of 32 bytes
    assert lib.aph_all_reduce_buffer_bytes(4, i64(1 << 20)) == 4096 + 2 * 4 * (1 << 20) + 4 * 512 * 128 * 32 + 512 * 512 * 32
    lib.aph_all_reduce_ll_max_bytes.restype = ctypes.c_size_t
    assert lib.aph_all_reduce_ll_max_bytes() == 2 << 20 and lib.aph_all_reduce_ll_max_passes() == 512
    assert lib.aph_ll_all_reduce(one, one, one, one, 0, 2, i64(4 << 20), 2, i64(16 << 20), 0, None) != 0
    assert b"LL lines" in lib.aph_b200_last_error()
    assert lib.aph_multimem_all_reduce(one, one, one, None, one, 0, 2, i64(1024), 2, i64(1 << 20), 0, None) != 0
    assert b"multicast" in lib.aph_b200_last_error()
    assert 1 <= lib.aph_all_reduce_max_blocks() <= 148
    assert lib.aph_one_shot_all_reduce(one, one, one, one, 0, 3, i64(1024), 2, i64(1 << 20), 0, None) != 0
    assert b"world size 3" in lib.aph_b200_last_error()
    assert lib.aph_one_shot_all_reduce(one, one, one, one, 2, 2, i64(1024), 2, i64(1 << 20), 0, None) != 0
    assert lib.aph_one_shot_all_reduce(one, one, one, one, 0, 2, i64(1 << 20), 2, i64(1 << 20), 0, None) != 0
    assert b"does not fit" in lib.aph_b200_last_error()
    assert lib.aph_one_shot_all_reduce(one, one, one, one, 0, 2, i64(0), 2, i64(1 << 20), 0, None) == 0
    assert lib.aph_all_reduce_add_rms_norm(one, one, one, f32(1e-5), one, one, 0, 2, i64(4), i64(4096), 0, i64(1 << 20), None) != 0
    assert b"float16 / bfloat16" in lib.aph_b200_last_error()
    assert lib.aph_all_reduce_add_rms_norm(one, one, one, f32(1e-5), one, one, 0, 2, i64(4), i64(32768), 2, i64(1 << 20), None) != 0


def test_noalloc_attention_entry_point_rejects_oversized_problems_on_the_host():
    """aph_paged_attention_v1_noalloc (static device scratch, no allocation): the size check is host-side."""
    from test_zz_no_alloc import reject_check
    _lib()
    reject_check()
