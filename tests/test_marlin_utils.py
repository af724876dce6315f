"""Host helpers of the Marlin format (aphrodite-fork_b200/marlin_utils.py) against fixtures evaluated by the
reference's own Python (tests/golden/make_golden.py marlin_utils -> marlin_utils_golden.npz). CPU only, exact equality."""
import os

import numpy as np
import pytest
import torch

from aphrodite_fork_b200 import marlin_utils as mu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "marlin_utils_golden.npz"))


def test_scale_permutations_and_type_ids():
    grouped, single = mu.get_scale_perms()
    assert grouped == G["scale_perm"].tolist() and single == G["scale_perm_single"].tolist()
    for name in ("uint4", "uint8", "uint4b8", "uint8b128"):
        assert getattr(mu.scalar_types, name).id == int(G["id_" + name][0])
    assert mu.scalar_types.uint4b8.has_bias() and not mu.scalar_types.uint4.has_bias()


@pytest.mark.parametrize("ci", range(int(G["num_cases"][0])))
def test_scales_zero_points_and_packers_match_the_reference(ci):
    pre = f"c{ci}_"
    k, n, gs, bits, groups = (int(v) for v in G[pre + "meta"])
    s, zp = torch.from_numpy(G[pre + "s"]), torch.from_numpy(G[pre + "zp"])
    assert np.array_equal(mu.marlin_permute_scales(s, k, n, gs).numpy(), G[pre + "s_marlin"])
    assert np.array_equal(mu.marlin_zero_points(zp, groups, n, bits).numpy(), G[pre + "zp_marlin"])
    packed = mu.pack_cols(zp, bits, groups, n)
    assert packed.dtype == torch.int32 and np.array_equal(packed.numpy(), G[pre + "zp_packed"])
    assert np.array_equal(mu.unpack_cols(packed, bits, groups, n).numpy(), G[pre + "zp_unpacked"])
    awq = torch.from_numpy(G[pre + "awq_zp_packed"])
    assert np.array_equal(mu.awq_to_marlin_zero_points(awq, groups, n, bits).numpy(), G[pre + "awq_zp_marlin"])
    # an AWQ checkpoint's zero points are the same values: both routes meet in the same Marlin layout
    assert np.array_equal(G[pre + "awq_zp_marlin"], G[pre + "zp_marlin"])


def test_g_idx_sort_moe_scales_and_empties():
    srt, order = mu.marlin_sort_g_idx(torch.from_numpy(G["g_idx"]))
    assert order.dtype == torch.int32 and np.array_equal(srt.numpy(), G["g_idx_sorted"])
    # argsort is not stable in general: the order must sort g_idx and be a permutation, as the reference's does
    assert np.array_equal(np.sort(order.numpy()), np.arange(len(G["g_idx"]))) and np.array_equal(G["g_idx"][order.numpy()], G["g_idx_sorted"])
    assert np.array_equal(mu.marlin_moe_permute_scales(torch.from_numpy(G["moe_s"]), 256, 128, 128).numpy(), G["moe_s_marlin"])
    for make in (mu.marlin_make_empty_g_idx, mu.marlin_make_empty_zp):
        t = make(torch.device("cpu"))
        assert t.numel() == 0 and t.dtype == torch.int32 and not t.requires_grad


def test_support_predicates_match_the_reference():
    names = ("uint4", "uint8", "uint4b8", "uint8b128")
    for ti, gs, has_zp, want in G["supported"].tolist():
        qt = getattr(mu.scalar_types, names[ti])
        assert mu.check_marlin_supported(qt, gs, bool(has_zp), 100) == bool(want), (names[ti], gs, has_zp)
        if not want:
            with pytest.raises(ValueError, match="Marlin does not support"):
                mu.verify_marlin_supported(qt, gs, bool(has_zp))
    assert mu.query_marlin_supported_quant_types(False, 75) == []
    for o, i, full, gs, want in G["shapes"].tolist():
        ok, msg = mu.check_marlin_supports_shape(o, i, full, gs)
        assert ok == bool(want) and (msg is None) == bool(want)
    assert mu.marlin_is_k_full(True, True) is False and mu.marlin_is_k_full(True, False) and mu.marlin_is_k_full(False, True)
    assert mu.marlin_repeat_scales_on_all_ranks(False, -1, True) and not mu.marlin_repeat_scales_on_all_ranks(False, 128, True)
    ws = mu.marlin_make_workspace(4096, torch.device("cpu"))
    assert ws.numel() == 4096 // 64 * 16 and ws.dtype == torch.int32 and int(ws.abs().sum()) == 0


def test_pack_cols_rejects_bad_shapes():
    with pytest.raises(AssertionError):
        mu.pack_cols(torch.zeros(4, 12, dtype=torch.int32), 4, 4, 12)  # 12 columns are not a multiple of 8 nibbles
    with pytest.raises(AssertionError):
        mu.unpack_cols(torch.zeros(4, 3, dtype=torch.int32), 4, 4, 16)
    with pytest.raises(Exception, match="num_bits must be 4 or 8"):
        mu.marlin_zero_points(torch.zeros(2, 64, dtype=torch.int32), 2, 64, 2)
