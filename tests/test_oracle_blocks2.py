"""oracle/blocks2.py (custom_mapping, randn_orthobasis_torch, gt_in_norm_range, is_gt_included)
against the reference: live where /root/reference exists, and against
tests/golden/blocks2_0.npz (outputs of the reference's own functions, make_golden.py)."""
import contextlib
import io
import os
import types

import numpy as np
import pytest
import torch

from oracle import blocks2 as ob, refshim

Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "blocks2_0.npz"))


def test_custom_mapping_matches_golden():
    for bi, beta in enumerate(Z["cm_betas"]):
        got = ob.custom_mapping([float(v) for v in Z["cm_x"]], float(beta))
        assert np.array_equal(np.asarray(got), Z["cm_out"][bi])          # same double arithmetic: exact


def test_randn_orthobasis_replay_matches_golden():
    got = ob.randn_orthobasis_from_draws(torch.from_numpy(Z["ro_z"]))
    assert np.array_equal(got.numpy(), Z["ro_out"])
    R = got.reshape(-1, 3, 3)
    eye = torch.eye(3).expand_as(R)
    assert torch.allclose(R @ R.transpose(1, 2), eye, atol=1e-5)


def test_gt_in_norm_range_matches_golden():
    lo, hi, gt = (torch.from_numpy(Z[k]) for k in ("gn_lo", "gn_hi", "gn_gt"))
    got = torch.stack([ob.gt_in_norm_range((lo[i], hi[i]), gt[i]) for i in range(len(gt))])
    assert np.array_equal(got.numpy(), Z["gn_out"])


def test_is_gt_included_matches_golden():
    inc = Z["ig_included"]
    assert 4 <= inc.sum() <= len(inc) - 4
    for i in range(len(inc)):
        r, p = Z["ig_ranges"][i], Z["ig_priors"][i]
        bits, _ = ob.is_gt_included_bits(Z["ig_center"][i], Z["ig_dims"][i], r[0], r[1], r[2], p[0], p[1], p[2])
        assert (bits == 0) == bool(inc[i]), i


@pytest.mark.ref
def test_blocks2_oracle_matches_reference_live():
    U = refshim.load().utils
    g = torch.Generator().manual_seed(77)
    xs = [float(v) for v in torch.rand(50, generator=g)] + [0.0, -1.0]
    for beta in (1.7, 2.5):
        assert ob.custom_mapping(xs, beta) == U.custom_mapping(xs, beta)
    torch.manual_seed(9)
    z = torch.randn(4, 5, 3, 3)
    torch.manual_seed(9)
    assert torch.equal(ob.randn_orthobasis_from_draws(z), U.randn_orthobasis_torch(5, 4))
    for _ in range(200):
        c, d = torch.randn(3, generator=g) * 2, 0.2 + 2 * torch.rand(3, generator=g)
        r = torch.sort(torch.randn(3, 2, generator=g) * 3, dim=-1)[0]
        p = torch.stack((0.2 + 2 * torch.rand(3, generator=g), 0.1 + 0.5 * torch.rand(3, generator=g)), -1)
        with contextlib.redirect_stdout(io.StringIO()) as buf:
            ref = U.is_gt_included(types.SimpleNamespace(center=c, dimensions=d), r[0], r[1], r[2], p[0], p[1], p[2])
        bits, by = ob.is_gt_included_bits(c.numpy(), d.numpy(), r[0].numpy(), r[1].numpy(), r[2].numpy(),
                                          p[0].numpy(), p[1].numpy(), p[2].numpy())
        assert (bits == 0) == bool(ref)
        if bits:
            want = [f"{ob.REASONS[i]} by {by[i]:.1f}" for i in range(3) if bits >> i & 1] + \
                   [ob.REASONS[i] for i in range(3, 9) if bits >> i & 1]
            assert str(want) in buf.getvalue()
