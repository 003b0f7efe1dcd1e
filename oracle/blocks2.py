"""CPU restatements of the small helpers of ProposalNetwork/utils/utils.py that round 2 adds to the
device path. TEST INFRASTRUCTURE ONLY (imported by tests/ and by
tests/golden/make_golden.py); pinned to the reference's own functions by
tests/test_oracle_blocks2.py (live, authoring container) and to tests/golden/blocks2_0.npz.

  custom_mapping          utils.py:207-223     randn_orthobasis_torch  utils.py:61-68
  gt_in_norm_range        utils.py:148-152     is_gt_included          utils.py:257-298
"""
import numpy as np
import torch

REASONS = ("x", "y", "z", "w-", "w+", "h-", "h+", "l-", "l+")


def custom_mapping(x, beta=1.7):
    """utils.py:207-223, in Python floats (double) like the reference."""
    out = []
    for v in x:
        v = float(v)
        out.append(0.0 if v <= 0 else (1 / (1 + (v / (1 - v)) ** (-beta))))
    return out


def randn_orthobasis_from_draws(z: torch.Tensor) -> torch.Tensor:
    """utils.py:61-68 from the point where ``z = torch.randn(num_instances, num_samples, 3, 3)``
    exists (the draws are an INPUT here: replay)."""
    z = z.clone()
    z = z / torch.norm(z, p=2, dim=-1, keepdim=True)
    z[:, :, 0] = torch.cross(z[:, :, 1], z[:, :, 2], dim=-1)
    z[:, :, 0] = z[:, :, 0] / torch.norm(z[:, :, 0], dim=-1, keepdim=True)
    z[:, :, 1] = torch.cross(z[:, :, 2], z[:, :, 0], dim=-1)
    z[:, :, 1] = z[:, :, 1] / torch.norm(z[:, :, 1], dim=-1, keepdim=True)
    return z


def gt_in_norm_range(rng, gt):
    """utils.py:148-152."""
    tmp = gt - rng[0]
    return tmp / abs(rng[1] - rng[0])


def is_gt_included_bits(center, dims, x_range, y_range, z_range, w_prior, h_prior, l_prior):
    """utils.py:257-298 as (bits, (x_by, y_by, z_by)); bits == 0 <=> the reference returns True.
    Bit i <-> REASONS[i], the order the reference appends to `because_of`. float32 arithmetic
    (the reference compares 0-d float32 tensors)."""
    f = np.float32
    stds_away = f(1.5)
    bits, by = 0, [f(0), f(0), f(0)]
    for a, r in enumerate((x_range, y_range, z_range)):
        v, lo, hi = f(center[a]), f(r[0]), f(r[1])
        if not (lo < v < hi):
            bits |= 1 << a
            by[a] = abs(lo - v) if v < lo else abs(v - hi)
    for a, p in enumerate((w_prior, h_prior, l_prior)):
        d, mean, sd = f(dims[a]), f(p[0]), f(p[1])
        if d < mean - stds_away * sd:
            bits |= 8 << (2 * a)
        if d > mean + stds_away * sd:
            bits |= 16 << (2 * a)
    return bits, [float(b) for b in by]
