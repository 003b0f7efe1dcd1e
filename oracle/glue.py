"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

Torch restatement of the caller glue in front of the loss block, line by line:
  roi_heads.py:748-781  Ks_scaled_per_box, focal / ratio / scale per box, virtual_to_real
  roi_heads.py:810-814  im_sizes per box
  roi_heads.py:823-840  class gather of the head outputs
  cube_head.py:174-176  rotation_6d_to_matrix over every class (pytorch3d; restated in standins.py)
  math_util.py:732-743  compute_virtual_scale_from_focal_spaces

Parity status: PINNED for the virtual-scale helper (tests call the reference's own
math_util.compute_virtual_scale_from_focal_spaces when /root/reference is present); the
remaining lines are plain torch indexing / arithmetic copied in meaning from the method body,
which cannot be called in isolation (it needs the detector's pooler and head).
"""
from __future__ import annotations

import torch

from .standins import rotation_6d_to_matrix  # noqa: F401  (re-exported for the tests)


def compute_virtual_scale_from_focal_spaces(f, H, f0, H0):
    return (H0 * f) / (f0 * H)


def box_constants(Ks, im_scales_ratio, im_current_dims, num_boxes_per_image, virtual_focal=512.0,
                  virtual_depth=True):
    Ks_scaled_per_box = torch.cat([
        (Ks[i] / im_scales_ratio[i]).unsqueeze(0).repeat([num, 1, 1])
        for (i, num) in enumerate(num_boxes_per_image)])
    Ks_scaled_per_box[:, -1, -1] = 1
    focal = torch.cat([(Ks[i][1, 1]).unsqueeze(0).repeat([num]) for (i, num) in enumerate(num_boxes_per_image)])
    ratios = torch.cat([torch.FloatTensor([im_scales_ratio[i]]).repeat(num)
                        for (i, num) in enumerate(num_boxes_per_image)])
    scales = torch.cat([torch.FloatTensor([im_current_dims[i][0]]).repeat(num)
                        for (i, num) in enumerate(num_boxes_per_image)])
    scales_original = scales * ratios
    if virtual_depth:
        v2r = compute_virtual_scale_from_focal_spaces(focal, scales_original, virtual_focal, scales)
    else:
        v2r = torch.ones_like(focal)
    im_sizes = []
    for i, j in enumerate(num_boxes_per_image):
        for _ in range(j):
            im_sizes.append(list(im_current_dims[i]))
    K4 = torch.stack((Ks_scaled_per_box[:, 0, 0], Ks_scaled_per_box[:, 1, 1], Ks_scaled_per_box[:, 0, 2],
                      Ks_scaled_per_box[:, 1, 2]), 1)
    return K4, v2r, torch.tensor(im_sizes, dtype=torch.float32)


def class_gather(x, box_classes):
    return x[torch.arange(x.shape[0]), box_classes]
