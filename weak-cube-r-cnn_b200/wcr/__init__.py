"""wcr -- host-side mirror of the Weak Cube R-CNN hot path on B200.

Python/PyTorch plumbing over libwcr_b200.so (include/wcr.h). Module layout
mirrors the reference so call sites can switch imports:

  ProposalNetwork.utils.plane        -> wcr.plane        (Plane_torch)
  ProposalNetwork.utils.spaces       -> wcr.spaces       (Cubes)
  ProposalNetwork.utils.conversions  -> wcr.conversions  (cubes_to_box)
  ProposalNetwork.utils.utils        -> wcr.utils        (iou_2d, mask_iou, ...)
  ROIHeads3DScore loss helpers       -> wcr.roi_heads
  new (no reference counterpart)     -> wcr.proposals    (propose, score_topk)
"""
from . import _abi  # noqa: F401
from ._abi import WcrError, LIB_PATH  # noqa: F401
