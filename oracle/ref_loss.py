"""TEST INFRASTRUCTURE ONLY (oracle). Authoring container only (needs /root/reference).

The weak-loss block of ROIHeads3DScore._forward_cube (roi_heads.py:843-1144) driven
line by line through the REFERENCE'S OWN functions (imported via oracle/refshim.py):
Cubes, cubes_to_box, R_from_allocentric, ROIHeads3DScore.pose_loss /
pseudo_gt_z_point_loss / dim_loss / safely_reduce_losses, torchvision's
generalized_box_iou_loss. It pins oracle/loss.py (tests/test_oracle_loss.py) and
generates tests/golden/loss_*.npz (tests/golden/make_golden.py). Only the glue that
the reference keeps inside the 450-line method body (gathers, Python loops) is
restated here, in the reference's order.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import refshim, standins


def reference_weak_loss(cfg, deltas, z_raw, dims_raw, pose, uncert, *, src_boxes, K4,
                        virtual_to_real, prior_mean, prior_std, gt_boxes, clamp,
                        num_boxes_per_image, depth_maps, normals, ground_conf, gt_masks=None,
                        at_which_mask_idx=None):
    from torchvision.ops import generalized_box_iou_loss
    ns = refshim.load()
    head = refshim.roi_head_instance()
    Cubes, cubes_to_box = ns.spaces.Cubes, ns.conversions.cubes_to_box
    n = deltas.shape[0]
    Ks = torch.zeros(n, 3, 3)
    Ks[:, 0, 0], Ks[:, 1, 1], Ks[:, 0, 2], Ks[:, 1, 2], Ks[:, 2, 2] = K4[:, 0], K4[:, 1], K4[:, 2], K4[:, 3], 1
    im_sizes = [[int(c[0]), int(c[1])] for c in clamp]
    # roi_heads.py:786-846
    src_widths = src_boxes[:, 2] - src_boxes[:, 0]
    src_heights = src_boxes[:, 3] - src_boxes[:, 1]
    src_ctr_x = src_boxes[:, 0] + 0.5 * src_widths
    src_ctr_y = src_boxes[:, 1] + 0.5 * src_heights
    cube_x = src_ctr_x + src_widths * deltas[:, 0]
    cube_y = src_ctr_y + src_heights * deltas[:, 1]
    cube_xy = torch.cat((cube_x.unsqueeze(1), cube_y.unsqueeze(1)), dim=1)
    # :848-866
    if cfg.dims_priors_func == "sigmoid":
        dmin = (prior_mean - 3 * prior_std).clip(0.0)
        dmax = prior_mean + 3 * prior_std
        cube_dims = ns.math_util.scaled_sigmoid(dims_raw, min=dmin, max=dmax)
    else:
        cube_dims = torch.exp(dims_raw.clip(max=5)) * prior_mean
    # :868-870
    cube_pose = ns.math_util.R_from_allocentric(Ks, pose, u=cube_x.detach(), v=cube_y.detach())
    # :872-905
    cube_z = z_raw.squeeze()
    if cfg.z_type == "sigmoid":
        cube_z = torch.sigmoid(cube_z) * 100
    elif cfg.z_type == "log":
        cube_z = torch.exp(cube_z)
    cube_z = cube_z * virtual_to_real
    # :927-931
    cube_x3d = cube_z * (cube_x - Ks[:, 0, 2]) / Ks[:, 0, 0]
    cube_y3d = cube_z * (cube_y - Ks[:, 1, 2]) / Ks[:, 1, 1]
    cubes_tensor = torch.cat((cube_x3d.unsqueeze(1), cube_y3d.unsqueeze(1), cube_z.unsqueeze(1),
                              cube_dims, cube_pose.reshape(n, 9)), axis=1).unsqueeze(1)
    cubes = Cubes(cubes_tensor)
    # :956-960
    proj_boxes = []
    for i in range(cubes.num_instances):
        proj_boxes.append(cubes_to_box(cubes[i], Ks[i], im_sizes[i])[0].tensor[0])
    proj = torch.stack(proj_boxes)
    lf = cfg.loss_functions
    terms = {}
    if "iou" in lf:   # :977-978
        terms["iou"] = generalized_box_iou_loss(gt_boxes, proj, reduction="none").view(n, -1).mean(dim=1)
    if "pose_alignment" in lf:   # :981-984
        lp = head.pose_loss(cube_pose, list(num_boxes_per_image))
        if lp is not None:
            terms["pose"] = lp.repeat(n)
    if "pose_ground" in lf:   # :987-995 (normals are an input here; A1.x pins them)
        nb = torch.tensor(list(num_boxes_per_image))
        nv = normals.repeat_interleave(nb, 0)
        conf = ground_conf.repeat_interleave(nb, 0)
        pred_normal = cube_pose[:, 1, :]
        terms["normal_vec"] = (1 - F.cosine_similarity(nv, pred_normal, dim=1).abs()) * conf
    if "pose_ground2" in lf:   # :997-1007
        nb = torch.tensor(list(num_boxes_per_image))
        nv = normals.repeat_interleave(nb, 0)
        conf = ground_conf.repeat_interleave(nb, 0)
        gt_R = head.normal_to_rotation(nv)
        terms["normal_vec"] = (1 - ns.roi_heads.so3_relative_angle(cube_pose, gt_R, cos_angle=True)) * conf
    if "z_pseudo_gt_patch" in lf:   # :1010-1011
        dm = standins.ImageList.from_tensors(list(depth_maps))
        terms["pseudo_gt_z"] = head.pseudo_gt_z_box_loss(dm, proj.split(list(num_boxes_per_image)), cube_z)
    elif "z_pseudo_gt_center" in lf:   # :1012-1013
        dm = standins.ImageList.from_tensors(list(depth_maps))
        terms["pseudo_gt_z"] = head.pseudo_gt_z_point_loss(dm, cube_xy, cube_z, list(num_boxes_per_image))
    if "segmentation" in lf:   # :948-954, 1016-1017
        bube = torch.zeros((n, 8, 2))
        for i in range(n):
            bc = cubes[i].get_bube_corners(Ks[i], im_sizes[i])
            x = torch.clamp(bc[..., 0], 0, int(im_sizes[i][0] - 1))
            y = torch.clamp(bc[..., 1], 0, int(im_sizes[i][1] - 1))
            bube[i] = torch.stack((x, y), dim=-1)
        terms["seg"] = head.segment_loss(gt_masks, bube, at_which_mask_idx)
    if "z" in lf:   # :1019-1021
        terms["z"] = head.z_loss(standins.Boxes(gt_boxes), cubes, Ks, im_sizes, standins.Boxes(proj))
    if "dims" in lf:   # :1024-1025
        lw, lh, ll = head.dim_loss((prior_mean, prior_std), cubes.dimensions.squeeze(1))
        if lw is not None:
            terms["dims_w"], terms["dims_h"], terms["dims_l"] = lw, lh, ll
    if "depth" in lf:   # :1027-1029
        dm = standins.ImageList.from_tensors(list(depth_maps))
        terms["depth"] = head.depth_range_loss(gt_masks, at_which_mask_idx, dm, cubes, standins.Boxes(gt_boxes),
                                               list(num_boxes_per_image))
    per_box = {"proj": proj.detach(), "cube_z": cube_z.detach(), "dims": cube_dims.detach(),
               "R": cube_pose.detach()}
    per_box.update({"raw_" + k: v.detach().clone() for k, v in terms.items()})
    losses = {}
    if cfg.use_confidence > 0:   # :1075-1105
        uncert_sf = 1.41421356 * torch.exp(-uncert)
        terms = {k: v * uncert_sf for k, v in terms.items()}
        losses["Cube/uncert"] = cfg.use_confidence * head.safely_reduce_losses(uncert.clone())
    w = {"iou": cfg.loss_w_iou, "pose": cfg.loss_w_pose, "normal_vec": cfg.loss_w_normal_vec,
         "pseudo_gt_z": cfg.loss_w_z, "dims_w": cfg.loss_w_dims, "dims_h": cfg.loss_w_dims,
         "dims_l": cfg.loss_w_dims, "seg": cfg.loss_w_seg, "z": cfg.loss_w_z, "depth": cfg.loss_w_depth}
    for k, v in terms.items():   # :1107-1144
        losses["Cube/loss_" + k] = head.safely_reduce_losses(v) * w[k] * cfg.loss_w_3d
    return losses, per_box
