"""The oracle's composition (oracle/glue.py + oracle/loss.py: class gather -> decode -> weak-loss
block) against the REFERENCE's own ROIHeads3DScore._forward_cube (roi_heads.py:698-1144):
against the committed goldens everywhere, and against a live run where /root/reference exists."""
import ast
import glob
import os
import random

import numpy as np
import pytest
import torch

from oracle import glue as og, ground as ogr, loss as ol, ref_forward_cube as rf, refshim

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "forward_cube_*.npz")))


def _oracle_losses(case, leaves):
    nb, n, cls = case["nb"], sum(case["nb"]), case["cls"]
    K4, v2r, clamp = og.box_constants(case["Ks"], case["ratio"], case["dims_img"], nb)
    p = case["priors"][cls]
    # normal_vector_from_maps zips the per-BOX Ks against images (roi_heads.py:472, :990)
    im_of_box = [i for i, c in enumerate(nb) for _ in range(c)]
    rng = random.Random(case["rng_seed"])
    normals = []
    for i in range(len(nb)):
        j = im_of_box[i]
        K = (case["Ks"][j] / case["ratio"][j]).numpy().copy()
        K[-1, -1] = 1
        g = case["ground_maps"][i]
        gm = None if tuple(g.shape) == (1, 1) else g.numpy()
        pts, _ = ogr.unproject_ground_frame(case["depth_maps"][i].numpy(), gm, K, 5)
        ids = ogr.draw_id_samples(len(pts), 1000, rng)
        eq, _, _ = ogr.fit_parallel(pts, ids, 0.05)
        normals.append(torch.from_numpy(ogr.fix_up_normal(eq[:3])))
    conf = torch.tensor([0.1 if tuple(g.shape) == (1, 1) else 1.0 for g in case["ground_maps"]])
    cfg = ol.LossConfig(loss_functions=case["loss_functions"], use_confidence=case["use_confidence"])
    return ol.weak_loss(cfg, og.class_gather(leaves["deltas"], cls), og.class_gather(leaves["z"], cls).reshape(n),
                        og.class_gather(leaves["dims"], cls), og.class_gather(leaves["pose"], cls),
                        og.class_gather(leaves["uncert"], cls), src_boxes=case["src_boxes"], K4=K4,
                        virtual_to_real=v2r, prior_mean=p[:, 0], prior_std=p[:, 1], gt_boxes=case["gt_boxes"],
                        clamp=clamp.long(), num_boxes_per_image=nb, depth_maps=case["depth_maps"],
                        normals=torch.stack(normals), ground_conf=conf, gt_masks=case["gt_masks"],
                        at_which_mask_idx=case["at_which_mask_idx"])[0]


@pytest.mark.parametrize("path", GOLDEN)
def test_oracle_composition_matches_forward_cube_golden(path):
    z = np.load(path)
    case = rf.make_case(**ast.literal_eval(bytes(z["kw"]).decode()))
    leaves = {k: v.clone().requires_grad_(True) for k, v in case["head"].items()}
    L = _oracle_losses(case, leaves)
    want = {k[5:]: float(z[k]) for k in z.files if k.startswith("loss:")}
    assert set(L) == set(want)
    for k, v in want.items():
        assert abs(float(L[k].detach()) - v) <= 1e-5 * abs(v) + 3e-6, (k, float(L[k].detach()), v)
    if "segmentation" not in case["loss_functions"]:
        sum(L.values()).backward()
        for k in leaves:
            g = leaves[k].grad if leaves[k].grad is not None else torch.zeros_like(leaves[k])
            b = torch.from_numpy(z["grad:" + k])
            assert (g - b).abs().max().item() <= 1e-4 * max(b.abs().max().item(), 1e-12), k


@pytest.mark.ref
def test_forward_cube_golden_is_reproducible_from_the_reference():
    z = np.load(GOLDEN[0])
    case = rf.make_case(**ast.literal_eval(bytes(z["kw"]).decode()))
    L, grads = rf.run_reference_with_grads(case)
    for k in L:
        assert float(L[k]) == float(z["loss:" + k])
    for k in grads:
        assert np.array_equal(grads[k].numpy(), z["grad:" + k])
