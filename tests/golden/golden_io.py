"""Read/write the committed golden fixtures (small .npz files)."""
import json

import numpy as np
import torch

TENSOR_KEYS = ("deltas", "z_raw", "dims_raw", "pose", "uncert", "src_boxes", "K4", "virtual_to_real",
               "prior_mean", "prior_std", "gt_boxes", "clamp", "normals", "ground_conf")


def save_loss_case(path, cfg_kw, inp, losses, grads):
    d = {"in_" + k: inp[k].numpy() for k in TENSOR_KEYS}
    d["in_depth"] = np.stack([x.numpy() for x in inp["depth_maps"]]).astype(np.float16)
    d["meta"] = np.frombuffer(json.dumps({"cfg": cfg_kw, "nb": inp["num_boxes_per_image"],
                                          "loss_keys": list(losses)}).encode(), dtype=np.uint8)
    d["loss_vals"] = np.array([float(v) for v in losses.values()], dtype=np.float64)
    for k, v in grads.items():
        d["grad_" + k] = v.numpy()
    np.savez_compressed(path, **d)


def load_loss_case(path):
    z = np.load(path)
    meta = json.loads(bytes(z["meta"]).decode())
    inp = {k: torch.from_numpy(z["in_" + k]) for k in TENSOR_KEYS}
    inp["depth_maps"] = [torch.from_numpy(x.astype(np.float32)) for x in z["in_depth"]]
    inp["num_boxes_per_image"] = meta["nb"]
    cfg_kw = dict(meta["cfg"])
    if "loss_functions" in cfg_kw:
        cfg_kw["loss_functions"] = tuple(cfg_kw["loss_functions"])
    losses = {k: torch.tensor(v, dtype=torch.float32) for k, v in zip(meta["loss_keys"], z["loss_vals"])}
    grads = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad_")}
    return cfg_kw, inp, losses, grads
