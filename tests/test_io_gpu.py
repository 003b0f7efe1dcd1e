"""On-disk formats -> device maps: npz files written exactly as the
reference writes them, decoded and resized on the device, against PIL on the host
(dataset_mapper.py:126-145). Bit-exact."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
PIL = pytest.importorskip("PIL")
from PIL import Image  # noqa: E402


def _mapper(a, hw):
    """dataset_mapper.py:127-128 / :138-139."""
    return np.array(Image.fromarray(a).resize((hw[1], hw[0]), Image.NEAREST))


@pytest.mark.parametrize("src_hw,dst_hw", [((480, 640), (512, 683)), ((530, 730), (480, 661)),
                                           ((427, 561), (427, 561)), ((375, 1242), (512, 1696)),
                                           ((64, 48), (640, 480))])
def test_resize_matches_pil(cuda_lib, src_hw, dst_hw):
    from wcr.io import resize_nearest
    rng = np.random.default_rng(src_hw[0] + dst_hw[1])
    depth = rng.uniform(0.3, 20.0, src_hw).astype(np.float32)
    mask = rng.random(src_hw) < 0.4
    got_d = resize_nearest(torch.from_numpy(depth).cuda(), dst_hw).cpu().numpy()
    got_m = resize_nearest(torch.from_numpy(mask).cuda(), dst_hw).cpu().numpy()
    assert np.array_equal(got_d, _mapper(depth, dst_hw))
    assert got_m.dtype == np.bool_ and np.array_equal(got_m, _mapper(mask, dst_hw))


def test_load_maps_from_reference_file_layout(cuda_lib, tmp_path):
    """generate_depth_maps.py:85-86 / generate_ground_segmentations.py:200, 205-206 file layout."""
    from wcr import io as wio
    from wcr.ground import normal_vector_from_maps
    from wcr import synth
    ddir, gdir = tmp_path / "depth_maps", tmp_path / "ground_maps"
    ddir.mkdir(); gdir.mkdir()
    ids, sizes, want_d, want_g, Ks = [101, 102, 103], [(512, 683), (480, 640), (384, 512)], [], [], []
    for k, (img_id, hw) in enumerate(zip(ids, sizes)):
        im = synth.make_image(40 + k, 480, 640, n_boxes=2, dummy_ground=False)
        depth, mask = im["depth"].numpy(), im["ground"].numpy().astype(bool)
        np.savez_compressed(ddir / f"{img_id}.npz", depth=depth)
        if img_id != 102:
            np.savez_compressed(gdir / f"{img_id}.npz", mask=mask)
        want_d.append(_mapper(depth, hw))
        want_g.append(None if img_id == 102 else _mapper(mask, hw))
        Ks.append(im["K"])
    (tmp_path / "no_ground_idx.csv").write_text("img_id\n102\n")
    no_ground = wio.read_no_ground_idx(str(tmp_path / "no_ground_idx.csv"))
    depths, grounds = wio.load_maps(ids, sizes, depth_dir=str(ddir), ground_dir=str(gdir), no_ground=no_ground)
    for d, g, wd, wg in zip(depths, grounds, want_d, want_g):
        assert np.array_equal(d.cpu().numpy(), wd)
        assert (g is None) == (wg is None)
        if g is not None:
            assert np.array_equal(g.cpu().numpy(), wg)
    # and the maps feed the path: up vectors come out finite for the images with a ground map
    nv = normal_vector_from_maps(grounds, depths, torch.stack(Ks).cuda())
    assert nv.shape == (3, 3) and torch.isfinite(nv).all()
