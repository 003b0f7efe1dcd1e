"""PIL's NEAREST index tables (host function of the C ABI) against PIL itself, and the
no_ground_idx.csv reader. dataset_mapper.py:126-145 resizes with Image.resize(size, Image.NEAREST)."""
import numpy as np
import pytest

PIL = pytest.importorskip("PIL")
from PIL import Image  # noqa: E402


@pytest.mark.parametrize("n_in,n_out", [(480, 512), (640, 683), (853, 640), (640, 853), (1, 7), (7, 1),
                                        (375, 1242), (1242, 375), (333, 1000), (1000, 333), (530, 730),
                                        (427, 512), (512, 427), (640, 640), (3, 2), (1277, 1279)])
def test_table_matches_pil(n_in, n_out):
    from wcr.io import pil_nearest_table
    # an image whose pixel value is its own column index: the resized row IS the table
    src = np.arange(n_in, dtype=np.float32)[None, :].repeat(2, 0)
    want = np.array(Image.fromarray(src).resize((n_out, 2), Image.NEAREST))[0].astype(np.int32)
    got = pil_nearest_table(n_in, n_out)
    assert np.array_equal(got, want)
    # and along y
    src = np.arange(n_in, dtype=np.float32)[:, None].repeat(2, 1)
    want = np.array(Image.fromarray(src).resize((2, n_out), Image.NEAREST))[:, 0].astype(np.int32)
    assert np.array_equal(got, want)


def test_no_ground_csv(tmp_path):
    from wcr.io import read_no_ground_idx
    p = tmp_path / "no_ground_idx.csv"
    p.write_text("img_id\n12\n7\n100003\n")
    assert read_no_ground_idx(str(p)) == {12, 7, 100003}
    assert read_no_ground_idx(str(tmp_path / "missing.csv")) == set()
