"""the presentation step (reference src/viewer/image.py:7-19)."""
import numpy as np
import pytest
from conftest import load_golden


def test_oracle_matches_reference_viewer_golden(oracle):
    g = load_golden("viewer_square")
    for n in (96, 130):
        assert np.array_equal(oracle.convert_array_to_image_array(g[f"x_{n}"]), g[f"img_{n}"])


@pytest.mark.gpu
def test_to_image_matches_reference_golden_and_oracle(oracle):
    from b200rt.viewer import to_image_array, convert_array_to_image
    g = load_golden("viewer_square")
    for n in (96, 130):
        assert np.array_equal(to_image_array(g[f"x_{n}"]), g[f"img_{n}"])
    rng = np.random.default_rng(3)
    for W, H in ((64, 64), (128, 192), (100, 36), (7, 5), (1, 1), (320, 180), (257, 64)):      # vector tiles, edge tiles, odd sizes
        x = rng.integers(0, 256, (3, W, H), dtype=np.uint8)
        exp = np.ascontiguousarray(x.transpose(2, 1, 0))                                         # out[y][x][c] = in[c][x][y]
        assert np.array_equal(to_image_array(x), exp), (W, H)
        assert np.array_equal(to_image_array(x, unswap_gb=True), exp[:, :, [0, 2, 1]]), (W, H)
    im = convert_array_to_image(g["x_96"])
    assert im.size == (96, 96) and im.mode == "RGB"


@pytest.mark.gpu
def test_to_image_device_array_path():
    import torch
    from b200rt import cuda
    from b200rt.viewer import to_image_array
    x = np.random.default_rng(1).integers(0, 256, (3, 192, 128), dtype=np.uint8)
    d = cuda.to_device(x)
    assert np.array_equal(to_image_array(d), x.transpose(2, 1, 0))
    assert np.array_equal(to_image_array(torch.from_numpy(x).cuda()), x.transpose(2, 1, 0))
