"""CPU suite: headless image output -- write_pfm / write_png of a float4[P] buffer, no GL."""
import numpy as np
import pytest

from mcrt_b200 import pt


def _image(w, h):
    rng = np.random.RandomState(3)
    img = rng.rand(h * w, 4).astype(np.float32) * 1.5 - 0.2        # values below 0 and above 1 are clamped
    img[5] = [np.nan, np.inf, -np.inf, 1.0]
    img[6] = [0.5, 127.5 / 255.0, 1.0, 0.0]
    return img


def test_png_is_the_rgba8_framebuffer_conversion_of_the_float_image(tmp_path):
    """Engine::takeScreenshot reads an RGBA8 framebuffer: clamp to [0, 1], x 255, round to nearest."""
    Image = pytest.importorskip("PIL.Image")
    w, h = 37, 19
    img = _image(w, h)
    path = str(tmp_path / "shot.png")
    pt.write_png(path, img, w, h)
    with Image.open(path) as im:
        assert im.size == (w, h) and im.mode == "RGBA"
        got = np.asarray(im, np.uint8).reshape(h * w, 4)
    rgb = np.nan_to_num(img[:, :3], nan=0.0, posinf=1.0, neginf=0.0)
    want = np.rint(np.clip(rgb, 0, 1) * 255).astype(np.uint8)
    assert np.array_equal(got[:, :3], want)
    assert (got[:, 3] == 255).all()
    assert tuple(got[5, :3]) == (0, 255, 0) and tuple(got[6, :3]) == (128, 128, 255)
    # row 0 of the buffer is the top row of the file, as in write_pfm (which stores bottom-up and therefore flips)
    assert np.array_equal(got.reshape(h, w, 4)[0, :, :3], want.reshape(h, w, 3)[0])


def test_pfm_round_trip_is_exact_and_bottom_up(tmp_path):
    w, h = 16, 9
    img = _image(w, h)
    img[5] = [0.25, 0.5, 0.75, 1.0]
    path = str(tmp_path / "acc.pfm")
    pt.write_pfm(path, img, w, h)
    with open(path, "rb") as f:
        assert f.readline() == b"PF\n"
        assert f.readline() == b"%d %d\n" % (w, h)
        assert float(f.readline()) < 0                      # negative scale = little endian
        data = np.frombuffer(f.read(), "<f4").reshape(h, w, 3)
    assert np.array_equal(data[::-1].reshape(-1, 3).view(np.uint32), img[:, :3].copy().view(np.uint32))
