"""CPU tests of the opt-in image ingest (cpu-ray-tracing-implementation_b200/images.py) and of the oracle's restatement of the
float / bilinear texture extensions (not part of the reference: the reference path stays the default)."""
import copy
import os

import numpy as np
import pytest
from conftest import abi, golden, pkg, same

images = pkg.images


def _write_exr_with_cv2(path, img, half=False, compression=None):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    flags = [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_HALF if half else cv2.IMWRITE_EXR_TYPE_FLOAT]
    if compression is not None:
        flags += [cv2.IMWRITE_EXR_COMPRESSION, compression]
    try:
        ok = cv2.imwrite(path, img[..., ::-1].copy(), flags)
    except Exception as e:
        pytest.skip(f"OpenCV without the OpenEXR codec: {e}")
    if not ok:
        pytest.skip("OpenCV could not write OpenEXR")
    return cv2


@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("comp", ["none", "zips", "zip"])
def test_read_exr_against_opencv(tmp_path, half, comp):
    rng = np.random.RandomState(3)
    img = (rng.rand(37, 53, 3) * 10 ** rng.uniform(-2, 2, size=(37, 53, 1))).astype(np.float32)
    path = str(tmp_path / "t.exr")
    import cv2 as _cv2

    code = {"none": _cv2.IMWRITE_EXR_COMPRESSION_NO, "zips": _cv2.IMWRITE_EXR_COMPRESSION_ZIPS, "zip": _cv2.IMWRITE_EXR_COMPRESSION_ZIP}[comp]
    cv2 = _write_exr_with_cv2(path, img, half=half, compression=code)
    got = images.read_exr(path)
    want = cv2.imread(path, cv2.IMREAD_UNCHANGED)[..., ::-1]
    assert got.shape == (37, 53, 3) and np.array_equal(got, want)
    if not half:
        assert np.array_equal(got, img)


def _earth_scene():
    g = golden("light_earth")
    s = copy.copy(g.scene)
    s.textures = g.scene.textures.copy()
    s.images = g.scene.images.copy()
    s.image_bytes = g.scene.image_bytes.copy()
    tex = int(np.nonzero(s.textures["kind"] == abi.TEX_PICTURE)[0][0])
    return g, s, tex


def test_float_copy_of_an_8bit_image_gives_identical_nearest_lookups(build_oracle):
    g, s, tex = _earth_scene()
    rows = g["tex_rows"] == tex
    uvp = g["tex_uvp"][rows]
    ref = g["tex_rgb"][rows]  # the reference's own lookups
    s.textures["aux"][tex] = images.float_copy_of_image(s, int(s.textures["aux"][tex]))
    assert s.images["bytes_per_pixel"][-1] == 12
    assert same(build_oracle.texture_sample(s, tex, uvp), ref).all()


def test_bilinear_lookup_properties(build_oracle):
    g, s, tex = _earth_scene()
    im = s.images[int(s.textures["aux"][tex])]
    W, H = int(im["width"]), int(im["height"])
    rng = np.random.RandomState(1)
    i, j = rng.randint(0, W, 500), rng.randint(0, H, 500)
    uvp = np.zeros((500, 5))
    uvp[:, 0], uvp[:, 1] = (i + 0.5) / W, 1 - (j + 0.5) / H  # texel centres
    nearest = build_oracle.texture_sample(s, tex, uvp)
    images.set_picture_filter(s, tex, True)
    linear = build_oracle.texture_sample(s, tex, uvp)
    assert np.allclose(linear, nearest, rtol=0, atol=1e-12)  # at a texel centre the filter returns the texel
    uvp[:, 0:2] = rng.rand(500, 2)
    mid = build_oracle.texture_sample(s, tex, uvp)
    assert (mid >= -1e-15).all() and (mid <= 255 / 256 + 1e-15).all()  # a convex combination of texels
    uvp[:, 0:2] = [[-0.3, 1.7]]  # far outside: clamped to the corner texel
    corner = build_oracle.texture_sample(s, tex, uvp[:1])
    uvp[:, 0:2] = [[0.5 / W, 1 - 0.5 / H]]
    assert np.allclose(corner, build_oracle.texture_sample(s, tex, uvp[:1]), atol=1e-12)


def test_adaptor_opt_ins_hand_over_the_float_image(build_oracle):
    """include/b200rt_adaptor.hpp: flattener::float_images / bilinear_pictures on the reference's own objects."""
    import refharness

    if not refharness.available() or not os.path.isdir(refharness.REFERENCE_ASSETS):
        pytest.skip("needs the reference checkout (assets/earthmap.jpg)")
    rs = refharness.RefScene("light_earth", width=32)
    plain = rs.scene
    ext = rs.reflatten(float_images=True, bilinear=True)
    t0 = int(np.nonzero(plain.textures["kind"] == abi.TEX_PICTURE)[0][0])
    assert ext.textures["kind"][t0] == abi.TEX_PICTURE_LINEAR and ext.images["bytes_per_pixel"][0] == 12
    im = plain.images[0]
    n = int(im["width"]) * int(im["height"]) * 3
    b = plain.image_bytes[int(im["offset"]):int(im["offset"]) + n]
    f = ext.image_bytes[int(ext.images["offset"][0]):int(ext.images["offset"][0]) + 4 * n].view(np.float32)
    # image::float_to_byte (src/image.h:105-109): the bytes the reference samples are the quantised floats
    q = np.where(f <= 0, 0, np.where(f >= 1, 255, (256.0 * f.astype(np.float64)).astype(np.int64))).astype(np.uint8)
    assert np.array_equal(q, b)
    rs.close()
