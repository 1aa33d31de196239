"""Rows a7 / a13 with the reference's own pixels (tests/golden/assets_small.npz: Skybox2.jpg cross and four Sponza textures,
down-sampled by tests/golden/make_golden_assets.py): cross -> cube faces (RayTracer.cpp:344-432), rescale to the texture
array (RayTracer.cpp:310-342, DX11.cpp:232-270), then a textured, sky-lit scene on the GPU against the oracle."""
import os

import numpy as np
import pytest

from tests import helpers

FIXTURE = os.path.join(os.path.dirname(__file__), "golden", "assets_small.npz")


def _assets(okrt):
    z = np.load(FIXTURE)
    textures = [np.ascontiguousarray(z[k]) for k in ("tex_sponza_floor_a_diff", "tex_spnza_bricks_a_diff", "tex_sponza_fabric_blue_diff", "tex_sponza_column_a_spec")]
    return np.ascontiguousarray(z["cross"]), textures


def test_real_assets_through_host_builders_equal_oracle(okrt, oracle):
    abi = okrt.abi
    cross, textures = _assets(okrt)
    faces = okrt.environmentCrossToFaces(cross)
    assert faces.shape == (6, 128, 128, 4)
    of = np.zeros_like(faces)
    oracle.lib.oracle_environment_cross_to_faces(abi.ptr(cross), cross.shape[1], cross.shape[0], abi.ptr(of))
    assert np.array_equal(faces, of)
    # +X is the cell at column 2, row 1 of the cross; -Z at column 3, row 1 (RayTracer.cpp:380-417) [Q13]
    assert np.array_equal(faces[0], cross[128:256, 256:384]) and np.array_equal(faces[5], cross[128:256, 384:512])
    arr = okrt.scaleTexturesToArray(textures)
    n = int(np.float32(np.sqrt(np.float32((160 * 160 + 128 * 128 + 96 * 128 + 64 * 64) // 4))))
    assert arr.shape == (4, n, n, 4)
    for i, t in enumerate(textures):
        o = np.zeros((n, n, 4), np.uint8)
        oracle.lib.oracle_scale_texture(abi.ptr(t), t.shape[1], t.shape[0], abi.ptr(o), n, n)
        assert np.array_equal(o, arr[i]), i


@pytest.mark.gpu
def test_textured_sky_lit_scene_with_real_assets(okrt, oracle):
    abi = okrt.abi
    cross, textures = _assets(okrt)
    faces = okrt.environmentCrossToFaces(cross)
    rm, sc, info = okrt.scenes.test_scene(with_textures=True)
    rm.textures = textures                                   # albedo 0 / 1, specular 3 of the TestScene now read Sponza's pixels
    W, H = 320, 180
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    rd = host["rd"].copy(); rd[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=faces), rd, 6, aov_dtype=abi.AOV)
    assert o["visits"]["textureTaps"] > 0
    for pipeline in (0, 1):
        r = helpers.Renderer(okrt, host, W, H, envFaces=faces)
        r.options(maxBounces=6, pipeline=pipeline, writeAov=1)
        r.render(host["rd"].copy())
        assert np.array_equal(r.aov()["triIdx"], o["aov"]["triIdx"])
        assert np.array_equal(r.accumulation().view(np.uint32), o["accum"].view(np.uint32)), pipeline
        r.close()
    # the sky is Skybox2's: a primary ray that misses everything returns the cube texel the oracle samples
    miss = o["aov"]["hitKind"] == 0
    assert miss.any() and o["accum"][miss][:, :3].max() > 0.1
