"""Generates tests/golden/assets_small.npz from assets the reference ships (run HERE, where /root/reference exists):
    python tests/golden/make_golden_assets.py
* cross      resources/environmentMaps/Skybox2.jpg (4096x3072 horizontal cross, RayTracer.cpp:344-432), box-filtered 8x -> 512x384 RGBA8
* textures   four of resources/Sponza/textures (sponza_floor_a_diff, spnza_bricks_a_diff, sponza_fabric_blue_diff,
             sponza_column_a_spec), box-filtered to different sizes so that loadTextureData's rescale to
             N = uint(sqrt(sum(w*h) / count)) (RayTracer.cpp:310-342) has work to do
The GPU box has no /root/reference: rows a7 / a13 are exercised there with these real pixels (tests/test_gpu_assets.py)."""
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/resources"


def rgba(path, size):
    img = Image.open(path).convert("RGBA").resize(size, Image.BOX)
    return np.ascontiguousarray(np.asarray(img, np.uint8))


def main():
    cross = rgba(os.path.join(REF, "environmentMaps", "Skybox2.jpg"), (512, 384))
    tex = {"sponza_floor_a_diff": (160, 160), "spnza_bricks_a_diff": (128, 128), "sponza_fabric_blue_diff": (96, 128), "sponza_column_a_spec": (64, 64)}
    arrays = {"cross": cross}
    for name, size in tex.items():
        arrays["tex_" + name] = rgba(os.path.join(REF, "Sponza", "textures", name + ".png"), size)
    np.savez_compressed(os.path.join(HERE, "assets_small.npz"), **arrays)
    print({k: v.shape for k, v in arrays.items()}, os.path.getsize(os.path.join(HERE, "assets_small.npz")))


if __name__ == "__main__":
    main()
