"""A small render through every kernel of the wavefront pipeline (generate, traversal, reference-order fallback, shade, accumulate,
overlay) for compute-sanitizer: python tools/sanitizer_case.py.  Prints the segment count and a digest of the accumulation image."""
import hashlib, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
okrt = importlib.import_module("gpu-raytracer_b200")
abi = okrt.abi
W, H = 128, 72
rm, sc, info = okrt.scenes.sponza_standin(4000, texture_size=32)
host = okrt.harness.build_host_side(okrt, rm, sc, info, W, H)
r = okrt.harness.Renderer(okrt, host, W, H, envFaces=okrt.scenes.procedural_sky_faces(16))
r.options(maxBounces=3)
rd = host["rd"].copy()
for frames in (2, 1, 3):                                       # three batches on three lanes
    rd[0]["numAccumulationFrames"] = r.L.okrt_get_num_accumulation_frames(r.h)
    abi.check(r.L.okrt_set_render_data(r.h, abi.ptr(rd)))
    abi.check(r.L.okrt_render_frames(r.h, frames))
abi.check(r.L.okrt_sync(r.h))
acc = r.accumulation()
print("segments", int(r.stats()["raySegments"]), "fallback", int(r.stats()["fallbackSegments"]), "u1 sha1", hashlib.sha1(acc.tobytes()).hexdigest()[:16])
r.close()
