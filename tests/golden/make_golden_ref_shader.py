"""Writes tests/golden/ref_shader_<scene>.npz: small frames rendered by the reference's own shader text compiled as C++
(oracle/_ref/libokrt_ref_shader.so; needs /root/reference).  The fixtures let a machine without the reference tree check the
oracle — and through it the CUDA kernels — against the reference's shader.  python tests/golden/make_golden_ref_shader.py"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
okrt = importlib.import_module("gpu-raytracer_b200")
from oracle import binding                                   # noqa: E402
from tests import test_cpu_ref_shader as t                   # noqa: E402

ref = binding.load_ref_shader()
if ref is None:
    raise SystemExit("the reference tree is not present")
oracle = binding.load()
cases = t.scene_cases(okrt)
for name in ("testscene", "sponza", "glass"):
    scene, rd, bounces = t._scene_and_rd(oracle, cases[name])
    frames, accum, result = [1, 3], [], []
    for f in frames:
        rd[0]["numAccumulationFrames"] = f
        r = ref.render(scene, rd, bounces)
        accum.append(r["accum"]); result.append(r["result"])
    out = os.path.join(ROOT, "tests", "golden", "ref_shader_%s.npz" % name)
    np.savez_compressed(out, frames=np.array(frames, np.uint32), bounces=np.uint32(bounces), accum=np.stack(accum), result=np.stack(result),
                        source=np.bytes_(ref.lib.ref_shader_description()))
    print(out, os.path.getsize(out), "bytes")
