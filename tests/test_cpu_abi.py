"""CPU tests of the C-ABI library: it loads without a GPU, exports every symbol include/okrt.h declares, validates
arguments, and refuses to render without a device (no CPU fallback)."""
import ctypes

import pytest
import os
import re
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "okrt.h")).read()
    return sorted(set(re.findall(r"OKRT_API[^;(]*?\b(okrt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(okrt):
    declared = _declared_symbols()
    assert len(declared) >= 55
    lib = ctypes.CDLL(okrt.abi.LIB_PATH)
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    bound = {name for name, _, _ in okrt.abi.SYMBOLS}
    assert set(declared) == bound, (set(declared) ^ bound)


def test_only_abi_symbols_are_exported(okrt):
    out = subprocess.run(["nm", "-D", "--defined-only", okrt.abi.LIB_PATH], capture_output=True, text=True).stdout
    names = [l.split()[-1] for l in out.splitlines() if " T " in l]
    extra = [n for n in names if not n.startswith("okrt_") and n not in ("_init", "_fini")]
    assert not extra, extra[:10]


def test_library_is_built_for_sm100a_only(okrt):
    out = subprocess.run(["cuobjdump", "--list-elf", okrt.abi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert "100a" in archs
    sass = subprocess.run(["cuobjdump", "-sass", "-arch", "sm_100a", okrt.abi.LIB_PATH], capture_output=True, text=True).stdout
    for kernel in ("traceFastKernel", "shadeKernel", "generateKernel", "pathTraceKernel", "traceExactKernel"):
        assert kernel in sass, kernel
    assert "LDG.E.ENL2.256" in sass                   # the 64-byte child-pair nodes are fetched as two 256-bit loads
    assert "TLD" in sass                              # triangle records go through the texture pipe


def test_no_device_means_no_renderer(okrt):
    abi = okrt.abi
    L = abi.lib()
    if L.okrt_device_count() > 0:
        return
    h = ctypes.c_void_p()
    rc = L.okrt_create(0, 64, 64, ctypes.byref(h))
    assert rc == 5 and not h.value                    # OKRT_ERR_NO_DEVICE: there is no CPU path
    assert b"no CUDA device" in L.okrt_last_error()


def test_no_device_means_no_device_builder(okrt):
    """okrt_geometry_build_device (row f2) has no CPU fallback either: without a GPU it fails loudly."""
    L = okrt.abi.lib()
    if L.okrt_device_count() > 0:
        return
    rm, _, _ = okrt.scenes.glass_scene(16, 8)
    with pytest.raises(okrt.abi.OkrtError) as err:
        okrt.Geometry(rm.meshes, 30, 5, device=0)
    assert err.value.code == 5 and "no CPU fallback" in str(err.value)


def test_argument_validation(okrt):
    abi = okrt.abi
    L = abi.lib()
    assert L.okrt_create(0, 0, 64, ctypes.byref(ctypes.c_void_p())) == 1
    assert L.okrt_build_bvh(None, 3, None, 30, 5, None, 0, None, None) == 1
    assert L.okrt_set_options(None, None) == 1 and L.okrt_set_render_data(None, None) == 1
    assert L.okrt_scale_texture(None, 1, 1, None, 1, 1) == 1
    assert L.okrt_get_num_accumulation_frames(None) == 0
    g = ctypes.c_void_p()
    assert L.okrt_geometry_build(None, 0, 30, 5, ctypes.byref(g)) == 0          # empty geometry is legal
    assert L.okrt_geometry_num_triangles(g) == 0
    t = ctypes.c_void_p()
    bad = np.zeros(1, abi.MESH_ENTITY); bad[0]["meshID"] = 3
    assert L.okrt_octree_build(g, abi.ptr(bad), 1, None, 0, 2, 50, ctypes.byref(t)) == 1
    assert b"meshID" in L.okrt_last_error()
    L.okrt_geometry_free(g)
    # node capacity too small -> OKRT_ERR_CAPACITY, nothing written past the end
    mesh = okrt.scenes.make_mesh(*okrt.scenes.icosphere(1, 1.0))
    bb = np.zeros(1, abi.AABB); bb[0]["min"], bb[0]["max"] = mesh.boundingBox
    nodes = np.zeros(2, abi.BVH_NODE); order = np.zeros(80, np.uint32); cnt = ctypes.c_uint32()
    assert L.okrt_build_bvh(abi.ptr(mesh.trianglesPos), 80, abi.ptr(bb), 30, 5, abi.ptr(nodes), 2, ctypes.byref(cnt), abi.ptr(order)) == 4


def test_defaults_match_reference(okrt):
    abi = okrt.abi
    rd = abi.default_render_data()[0]
    assert rd["accumulationEnabled"] == 1 and rd["debugMaxCount"] == 500 and rd["numAccumulationFrames"] == 0    # RayTracer.h:131-155
    o = abi.default_options()[0]
    assert o["maxBounces"] == 1 and o["traversal"] == 0 and o["coverAllPixels"] == 0 and o["frameStride"] == 1      # NUM_BOUNCES (1) at HEAD
    m = abi.default_material()
    assert m["roughness"] == 1 and m["specular"] == 1 and m["indexOfRefraction"] == 1 and m["normalMapIdx"] == abi.INVALID_UINT


def test_empty_and_degenerate_meshes(okrt):
    """Edge cases of the builder: no triangles, one triangle, identical triangles (SAH refuses), maxDepth 1."""
    abi = okrt.abi
    sc = okrt.scenes
    g = okrt.Geometry([sc.make_mesh(np.zeros((0, 3, 3))), sc.make_mesh(np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], float)),
                       sc.make_mesh(np.repeat(np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], float), 20, axis=0))], 1, 5)
    assert list(g.meshDescs["numBvhNodes"]) == [1, 1, 1]
    assert g.bvhNodes[0]["triStart"] == 0 and g.bvhNodes[0]["triEnd"] == 0
    assert g.bvhNodes[2]["triEnd"] - g.bvhNodes[2]["triStart"] == 20


def test_cmake_lists_every_source_the_in_tree_build_compiles():
    """CMakeLists.txt and gpu-raytracer_b200/build.py are two descriptions of one library: same translation units."""
    import importlib.util
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("okrt_build_for_test", os.path.join(root, "gpu-raytracer_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    cmake = open(os.path.join(root, "CMakeLists.txt")).read()
    listed = set(re.findall(r"\$\{CSRC\}/(\w+\.(?:cu|cpp))", cmake))
    built = {os.path.basename(s) for s in list(b.HOST_SOURCES) + list(b.DEVICE_SOURCES)}
    assert listed == built, (sorted(listed), sorted(built))
