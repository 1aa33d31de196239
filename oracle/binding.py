"""ctypes binding of the CPU oracle (oracle/_build/libokrt_oracle.so) and of the compiled reference
(oracle/_ref/libokrt_ref.so).  TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__ (build of the
checker, smoke()) and bench.py's CPU legs (cpu_baseline, parity check, --impl reference).  Nothing under
gpu-raytracer_b200/ imports this package."""
import ctypes
import os
import subprocess

import numpy as np

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(ORACLE_DIR)
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "libokrt_oracle.so")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "libokrt_ref.so")
REF_OCTREE_LIB = os.path.join(ORACLE_DIR, "_ref", "libokrt_ref_octree.so")
REF_SHADER_LIB = os.path.join(ORACLE_DIR, "_ref", "libokrt_ref_shader.so")

MATH_CANONICAL, MATH_LIBM = 0, 1


class OracleScene(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("triPos", "triInfo", "bvhNodes", "textures", "envFaces", "octNodes", "spheres",
                                              "meshes", "dirLights", "pointLights", "spotLights")] + \
               [(n, ctypes.c_uint32) for n in ("numTris", "numBvhNodes", "texSize", "texLayers", "envW", "envH", "numOctNodes",
                                              "numSpheres", "numMeshes", "numDirLights", "numPointLights", "numSpotLights")]


def build_oracle(force=False):
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("okrt_oracle.cpp", "oracle_math.h")]
    stale = force or not os.path.exists(ORACLE_LIB) or any(os.path.getmtime(s) > os.path.getmtime(ORACLE_LIB) for s in srcs)
    if stale:
        subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True)
    return ORACLE_LIB


def build_ref():
    """Compiles the reference's BvhBuilder.cpp where /root/reference exists; otherwise keeps the prebuilt file."""
    if os.path.exists("/root/reference/source/Graphics/BvhBuilder.cpp"):
        shims = [os.path.join(ORACLE_DIR, f) for f in ("ref_shim.cpp", "ref_octree_shim.cpp", "ref_shader_driver.cpp",
                                                       os.path.join("ref_stubs", "hlsl_shim.h"), "oracle_math.h", "Makefile")]
        libs = (REF_LIB, REF_OCTREE_LIB, REF_SHADER_LIB)
        stale = not all(os.path.exists(l) for l in libs)
        stale = stale or any(os.path.getmtime(f) > min(os.path.getmtime(l) for l in libs) for f in shims)
        if stale:
            subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True)
    return REF_LIB if os.path.exists(REF_LIB) else None


_P, _U, _I, _F = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_float


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        lib.oracle_render.restype = _U
        lib.oracle_render.argtypes = [ctypes.POINTER(OracleScene), _P, _U, _U, _U, _P, _P, _P, _U, _U, _I, _I, ctypes.POINTER(ctypes.c_uint64)]
        lib.oracle_render_rows.restype = _U
        lib.oracle_render_rows.argtypes = [ctypes.POINTER(OracleScene), _P, _U, _U, _U, _P, _P, _P, _P, _U, _I, _I, ctypes.POINTER(ctypes.c_uint64)]
        lib.oracle_trace_rays.restype = None
        lib.oracle_trace_rays.argtypes = [ctypes.POINTER(OracleScene), _P, _P, _U, _P, _I]
        lib.oracle_last_visit_totals.restype = None
        lib.oracle_last_visit_totals.argtypes = [_P]
        lib.oracle_pcg_hash.restype = _U
        lib.oracle_pcg_hash.argtypes = [ctypes.POINTER(_U)]
        lib.oracle_random_float.restype = _F
        lib.oracle_random_float.argtypes = [ctypes.POINTER(_U)]
        lib.oracle_ray_triangle.restype = _F
        lib.oracle_ray_triangle.argtypes = [_P] * 6
        lib.oracle_ray_aabb_dist.restype = _F
        lib.oracle_ray_aabb_dist.argtypes = [_P] * 4
        lib.oracle_ray_sphere.restype = _F
        lib.oracle_ray_sphere.argtypes = [_P, _P, _P, _F]
        lib.oracle_math_eval.restype = None
        lib.oracle_math_eval.argtypes = [_I, _P, _P, _P, _U]
        lib.oracle_sample_texture.restype = None
        lib.oracle_sample_texture.argtypes = [ctypes.POINTER(OracleScene), _U, _F, _F, _P]
        lib.oracle_sample_environment.restype = None
        lib.oracle_sample_environment.argtypes = [ctypes.POINTER(OracleScene), _P, _P]
        lib.oracle_build_bvh.restype = _U
        lib.oracle_build_bvh.argtypes = [_P, _U, _P, _P, _U, _U, _P, _P]
        lib.oracle_transform_matrix.restype = None
        lib.oracle_transform_matrix.argtypes = [_P, _P, _P, _P]
        lib.oracle_calculate_projection_data.restype = None
        lib.oracle_calculate_projection_data.argtypes = [_P, _P, _F, _F, _F, _U, _U, _P]
        lib.oracle_build_octree.restype = _U
        lib.oracle_build_octree.argtypes = [_P, _P, _U, _P, _U, _P, _U, _U, _U, _P, _U, _P, ctypes.POINTER(_U)]
        lib.oracle_texture_array_size.restype = _U
        lib.oracle_texture_array_size.argtypes = [_P, _P, _U]
        lib.oracle_scale_texture.restype = None
        lib.oracle_scale_texture.argtypes = [_P, _U, _U, _P, _U, _U]
        lib.oracle_environment_cross_to_faces.restype = None
        lib.oracle_environment_cross_to_faces.argtypes = [_P, _U, _U, _P]

    # ---- helpers over numpy arrays (dtypes from the product's _abi module are byte-compatible) ----
    @staticmethod
    def _p(a):
        return None if a is None or len(a) == 0 else a.ctypes.data_as(ctypes.c_void_p)

    def make_scene(self, triPos, triInfo, bvhNodes, octNodes, instances, spheres=None, textures=None, envFaces=None,
                   dirLights=None, pointLights=None, spotLights=None):
        keep = dict(triPos=triPos, triInfo=triInfo, bvhNodes=bvhNodes, octNodes=octNodes, instances=instances, spheres=spheres,
                    textures=textures, envFaces=envFaces, dirLights=dirLights, pointLights=pointLights, spotLights=spotLights)
        s = OracleScene()
        s.triPos, s.triInfo, s.bvhNodes = self._p(triPos), self._p(triInfo), self._p(bvhNodes)
        s.octNodes, s.meshes, s.spheres = self._p(octNodes), self._p(instances), self._p(spheres)
        s.textures = self._p(textures.reshape(-1)) if textures is not None and textures.size else None
        s.envFaces = self._p(envFaces.reshape(-1)) if envFaces is not None and envFaces.size else None
        s.dirLights, s.pointLights, s.spotLights = self._p(dirLights), self._p(pointLights), self._p(spotLights)
        s.numTris, s.numBvhNodes, s.numOctNodes = len(triPos), len(bvhNodes), len(octNodes)
        s.numMeshes = len(instances)
        s.numSpheres = 0 if spheres is None else len(spheres)
        s.numDirLights = 0 if dirLights is None else len(dirLights)
        s.numPointLights = 0 if pointLights is None else len(pointLights)
        s.numSpotLights = 0 if spotLights is None else len(spotLights)
        if textures is not None and textures.size:
            s.texLayers, s.texSize = textures.shape[0], textures.shape[1]
        if envFaces is not None and envFaces.size:
            s.envH, s.envW = envFaces.shape[1], envFaces.shape[2]
        s._keep = keep
        return s

    def trace_rays(self, scene, rd, rays, threads=0):
        """findClosestHit for an [n, 6] float32 array of rays; returns an [n, 6] uint32 array
        (distance, u, v as float bits, triangle, kind, index)."""
        rays = np.ascontiguousarray(rays, np.float32)
        out = np.zeros((len(rays), 6), np.uint32)
        self.lib.oracle_trace_rays(ctypes.byref(scene), rd.ctypes.data_as(_P), rays.ctypes.data_as(_P), len(rays), out.ctypes.data_as(_P), threads)
        return out

    def render(self, scene, rd, maxBounces, seedFrame=None, coverAll=0, accum=None, result=None, aov_dtype=None,
               rows=None, threads=0, math=MATH_CANONICAL, row_list=None):
        """One frame.  rd: 1-element RENDER_DATA array.  Returns dict(accum, result, aov, segments, overflow)."""
        W, H = int(rd[0]["textureDims"][0]), int(rd[0]["textureDims"][1])
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        if result is None:
            result = np.zeros((H, W, 4), np.uint8)
        aov = np.zeros((H, W), aov_dtype) if aov_dtype is not None else None
        seg = ctypes.c_uint64(0)
        if seedFrame is None:
            seedFrame = int(rd[0]["numAccumulationFrames"])
        if row_list is not None:
            rl = np.ascontiguousarray(row_list, np.uint32)
            ovf = self.lib.oracle_render_rows(ctypes.byref(scene), rd.ctypes.data_as(_P), maxBounces, seedFrame, coverAll,
                                              accum.ctypes.data_as(_P), result.ctypes.data_as(_P),
                                              None if aov is None else aov.ctypes.data_as(_P), rl.ctypes.data_as(_P), len(rl), threads, math, ctypes.byref(seg))
        else:
            y0, y1 = (0, H) if rows is None else rows
            ovf = self.lib.oracle_render(ctypes.byref(scene), rd.ctypes.data_as(_P), maxBounces, seedFrame, coverAll,
                                         accum.ctypes.data_as(_P), result.ctypes.data_as(_P),
                                         None if aov is None else aov.ctypes.data_as(_P), y0, y1, threads, math, ctypes.byref(seg))
        v = np.zeros(7, np.uint64)
        self.lib.oracle_last_visit_totals(v.ctypes.data_as(_P))
        visits = dict(zip(("segments", "octNodes", "instances", "bvhNodes", "triangles", "hits", "textureTaps"), (int(x) for x in v)))
        return dict(accum=accum, result=result, aov=aov, segments=seg.value, overflow=ovf, visits=visits)


def oracle_scene(oracle, host, envFaces=None, dirLights=None, pointLights=None, spotLights=None):
    """The oracle's view of the SAME arrays the product's harness uploads (harness.build_host_side)."""
    g, t = host["geometry"], host["octree"]
    return oracle.make_scene(g.trianglesPos, g.trianglesInfo, g.bvhNodes, t.nodes, t.instances, spheres=host["spheres"],
                             textures=host["textures"] if host["textures"].size else None, envFaces=envFaces,
                             dirLights=dirLights, pointLights=pointLights, spotLights=spotLights)


_oracle = None


def load():
    global _oracle
    if _oracle is None:
        _oracle = Oracle(ctypes.CDLL(build_oracle()))
    return _oracle


class RefShader:
    """The reference's own compute shader text compiled as C++ (oracle/ref_shader_driver.cpp, oracle/ref_stubs/hlsl_shim.h)."""

    def __init__(self, lib):
        self.lib = lib
        lib.ref_shader_render.restype = _U
        lib.ref_shader_render.argtypes = [ctypes.POINTER(OracleScene), _P, _U, _U, _P, _P, _P, _I]
        lib.ref_shader_render_rows.restype = _U
        lib.ref_shader_render_rows.argtypes = [ctypes.POINTER(OracleScene), _P, _U, _U, _P, _P, _P, _P, _U, _I]
        lib.ref_shader_description.restype = ctypes.c_char_p

    def render(self, scene, rd, maxBounces, coverAll=0, accum=None, threads=0, row_list=None):
        """One RayTracer::render(): rd as uploaded (numAccumulationFrames already advanced).  accum (H, W, 4) float32 is u1, read
        and written.  Returns dict(accum, result_float, result): result_float is what main() stores into u0 before the UNORM8
        conversion, result its RGBA8 form."""
        W, H = int(rd[0]["textureDims"][0]), int(rd[0]["textureDims"][1])
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        result_f = np.zeros((H, W, 4), np.float32)
        result = np.zeros((H, W, 4), np.uint8)
        rl = None if row_list is None else np.ascontiguousarray(row_list, np.uint32)
        self.lib.ref_shader_render_rows(ctypes.byref(scene), rd.ctypes.data_as(_P), int(maxBounces), int(coverAll),
                                        accum.ctypes.data_as(_P), result_f.ctypes.data_as(_P), result.ctypes.data_as(_P),
                                        None if rl is None else rl.ctypes.data_as(_P), 0 if rl is None else len(rl), int(threads))
        return dict(accum=accum, result_float=result_f, result=result)


_ref_shader = None


def load_ref_shader():
    """None where neither /root/reference nor a prebuilt oracle/_ref/libokrt_ref_shader.so exists."""
    global _ref_shader
    if _ref_shader is None:
        build_ref()
        if not os.path.exists(REF_SHADER_LIB):
            return None
        _ref_shader = RefShader(ctypes.CDLL(REF_SHADER_LIB))
    return _ref_shader


_ref_octree = None


def load_ref_octree():
    """The reference's own RayTracer::createOctTree / loadOctTree / refitOctTreeNode (oracle/ref_octree_shim.cpp); None where
    neither /root/reference nor a prebuilt oracle/_ref/libokrt_ref_octree.so exists."""
    global _ref_octree
    if _ref_octree is None:
        build_ref()
        if not os.path.exists(REF_OCTREE_LIB):
            return None
        lib = ctypes.CDLL(REF_OCTREE_LIB)
        lib.ref_build_octree.restype = _U
        lib.ref_build_octree.argtypes = [_P, _P, _U, _P, _U, _P, _U, _U, _U, _P, _U, _P, _U, ctypes.POINTER(_U)]
        lib.ref_render_data_after_projection.restype = None
        lib.ref_render_data_after_projection.argtypes = [_P, _P, _F, _F, _F, _U, _U, _P]
        lib.ref_load_mesh_and_bvh.restype = _U
        lib.ref_load_mesh_and_bvh.argtypes = [_P, _U, _U, _U, _P, _P, _U, _P, _U, _P, ctypes.POINTER(_U)]
        lib.ref_environment_cross_to_faces.restype = _U
        lib.ref_environment_cross_to_faces.argtypes = [_P, _U, _U, _P]
        _ref_octree = lib
    return _ref_octree


_ref = None


def load_ref():
    global _ref
    if _ref is None:
        path = build_ref()
        if path is None:
            return None
        lib = ctypes.CDLL(path)
        lib.ref_build_bvh.restype = _U
        lib.ref_build_bvh.argtypes = [_P, _U, _P, _P, _U, _U, _P, _U, _P]
        lib.ref_transform_matrices.restype = None
        lib.ref_transform_matrices.argtypes = [_P] * 5
        lib.ref_projection_data.restype = None
        lib.ref_projection_data.argtypes = [_P, _P, _F, _F, _F, _U, _U, _P, _P, _P, _P, _P]
        lib.ref_aabb_area.restype = _F
        lib.ref_aabb_area.argtypes = [_P, _P]
        lib.ref_aabb_intersects.restype = _I
        lib.ref_aabb_intersects.argtypes = [_P] * 4
        lib.ref_struct_sizes.restype = None
        lib.ref_struct_sizes.argtypes = [_P]
        lib.ref_material_default.restype = None
        lib.ref_material_default.argtypes = [_P]
        _ref = lib
    return _ref
