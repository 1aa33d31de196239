"""ctypes binding of include/okrt.h — the only way the Python host side talks to the product.

The numpy dtypes below are byte-identical to the wire structs (and therefore to the reference's
GPU structs, resources/shaders/GPU-Structs.hlsli); arrays of them are passed straight through as
pointers.  The library must exist (gpu-raytracer_b200/libokrt.so, built by build.py): there is no
Python or CPU fallback for any of these calls.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OKRT_LIB") or os.path.join(HERE, "libokrt.so")   # OKRT_LIB: alternative build (tuning experiments)

INVALID_UINT = 0xFFFFFFFF

# ---------------------------------------------------------------- wire dtypes
TRIANGLE = np.dtype([("position", "<f4", (3, 3))])
VERTEX_INFO = np.dtype([("normal", "<f4", 3), ("uv", "<f4", 2), ("tangent", "<f4", 3), ("bitangent", "<f4", 3)])
TRIANGLE_INFO = np.dtype([("vertexInfo", VERTEX_INFO, 3)])
AABB = np.dtype([("min", "<f4", 3), ("max", "<f4", 3)])
BVH_NODE = np.dtype([("boundingBox", AABB), ("triStart", "<u4"), ("triEnd", "<u4"), ("firstChildIdx", "<u4")])
OCTREE_NODE = np.dtype([("boundingBox", AABB), ("meshesStartIdx", "<u4"), ("meshesEndIdx", "<u4"),
                        ("spheresStartIdx", "<u4"), ("spheresEndIdx", "<u4"), ("firstChildIdx", "<u4"), ("numChildren", "<u4")])
MATERIAL = np.dtype([("albedo", "<f4", 3), ("albedoTextureIdx", "<u4"), ("roughness", "<f4"), ("roughnessTextureIdx", "<u4"),
                     ("metallic", "<f4"), ("metallicTextureIdx", "<u4"), ("specular", "<f4"), ("specularTextureIdx", "<u4"),
                     ("normalMapIdx", "<u4"), ("emissionColour", "<f4", 3), ("emissionPower", "<f4"),
                     ("transparency", "<f4"), ("indexOfRefraction", "<f4")])
MESH_INSTANCE = np.dtype([("transformMatrix", "<f4", 16), ("inverseTransformMatrix", "<f4", 16), ("triStart", "<u4"),
                          ("triEnd", "<u4"), ("boundingBox", AABB), ("material", MATERIAL), ("bvhNodeStartIdx", "<u4")])
SPHERE = np.dtype([("position", "<f4", 3), ("material", MATERIAL), ("radius", "<f4")])
DIRECTIONAL_LIGHT = np.dtype([("colour", "<f4", 3), ("intensity", "<f4"), ("effectiveAngle", "<f4"), ("direction", "<f4", 3)])
POINT_LIGHT = np.dtype([("colour", "<f4", 3), ("intensity", "<f4"), ("radius", "<f4"), ("position", "<f4", 3)])
SPOT_LIGHT = np.dtype([("colour", "<f4", 3), ("intensity", "<f4"), ("radius", "<f4"), ("maxAngle", "<f4"),
                       ("position", "<f4", 3), ("direction", "<f4", 3)])
RENDER_DATA = np.dtype([("accumulationEnabled", "<u4"), ("numAccumulationFrames", "<u4"), ("numSpheres", "<u4"),
                        ("numMeshes", "<u4"), ("numDirLights", "<u4"), ("numPointLights", "<u4"), ("numSpotLights", "<u4"),
                        ("debugMode", "<u4"), ("textureDims", "<u4", 2), ("viewPlaneDims", "<f4", 2),
                        ("cameraInverseProjectionMatrix", "<f4", 16), ("cameraInverseViewMatrix", "<f4", 16),
                        ("cameraPosition", "<f4", 3), ("cameraNearZ", "<f4"), ("cameraUpDir", "<f4", 3), ("dofStrength", "<f4"),
                        ("cameraRightDir", "<f4", 3), ("dofDistance", "<f4"), ("debugMaxCount", "<u4"), ("pad0", "<f4", 3)])
MESH_DESC = np.dtype([("startIdx", "<u4"), ("endIdx", "<u4"), ("bvhTreeStartIdx", "<u4"), ("numBvhNodes", "<u4")])
TRANSFORM = np.dtype([("position", "<f4", 3), ("rotation", "<f4", 3), ("scale", "<f4", 3)])
MESH_ENTITY = np.dtype([("transform", TRANSFORM), ("material", MATERIAL), ("meshID", "<u4")])
SPHERE_ENTITY = np.dtype([("transform", TRANSFORM), ("material", MATERIAL), ("radius", "<f4")])
CAMERA = np.dtype([("position", "<f4", 3), ("rotation", "<f4", 3), ("fov", "<f4"), ("nearZ", "<f4"), ("farZ", "<f4")])
OPTIONS = np.dtype([("maxBounces", "<u4"), ("traversal", "<u4"), ("coverAllPixels", "<u4"), ("writeAov", "<u4"),
                    ("shardMode", "<u4"), ("shardRank", "<u4"), ("shardCount", "<u4"), ("shardBandRows", "<u4"),
                    ("frameOffset", "<u4"), ("frameStride", "<u4"), ("pipeline", "<u4"), ("batchesInFlight", "<u4"),
                    ("reserved", "<u4", 4)])
AOV = np.dtype([("hitKind", "<u4"), ("hitIdx", "<u4"), ("triIdx", "<u4"), ("baryU", "<f4"), ("baryV", "<f4"),
                ("distance", "<f4"), ("bbCheckCount", "<u4"), ("triCheckCount", "<u4"), ("numSegments", "<u4"),
                ("flags", "<u4"), ("radiance", "<f4", 3), ("reserved", "<u4")])
STATS = np.dtype([("raySegments", "<u8"), ("fallbackSegments", "<u8"), ("pixelsTraced", "<u8"), ("kernelLaunches", "<u8"),
                  ("lastRenderMs", "<f4"), ("traceKernelMs", "<f4"), ("traceKernelLaunches", "<f4"), ("reserved", "<f4", 1)])

LAYOUT_BUILD = np.dtype([("wallMs", "<f4"), ("deviceMs", "<f4"), ("refineMs", "<f4"), ("builder", "<u4"), ("numPairNodes", "<u4"),
                         ("fastStackDepth", "<u4"), ("rounds", "<u4"), ("why", "S100")])
UNIT_PEAKS = np.dtype([("sectorsPerSecond", "<f8"), ("sectorsPerSecondL1", "<f8"), ("warpInstructionsPerSecond", "<f8"), ("smClockMHz", "<f8"), ("numSms", "<u4"), ("reserved", "<u4")])
ALLREDUCE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p)
LAYOUT_INFO = np.dtype([("numPairNodes", "<u4"), ("numFlattenedLeaves", "<u4"), ("numMergedInstances", "<u4"),
                        ("numInstanceRecords", "<u4"), ("fastStackDepth", "<u4"), ("exactBvhDepth", "<u4"), ("exactOctDepth", "<u4"),
                        ("reserved", "<u4"), ("sahCost", "<f8"), ("buildMs", "<f8")])

WIRE_SIZES = {"TRIANGLE": 36, "VERTEX_INFO": 44, "TRIANGLE_INFO": 132, "AABB": 24, "BVH_NODE": 36, "OCTREE_NODE": 48,
              "MATERIAL": 68, "MESH_INSTANCE": 232, "SPHERE": 84, "DIRECTIONAL_LIGHT": 32, "POINT_LIGHT": 32,
              "SPOT_LIGHT": 48, "RENDER_DATA": 240, "AOV": 56}
for _name, _size in WIRE_SIZES.items():
    assert globals()[_name].itemsize == _size, (_name, globals()[_name].itemsize, _size)


class MeshView(ctypes.Structure):
    _fields_ = [("trianglesPos", ctypes.c_void_p), ("trianglesInfo", ctypes.c_void_p),
                ("numTriangles", ctypes.c_uint32), ("boundingBox", ctypes.c_float * 6)]


TRAVERSAL_FAST, TRAVERSAL_EXACT, TRAVERSAL_WIDE = 0, 1, 2
SHARD_NONE, SHARD_TILE = 0, 1
PIPELINE_WAVEFRONT, PIPELINE_MEGAKERNEL = 0, 1

OK = 0
ERROR_NAMES = {1: "OKRT_ERR_INVALID_ARGUMENT", 2: "OKRT_ERR_CUDA", 3: "OKRT_ERR_STATE", 4: "OKRT_ERR_CAPACITY",
               5: "OKRT_ERR_NO_DEVICE"}

# every symbol include/okrt.h declares: (name, restype, argtypes)
_P, _U, _I = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int
SYMBOLS = [
    ("okrt_version", ctypes.c_char_p, []),
    ("okrt_last_error", ctypes.c_char_p, []),
    ("okrt_device_count", _I, []),
    ("okrt_material_default", None, [_P]),
    ("okrt_options_default", None, [_P]),
    ("okrt_render_data_default", None, [_P]),
    ("okrt_geometry_build", _I, [_P, _U, _U, _U, ctypes.POINTER(_P)]),
    ("okrt_geometry_build_device", _I, [_P, _U, _U, _I, ctypes.POINTER(_P), ctypes.POINTER(ctypes.c_float)]),
    ("okrt_geometry_free", None, [_P]),
    ("okrt_geometry_num_triangles", _U, [_P]),
    ("okrt_geometry_num_nodes", _U, [_P]),
    ("okrt_geometry_num_meshes", _U, [_P]),
    ("okrt_geometry_triangles_pos", _P, [_P]),
    ("okrt_geometry_triangles_info", _P, [_P]),
    ("okrt_geometry_bvh_nodes", _P, [_P]),
    ("okrt_geometry_mesh_descs", _P, [_P]),
    ("okrt_geometry_mesh_bounds", _P, [_P]),
    ("okrt_build_bvh", _I, [_P, _U, _P, _U, _U, _P, _U, ctypes.POINTER(_U), _P]),
    ("okrt_octree_build", _I, [_P, _P, _U, _P, _U, _U, _U, ctypes.POINTER(_P)]),
    ("okrt_octree_free", None, [_P]),
    ("okrt_octree_num_nodes", _U, [_P]),
    ("okrt_octree_num_instances", _U, [_P]),
    ("okrt_octree_nodes", _P, [_P]),
    ("okrt_octree_instances", _P, [_P]),
    ("okrt_transform_matrix", None, [_P, _P]),
    ("okrt_calculate_projection_data", _I, [_P, _U, _U, _P]),
    ("okrt_texture_array_size", _U, [_P, _P, _U]),
    ("okrt_scale_texture", _I, [_P, _U, _U, _P, _U, _U]),
    ("okrt_environment_cross_to_faces", _I, [_P, _U, _U, _P]),
    ("okrt_obj_load", _I, [ctypes.c_char_p, ctypes.c_float, ctypes.POINTER(_P)]),
    ("okrt_obj_load_memory", _I, [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_float, ctypes.POINTER(_P)]),
    ("okrt_obj_free", None, [_P]),
    ("okrt_obj_num_meshes", _U, [_P]),
    ("okrt_obj_mesh", _I, [_P, _U, _P, ctypes.POINTER(ctypes.c_char_p)]),
    ("okrt_obj_texture_path", ctypes.c_char_p, [_P, _U, _U]),
    ("okrt_write_png", _I, [ctypes.c_char_p, _P, _U, _U]),
    ("okrt_write_accumulation_file", _I, [ctypes.c_char_p, _P, _U, _U, _U]),
    ("okrt_read_accumulation_file", _I, [ctypes.c_char_p, _P, ctypes.POINTER(_U), ctypes.POINTER(_U), ctypes.POINTER(_U)]),
    ("okrt_overlay_build", _I, [_U, _P, _U, _U, _U, _P, _P, _U, _U, ctypes.POINTER(_P)]),
    ("okrt_overlay_free", None, [_P]),
    ("okrt_overlay_num_segments", _U, [_P]),
    ("okrt_overlay_num_draws", _U, [_P]),
    ("okrt_overlay_segments", _P, [_P]),
    ("okrt_overlay_keys", _P, [_P]),
    ("okrt_overlay_colours", _P, [_P]),
    ("okrt_layout_report", _I, [_P, _U, _P, _U, _P, _U, _P, _U, _P]),
    ("okrt_create", _I, [_I, _U, _U, ctypes.POINTER(_P)]),
    ("okrt_destroy", None, [_P]),
    ("okrt_resize", _I, [_P, _U, _U]),
    ("okrt_set_triangles", _I, [_P, _P, _P, _U]),
    ("okrt_set_bvh_nodes", _I, [_P, _P, _U]),
    ("okrt_set_textures", _I, [_P, _P, _U, _U]),
    ("okrt_set_environment_map", _I, [_P, _P, _U, _U]),
    ("okrt_set_octree", _I, [_P, _P, _U]),
    ("okrt_set_spheres", _I, [_P, _P, _U]),
    ("okrt_set_mesh_instances", _I, [_P, _P, _U]),
    ("okrt_set_directional_lights", _I, [_P, _P, _U]),
    ("okrt_set_point_lights", _I, [_P, _P, _U]),
    ("okrt_set_spot_lights", _I, [_P, _P, _U]),
    ("okrt_set_render_data", _I, [_P, _P]),
    ("okrt_get_render_data", _I, [_P, _P]),
    ("okrt_set_options", _I, [_P, _P]),
    ("okrt_render", _I, [_P]),
    ("okrt_render_frames", _I, [_P, _U]),
    ("okrt_sync", _I, [_P]),
    ("okrt_toggle_accumulation", _I, [_P, _I]),
    ("okrt_reset_accumulation", _I, [_P]),
    ("okrt_get_num_accumulation_frames", _U, [_P]),
    ("okrt_read_result", _I, [_P, _P]),
    ("okrt_draw_overlay", _I, [_P, _P]),
    ("okrt_draw_overlay_static", _I, [_P, _P]),
    ("okrt_read_result_owned", _I, [_P, _P, ctypes.POINTER(ctypes.c_uint64)]),
    ("okrt_read_accumulation", _I, [_P, _P]),
    ("okrt_read_aov", _I, [_P, _P]),
    ("okrt_write_accumulation", _I, [_P, _P, _U]),
    ("okrt_resolve", _I, [_P, _U]),
    ("okrt_host_register", _I, [_P, ctypes.c_size_t]),
    ("okrt_host_unregister", _I, [_P]),
    ("okrt_accumulation_device_ptr", _P, [_P]),
    ("okrt_result_device_ptr", _P, [_P]),
    ("okrt_cuda_stream", _P, [_P]),
    ("okrt_shard_owner_of_row", _U, [_U, _U, _U]),
    ("okrt_shard_seed_frame", _U, [_P, _U, _U]),
    ("okrt_get_stats", _I, [_P, _P]),
    ("okrt_reset_stats", _I, [_P]),
    ("okrt_get_trace_stats", _I, [_P, _P, _U]),
    ("okrt_read_result_async", _I, [_P, _P, ctypes.POINTER(ctypes.c_uint64), _U]),
    ("okrt_wait_read", _I, [_P, _U]),
    ("okrt_reduce_accumulation", _I, [_P, _P, _P, _U]),
    ("okrt_measure_unit_peaks", _I, [_I, _P]),
    ("okrt_get_layout_build", _I, [_P, _P]),
    ("okrt_wait_layout", _I, [_P]),
    ("okrt_set_shard_bands", _I, [_P, _P, _U]),
    ("okrt_shard_balance", _I, [_P, _U, _U, _P]),
    ("okrt_trace_rays_host", _I, [_P, _U, _P, _U, _P, _U, _P, _U, _U, _P, _U, _P]),
]


class OkrtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("%s: %s" % (ERROR_NAMES.get(code, "error %d" % code), message))
        self.code = code


def load_library(path):
    """ctypes handle of a build of the library at `path` with every symbol typed (lib() below is the product build)."""
    handle = ctypes.CDLL(path)
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(handle, name)
        fn.restype = restype
        fn.argtypes = argtypes
    return handle


_lib = None


def lib():
    """Loads libokrt.so once.  Raises if the library has not been built — nothing substitutes for it."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("gpu-raytracer_b200/libokrt.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(or python gpu-raytracer_b200/build.py); there is no fallback implementation")
        _lib = load_library(LIB_PATH)
    return _lib


def check(code):
    if code != OK:
        raise OkrtError(code, lib().okrt_last_error().decode("utf-8", "replace"))


def ptr(array):
    """Pointer to a C-contiguous numpy array (None -> NULL)."""
    if array is None:
        return None
    assert array.flags["C_CONTIGUOUS"]
    return array.ctypes.data_as(ctypes.c_void_p)


def as_array(pointer, dtype, count):
    """Copies `count` elements of `dtype` from a library-owned pointer into a numpy array."""
    if not count or not pointer:
        return np.zeros(0, dtype=dtype)
    buf = (ctypes.c_char * (dtype.itemsize * count)).from_address(pointer)
    return np.frombuffer(buf, dtype=dtype, count=count).copy()


def default_material(n=None):
    m = np.zeros(1, dtype=MATERIAL)
    lib().okrt_material_default(ptr(m))
    return m[0] if n is None else np.repeat(m, n)


def default_options():
    o = np.zeros(1, dtype=OPTIONS)
    lib().okrt_options_default(ptr(o))
    return o


def default_render_data():
    rd = np.zeros(1, dtype=RENDER_DATA)
    lib().okrt_render_data_default(ptr(rd))
    return rd
