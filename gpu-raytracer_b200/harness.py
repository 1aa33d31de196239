"""Low-level harness over the C ABI: the binding-table setters exactly as include/okrt.h declares them.

`build_host_side` runs the host builders (okrt_geometry_build, okrt_octree_build, okrt_calculate_projection_data,
okrt_scale_texture) and returns every array of the reference's binding table (t0..t7, b0); `Renderer` uploads them
and drives okrt_render_frames / okrt_read_*.  bench.py, tools/ and the tests use it; nothing here touches oracle/."""
import ctypes

import numpy as np



def build_host_side(okrt, rm, sc, info, width, height, spheres=False, bvhDevice=None):
    """Runs the product's host builders (no GPU needed) and returns every array of the binding table."""
    abi = okrt.abi
    L = abi.lib()
    geo = okrt.Geometry(rm.meshes, *info.get("bvh", (30, 5)), device=bvhDevice)
    oct_ = okrt.OctTree(geo, sc, *info.get("octree", (2, 50)))
    rd = abi.default_render_data()
    rd[0]["textureDims"] = (width, height)
    cam = sc.cameraArray()
    abi.check(L.okrt_calculate_projection_data(abi.ptr(cam), width, height, abi.ptr(rd)))
    sph = sc.sphereArray() if spheres else np.zeros(0, abi.SPHERE)
    rd[0]["numSpheres"] = len(sph)
    tex = okrt.scaleTexturesToArray(rm.textures)
    return dict(geometry=geo, octree=oct_, rd=rd, spheres=sph, textures=tex)


def scene_digest(host, env=None):
    """sha256 over every array of the binding table: numpy's transcendental kernels differ in the last bit between
    CPU generations, so a golden render is only comparable where the generated scene is bit-identical."""
    import hashlib
    h = hashlib.sha256()
    g, t = host["geometry"], host["octree"]
    for a in (g.trianglesPos, g.trianglesInfo, g.bvhNodes, t.nodes, t.instances, host["spheres"], host["textures"], host["rd"]):
        h.update(np.ascontiguousarray(a).tobytes())
    if env is not None:
        h.update(np.ascontiguousarray(env).tobytes())
    return h.hexdigest()


DEFAULT_PIPELINE = 0      # 0 wavefront, 1 persistent megakernel (the tests parametrize this)


class Renderer:
    """Low-level use of the C ABI (binding-table setters), as the parity tests drive it."""

    def __init__(self, okrt, host, width, height, envFaces=None, device=0, lib=None):
        """lib: a ctypes handle from abi.load_library (e.g. the -DOKRT_TRACE_STATS build); default: the product library."""
        abi = okrt.abi
        self.abi, self.L = abi, (lib if lib is not None else abi.lib())
        self.W, self.H = width, height
        h = ctypes.c_void_p()
        abi.check(self.L.okrt_create(device, width, height, ctypes.byref(h)))
        self.h = h
        g, t = host["geometry"], host["octree"]
        abi.check(self.L.okrt_set_triangles(h, abi.ptr(g.trianglesPos), abi.ptr(g.trianglesInfo), len(g.trianglesPos)))
        abi.check(self.L.okrt_set_bvh_nodes(h, abi.ptr(g.bvhNodes), len(g.bvhNodes)))
        abi.check(self.L.okrt_set_mesh_instances(h, abi.ptr(t.instances) if len(t.instances) else None, len(t.instances)))
        abi.check(self.L.okrt_set_octree(h, abi.ptr(t.nodes), len(t.nodes)))
        if len(host["spheres"]):
            abi.check(self.L.okrt_set_spheres(h, abi.ptr(host["spheres"]), len(host["spheres"])))
        if host["textures"].size:
            tex = host["textures"]
            abi.check(self.L.okrt_set_textures(h, abi.ptr(tex), tex.shape[1], tex.shape[0]))
        if envFaces is not None:
            f = np.ascontiguousarray(envFaces)
            abi.check(self.L.okrt_set_environment_map(h, abi.ptr(f), f.shape[2], f.shape[1]))
        self.opt = abi.default_options()
        self.opt[0]["pipeline"] = DEFAULT_PIPELINE

    def set_lights(self, dirLights=None, pointLights=None, spotLights=None):
        abi = self.abi
        for arr, fn in ((dirLights, self.L.okrt_set_directional_lights), (pointLights, self.L.okrt_set_point_lights), (spotLights, self.L.okrt_set_spot_lights)):
            if arr is not None and len(arr):
                abi.check(fn(self.h, abi.ptr(arr), len(arr)))

    def options(self, **kw):
        for k, v in kw.items():
            self.opt[0][k] = v
        self.abi.check(self.L.okrt_set_options(self.h, self.abi.ptr(self.opt)))

    def render(self, rd, frames=1):
        self.abi.check(self.L.okrt_set_render_data(self.h, self.abi.ptr(rd)))
        self.abi.check(self.L.okrt_render_frames(self.h, frames))
        self.abi.check(self.L.okrt_sync(self.h))

    def wait_layout(self):
        """Adopts the background-refined traversal layout (okrt_wait_layout); returns the okrt_layout_build record."""
        self.abi.check(self.L.okrt_wait_layout(self.h))
        b = np.zeros(1, self.abi.LAYOUT_BUILD)
        self.abi.check(self.L.okrt_get_layout_build(self.h, self.abi.ptr(b)))
        return b[0]

    def result(self):
        out = np.zeros((self.H, self.W, 4), np.uint8)
        self.abi.check(self.L.okrt_read_result(self.h, self.abi.ptr(out)))
        return out

    def accumulation(self):
        out = np.zeros((self.H, self.W, 4), np.float32)
        self.abi.check(self.L.okrt_read_accumulation(self.h, self.abi.ptr(out)))
        return out

    def aov(self):
        out = np.zeros((self.H, self.W), self.abi.AOV)
        self.abi.check(self.L.okrt_read_aov(self.h, self.abi.ptr(out)))
        return out

    def stats(self):
        s = np.zeros(1, self.abi.STATS)
        self.abi.check(self.L.okrt_get_stats(self.h, self.abi.ptr(s)))
        return s[0]

    def trace_stats(self, bounces):
        """[bounces, 8] step counters of the traversal kernel since okrt_reset_stats (a -DOKRT_TRACE_STATS library only)."""
        out = np.zeros((bounces, 8), np.uint64)
        rc = self.L.okrt_get_trace_stats(self.h, self.abi.ptr(out), bounces)
        if rc != self.abi.OK:
            raise self.abi.OkrtError(rc, self.L.okrt_last_error().decode("utf-8", "replace"))
        return out

    def close(self):
        if self.h:
            self.L.okrt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
