"""Host-side mirror of the reference's `RayTracer` class for the path-tracing hot path.

Same method names, argument meaning and call order as source/Graphics/RayTracer.h:82-116:
    initiate, loadMeshAndBvhData, createOctTree, setScene, render, toggleAccumulation,
    resetAccumulation, getNumAccumulationFrames, onResize, getDOFStrength/Distance (properties here),
    setDebugMode, getDebugMaxCount
`ResourceManager` and `Scene` are the plain-array stand-ins for source/Graphics/ResourceManager.h and
the EnTT registry (source/Scene): just enough to drive the path the way Application.cpp does
(Application.cpp:28-48,65-104).  Every computation happens in libokrt.so through the C ABI.
"""
import ctypes
import os

import numpy as np

from . import _abi as abi


class Mesh:
    """De-indexed triangle soup + bounding box (source/Graphics/Mesh.h:17-58)."""

    def __init__(self, triangles_pos, triangles_info=None, bounding_box=None, name=""):
        self.trianglesPos = np.ascontiguousarray(triangles_pos, dtype=abi.TRIANGLE)
        n = len(self.trianglesPos)
        if triangles_info is None:
            triangles_info = np.zeros(n, dtype=abi.TRIANGLE_INFO)
        self.trianglesInfo = np.ascontiguousarray(triangles_info, dtype=abi.TRIANGLE_INFO)
        assert len(self.trianglesInfo) == n
        if bounding_box is None:
            # Importer::aiMeshToMeshData (Importer.cpp:17-18,52-53): min/max over all positions
            pos = self.trianglesPos["position"].reshape(-1, 3)
            lo = pos.min(axis=0) if n else np.full(3, np.finfo(np.float32).max, np.float32)
            hi = pos.max(axis=0) if n else np.full(3, -np.finfo(np.float32).max, np.float32)
            bounding_box = (lo, hi)
        self.boundingBox = (np.asarray(bounding_box[0], np.float32), np.asarray(bounding_box[1], np.float32))
        self.name = name


class ResourceManager:
    """Mesh and texture stores; asset id = load order (source/Graphics/ResourceManager.h)."""

    def __init__(self):
        self.meshes = []
        self.textures = []        # list of HxWx4 uint8 arrays

    def addMesh(self, mesh):
        self.meshes.append(mesh)
        return len(self.meshes) - 1

    def addTexture(self, rgba8):
        rgba8 = np.ascontiguousarray(rgba8, dtype=np.uint8)
        assert rgba8.ndim == 3 and rgba8.shape[2] == 4
        self.textures.append(rgba8)
        return len(self.textures) - 1


class Scene:
    """Entities in the order the reference's registry views would enumerate them."""

    def __init__(self):
        self.meshEntities = []    # (transform dict, material, meshID)
        self.sphereEntities = []
        self.directionalLights = []
        self.pointLights = []
        self.spotLights = []
        self.camera = dict(position=(0.0, 0.0, 0.0), rotation=(0.0, 0.0, 0.0), fov=90.0, nearZ=0.1, farZ=1000.0)

    @staticmethod
    def _transform(position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1)):
        return dict(position=tuple(position), rotation=tuple(rotation), scale=tuple(scale))

    def addMeshEntity(self, meshID, material=None, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1)):
        material = abi.default_material() if material is None else material
        self.meshEntities.append((self._transform(position, rotation, scale), material, int(meshID)))
        return len(self.meshEntities) - 1

    def addSphereEntity(self, radius=4.0, material=None, position=(0, 0, 0)):
        material = abi.default_material() if material is None else material
        self.sphereEntities.append((self._transform(position), material, float(radius)))
        return len(self.sphereEntities) - 1

    def setCamera(self, position, rotation=(0, 0, 0), fov=90.0, nearZ=0.1, farZ=1000.0):
        self.camera = dict(position=tuple(position), rotation=tuple(rotation), fov=fov, nearZ=nearZ, farZ=farZ)

    # plain-array forms consumed by the C ABI
    def meshEntityArray(self):
        a = np.zeros(len(self.meshEntities), dtype=abi.MESH_ENTITY)
        for i, (t, m, mesh_id) in enumerate(self.meshEntities):
            a[i]["transform"]["position"] = t["position"]
            a[i]["transform"]["rotation"] = t["rotation"]
            a[i]["transform"]["scale"] = t["scale"]
            a[i]["material"] = m
            a[i]["meshID"] = mesh_id
        return a

    def sphereEntityArray(self):
        a = np.zeros(len(self.sphereEntities), dtype=abi.SPHERE_ENTITY)
        for i, (t, m, radius) in enumerate(self.sphereEntities):
            a[i]["transform"]["position"] = t["position"]
            a[i]["transform"]["rotation"] = t["rotation"]
            a[i]["transform"]["scale"] = t["scale"]
            a[i]["material"] = m
            a[i]["radius"] = radius
        return a

    def sphereArray(self):
        """t6 elements: position + Sphere (RayTracer.cpp:88)."""
        a = np.zeros(len(self.sphereEntities), dtype=abi.SPHERE)
        for i, (t, m, radius) in enumerate(self.sphereEntities):
            a[i]["position"] = t["position"]
            a[i]["material"] = m
            a[i]["radius"] = radius
        return a

    def cameraArray(self):
        c = np.zeros(1, dtype=abi.CAMERA)
        c[0]["position"] = self.camera["position"]
        c[0]["rotation"] = self.camera["rotation"]
        c[0]["fov"] = self.camera["fov"]
        c[0]["nearZ"] = self.camera["nearZ"]
        c[0]["farZ"] = self.camera["farZ"]
        return c


class ObjectDescription:
    """Importer::ObjectDecriptionStr (Importer.h): one aiMesh of an imported file + the texture paths of its material."""

    def __init__(self, mesh, name, texturePaths):
        self.mesh, self.name = mesh, name
        (self.albedoTexturePath, self.rougnessTexturePath, self.metallicTexturePath,
         self.specularTexturePath, self.normalTexturePath) = texturePaths


def loadObjects(path=None, scale=1.0, objText=None, mtlText=None):
    """Importer::loadObjects (Importer.cpp:77-120) for Wavefront OBJ, through okrt_obj_load (csrc/okrt_obj.cpp).
    Either `path` (the .mtl files it names are read from its directory) or `objText` (+ optional `mtlText`)."""
    import ctypes
    L = abi.lib()
    h = ctypes.c_void_p()
    if objText is not None:
        ob = objText.encode() if isinstance(objText, str) else objText
        mb = (mtlText.encode() if isinstance(mtlText, str) else mtlText) if mtlText else b""
        abi.check(L.okrt_obj_load_memory(ob, len(ob), mb if mb else None, len(mb), scale, ctypes.byref(h)))
    else:
        abi.check(L.okrt_obj_load(str(path).encode(), scale, ctypes.byref(h)))
    try:
        out = []
        for i in range(L.okrt_obj_num_meshes(h)):
            view = abi.MeshView()
            name = ctypes.c_char_p()
            abi.check(L.okrt_obj_mesh(h, i, ctypes.cast(ctypes.pointer(view), ctypes.c_void_p), ctypes.byref(name)))
            n = int(view.numTriangles)
            pos = abi.as_array(view.trianglesPos, abi.TRIANGLE, n)
            info = abi.as_array(view.trianglesInfo, abi.TRIANGLE_INFO, n)
            box = (np.array([view.boundingBox[k] for k in range(3)], np.float32), np.array([view.boundingBox[3 + k] for k in range(3)], np.float32))
            nm = (name.value or b"").decode("utf-8", "replace")
            tex = tuple(L.okrt_obj_texture_path(h, i, s).decode("utf-8", "replace") for s in range(5))
            out.append(ObjectDescription(Mesh(pos, info, box, nm), nm, tex))
        return out
    finally:
        L.okrt_obj_free(h)


class Geometry:
    """Result of loadMeshAndBvhData: the three GPU arrays + mesh descriptors (RayTracer.cpp:120-217)."""

    def __init__(self, meshes, maxDepth, maxLeafTriangles, device=None):
        """device=None: the reference's SAH tree, bit for bit (CPU).  device=k: LBVH built on cuda:k (row f2) — the
        reference's node format with a different tree; self.buildMs is the device-timed duration."""
        L = abi.lib()
        views = (abi.MeshView * max(len(meshes), 1))()
        self._keep = meshes
        for i, m in enumerate(meshes):
            views[i].trianglesPos = m.trianglesPos.ctypes.data
            views[i].trianglesInfo = m.trianglesInfo.ctypes.data
            views[i].numTriangles = len(m.trianglesPos)
            for k in range(3):
                views[i].boundingBox[k] = float(m.boundingBox[0][k])
                views[i].boundingBox[3 + k] = float(m.boundingBox[1][k])
        handle = ctypes.c_void_p()
        self.buildMs = None
        if device is None:
            abi.check(L.okrt_geometry_build(ctypes.cast(views, ctypes.c_void_p), len(meshes), maxDepth, maxLeafTriangles,
                                            ctypes.byref(handle)))
        else:
            ms = ctypes.c_float(0.0)
            abi.check(L.okrt_geometry_build_device(ctypes.cast(views, ctypes.c_void_p), len(meshes), maxLeafTriangles, int(device),
                                                   ctypes.byref(handle), ctypes.byref(ms)))
            self.buildMs = float(ms.value)
        self.handle = handle
        nt, nn, nm = L.okrt_geometry_num_triangles(handle), L.okrt_geometry_num_nodes(handle), L.okrt_geometry_num_meshes(handle)
        self.trianglesPos = abi.as_array(L.okrt_geometry_triangles_pos(handle), abi.TRIANGLE, nt)
        self.trianglesInfo = abi.as_array(L.okrt_geometry_triangles_info(handle), abi.TRIANGLE_INFO, nt)
        self.bvhNodes = abi.as_array(L.okrt_geometry_bvh_nodes(handle), abi.BVH_NODE, nn)
        self.meshDescs = abi.as_array(L.okrt_geometry_mesh_descs(handle), abi.MESH_DESC, nm)
        self.meshBounds = abi.as_array(L.okrt_geometry_mesh_bounds(handle), abi.AABB, nm)

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                abi.lib().okrt_geometry_free(self.handle)
                self.handle = None
        except Exception:
            pass


class OctTree:
    """Result of createOctTree/loadOctTree: GPU_OctTreeNode[] + GPU_MeshComponent[] (RayTracer.cpp:523-788)."""

    def __init__(self, geometry, scene, maxDepth, maxLeafObjects):
        L = abi.lib()
        me, se = scene.meshEntityArray(), scene.sphereEntityArray()
        handle = ctypes.c_void_p()
        abi.check(L.okrt_octree_build(geometry.handle, abi.ptr(me), len(me), abi.ptr(se), len(se), maxDepth, maxLeafObjects,
                                      ctypes.byref(handle)))
        self.nodes = abi.as_array(L.okrt_octree_nodes(handle), abi.OCTREE_NODE, L.okrt_octree_num_nodes(handle))
        self.instances = abi.as_array(L.okrt_octree_instances(handle), abi.MESH_INSTANCE, L.okrt_octree_num_instances(handle))
        L.okrt_octree_free(handle)


def scaleTexturesToArray(textures):
    """RayTracer::loadTextureData (RayTracer.cpp:310-342): one N x N x L RGBA8 array."""
    L = abi.lib()
    if not textures:
        return np.zeros((0, 0, 0, 4), np.uint8)
    widths = np.array([t.shape[1] for t in textures], np.uint32)
    heights = np.array([t.shape[0] for t in textures], np.uint32)
    n = L.okrt_texture_array_size(abi.ptr(widths), abi.ptr(heights), len(textures))
    out = np.zeros((len(textures), n, n, 4), np.uint8)
    for i, t in enumerate(textures):
        t = np.ascontiguousarray(t)
        abi.check(L.okrt_scale_texture(abi.ptr(t), t.shape[1], t.shape[0], abi.ptr(out[i]), n, n))
    return out


def environmentCrossToFaces(cross_rgba8):
    """RayTracer::loadEnvironmentMap (RayTracer.cpp:344-432): 4x3 cross image -> 6 faces."""
    L = abi.lib()
    cross = np.ascontiguousarray(cross_rgba8, np.uint8)
    h, w = cross.shape[:2]
    faces = np.zeros((6, h // 3, w // 4, 4), np.uint8)
    abi.check(L.okrt_environment_cross_to_faces(abi.ptr(cross), w, h, abi.ptr(faces)))
    return faces


class RayTracer:
    DebugDisplayMode_None, DebugDisplayMode_BBCheckCount, DebugDisplayMode_TriCheckCount = 0, 1, 2

    def __init__(self, device=0):
        self._device = device
        self._handle = None
        self._scene = None
        self._resources = None
        self.geometry = None
        self.octTree = None
        self._rd = None
        self._opt = None
        self._sphereBinding = False

    # ------------------------------------------------------------------ lifetime
    def initiate(self, width, height, resourceManager, environmentMapFaces=None, maxBounces=1):
        """RayTracer::initiate (RayTracer.cpp:60-118).  `environmentMapFaces`: 6 x H x W x 4 uint8 or None."""
        L = abi.lib()
        self.shutdown()
        handle = ctypes.c_void_p()
        abi.check(L.okrt_create(self._device, width, height, ctypes.byref(handle)))
        self._handle = handle
        self._resources = resourceManager
        self._rd = abi.default_render_data()
        self._rd[0]["textureDims"] = (width, height)
        self._opt = abi.default_options()
        self._opt[0]["maxBounces"] = maxBounces
        abi.check(L.okrt_set_options(self._handle, abi.ptr(self._opt)))
        self.loadTextureData()
        if environmentMapFaces is not None:
            faces = np.ascontiguousarray(environmentMapFaces, np.uint8)
            abi.check(L.okrt_set_environment_map(self._handle, abi.ptr(faces), faces.shape[2], faces.shape[1]))
        self.loadMeshAndBvhData(30, 5)                       # RayTracer.cpp:97

    def shutdown(self):
        if self._handle:
            abi.lib().okrt_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self.shutdown()
        except Exception:
            pass

    # --------------------------------------------------------------- scene feed
    def loadTextureData(self):
        arr = scaleTexturesToArray(self._resources.textures)
        if arr.shape[0]:
            abi.check(abi.lib().okrt_set_textures(self._handle, abi.ptr(arr), arr.shape[1], arr.shape[0]))
        self.textureArray = arr

    def loadMeshAndBvhData(self, maxDepth, maxLeafTriangles):
        """RayTracer::loadMeshAndBvhData (RayTracer.cpp:120-217)."""
        L = abi.lib()
        if not self._resources.meshes:
            return
        self.geometry = Geometry(self._resources.meshes, maxDepth, maxLeafTriangles)
        g = self.geometry
        abi.check(L.okrt_set_triangles(self._handle, abi.ptr(g.trianglesPos), abi.ptr(g.trianglesInfo), len(g.trianglesPos)))
        abi.check(L.okrt_set_bvh_nodes(self._handle, abi.ptr(g.bvhNodes), len(g.bvhNodes)))

    def setScene(self, scene):
        self._scene = scene

    def bindSpheres(self, enable=True):
        """At HEAD the reference never uploads spheres and forces numSpheres = 0 (RayTracer.cpp:780-781);
        the shader contract (t6) still has them.  enable=True feeds them."""
        self._sphereBinding = bool(enable)

    def createOctTree(self, scene, maxDepth, maxLeafObjects):
        """RayTracer::createOctTree (RayTracer.cpp:523-704) + loadOctTree (:706-788)."""
        L = abi.lib()
        self._scene = scene
        self.octTree = OctTree(self.geometry, scene, maxDepth, maxLeafObjects)
        t = self.octTree
        abi.check(L.okrt_set_mesh_instances(self._handle, abi.ptr(t.instances), len(t.instances)))
        abi.check(L.okrt_set_octree(self._handle, abi.ptr(t.nodes), len(t.nodes)))
        spheres = scene.sphereArray() if self._sphereBinding else np.zeros(0, abi.SPHERE)
        abi.check(L.okrt_set_spheres(self._handle, abi.ptr(spheres) if len(spheres) else None, len(spheres)))
        self._rd[0]["numMeshes"] = 0                          # RayTracer.cpp:780
        self._rd[0]["numSpheres"] = len(spheres)              # 0 at HEAD (:781)

    # ------------------------------------------------------------------- frame
    def _updateBuffers(self):
        """calculateProjectionData + the light part of updateBuffers (RayTracer.cpp:268-308,434-488)."""
        L = abi.lib()
        s = self._scene
        # keep the library's frame counter, refresh everything else
        cur = np.zeros(1, abi.RENDER_DATA)
        abi.check(L.okrt_get_render_data(self._handle, abi.ptr(cur)))
        self._rd[0]["numAccumulationFrames"] = cur[0]["numAccumulationFrames"]
        self._rd[0]["accumulationEnabled"] = cur[0]["accumulationEnabled"]
        cam = s.cameraArray()
        w, h = int(self._rd[0]["textureDims"][0]), int(self._rd[0]["textureDims"][1])
        abi.check(L.okrt_calculate_projection_data(abi.ptr(cam), w, h, abi.ptr(self._rd)))
        for name, items, dtype, setter in (("numDirLights", s.directionalLights, abi.DIRECTIONAL_LIGHT, L.okrt_set_directional_lights),
                                           ("numPointLights", s.pointLights, abi.POINT_LIGHT, L.okrt_set_point_lights),
                                           ("numSpotLights", s.spotLights, abi.SPOT_LIGHT, L.okrt_set_spot_lights)):
            arr = np.array(items, dtype=dtype) if items else np.zeros(0, dtype)
            abi.check(setter(self._handle, abi.ptr(arr) if len(arr) else None, len(arr)))
            self._rd[0][name] = len(arr)
        abi.check(L.okrt_set_render_data(self._handle, abi.ptr(self._rd)))

    def render(self, frames=1):
        """RayTracer::render (RayTracer.cpp:219-260), `frames` times."""
        self._updateBuffers()
        abi.check(abi.lib().okrt_render_frames(self._handle, frames))

    def sync(self):
        abi.check(abi.lib().okrt_sync(self._handle))

    def toggleAccumulation(self, enable):
        abi.check(abi.lib().okrt_toggle_accumulation(self._handle, 1 if enable else 0))

    def resetAccumulation(self):
        abi.check(abi.lib().okrt_reset_accumulation(self._handle))

    def getNumAccumulationFrames(self):
        return abi.lib().okrt_get_num_accumulation_frames(self._handle)

    def onResize(self, width, height):
        abi.check(abi.lib().okrt_resize(self._handle, width, height))
        self._rd[0]["textureDims"] = (width, height)

    def setDebugMode(self, mode):
        self._rd[0]["debugMode"] = mode

    @property
    def debugMaxCount(self):
        return int(self._rd[0]["debugMaxCount"])

    @debugMaxCount.setter
    def debugMaxCount(self, v):
        self._rd[0]["debugMaxCount"] = v

    @property
    def dofStrength(self):
        return float(self._rd[0]["dofStrength"])

    @dofStrength.setter
    def dofStrength(self, v):
        self._rd[0]["dofStrength"] = v

    @property
    def dofDistance(self):
        return float(self._rd[0]["dofDistance"])

    @dofDistance.setter
    def dofDistance(self, v):
        self._rd[0]["dofDistance"] = v

    # ------------------------------------------------- new (not in the reference)
    def setOptions(self, **kw):
        for k, v in kw.items():
            self._opt[0][k] = v
        abi.check(abi.lib().okrt_set_options(self._handle, abi.ptr(self._opt)))

    def renderData(self):
        self._updateBuffers()
        return self._rd.copy()

    def readResult(self):
        w, h = int(self._rd[0]["textureDims"][0]), int(self._rd[0]["textureDims"][1])
        out = np.zeros((h, w, 4), np.uint8)
        abi.check(abi.lib().okrt_read_result(self._handle, abi.ptr(out)))
        return out

    def readAccumulation(self):
        w, h = int(self._rd[0]["textureDims"][0]), int(self._rd[0]["textureDims"][1])
        out = np.zeros((h, w, 4), np.float32)
        abi.check(abi.lib().okrt_read_accumulation(self._handle, abi.ptr(out)))
        return out

    # ---- DebugRenderer's box overlays (row f4)
    DrawMode_None, DrawMode_DrawSingle, DrawMode_DrawWithChildren, DrawMode_DrawWithDecendants = 0, 1, 2, 3

    def getGlobalNodeIdx(self, meshEntityIdx, localNodeIdx):
        """RayTracer::getGlobalNodeIdx (RayTracer.h:226-230)."""
        mesh_id = self._scene.meshEntities[meshEntityIdx][2]
        return int(self.geometry.meshDescs[mesh_id]["bvhTreeStartIdx"]) + int(localNodeIdx)

    def _drawOverlay(self, kind, nodes, nodeIdx, drawMode, transform):
        L = abi.lib()
        cam = self._scene.cameraArray()
        w, h = int(self._rd[0]["textureDims"][0]), int(self._rd[0]["textureDims"][1])
        handle = ctypes.c_void_p()
        abi.check(L.okrt_overlay_build(kind, abi.ptr(nodes), len(nodes), int(nodeIdx), int(drawMode),
                                       abi.ptr(transform) if transform is not None else None, abi.ptr(cam), w, h, ctypes.byref(handle)))
        try:
            abi.check(L.okrt_draw_overlay(self._handle, handle))
        finally:
            L.okrt_overlay_free(handle)

    def renderBvhNodeBBs(self, meshEntityIdx, localNodeIdx, drawMode):
        """DebugRenderer::renderBvhNodeBBs (DebugRenderer.cpp:264-303) over the result image."""
        tr = np.zeros(1, abi.TRANSFORM)
        t = self._scene.meshEntities[meshEntityIdx][0]
        tr[0]["position"], tr[0]["rotation"], tr[0]["scale"] = t["position"], t["rotation"], t["scale"]
        self._drawOverlay(0, self.geometry.bvhNodes, self.getGlobalNodeIdx(meshEntityIdx, localNodeIdx), drawMode, tr)

    def renderOctTreeNodeBBs(self, nodeIdx, drawMode):
        """DebugRenderer::renderOctTreeNodeBBs (DebugRenderer.cpp:335-364)."""
        self._drawOverlay(1, self.octTree.nodes, nodeIdx, drawMode, None)

    def saveScreenshot(self, path="render.png"):
        """Application::saveScreenshot (Application.cpp:661-674): the RGBA8 target as a PNG ("render.png" there)."""
        img = self.readResult()
        abi.check(abi.lib().okrt_write_png(os.fsencode(path), abi.ptr(img), img.shape[1], img.shape[0]))

    def saveAccumulation(self, path):
        """Progressive state to disk (row f3): RGBA32F accumulation + frame count; loadAccumulation resumes from it."""
        acc = self.readAccumulation()
        abi.check(abi.lib().okrt_write_accumulation_file(os.fsencode(path), abi.ptr(acc), acc.shape[1], acc.shape[0],
                                                         self.getNumAccumulationFrames()))

    def loadAccumulation(self, path):
        w, h, n = ctypes.c_uint32(0), ctypes.c_uint32(0), ctypes.c_uint32(0)
        L = abi.lib()
        abi.check(L.okrt_read_accumulation_file(os.fsencode(path), None, ctypes.byref(w), ctypes.byref(h), ctypes.byref(n)))
        if (w.value, h.value) != (int(self._rd[0]["textureDims"][0]), int(self._rd[0]["textureDims"][1])):
            raise ValueError("accumulation file is %dx%d, the renderer %dx%d" % (w.value, h.value, *self._rd[0]["textureDims"]))
        acc = np.zeros((h.value, w.value, 4), np.float32)
        abi.check(L.okrt_read_accumulation_file(os.fsencode(path), abi.ptr(acc), ctypes.byref(w), ctypes.byref(h), ctypes.byref(n)))
        abi.check(L.okrt_write_accumulation(self._handle, abi.ptr(acc), n.value))

    def readAov(self):
        w, h = int(self._rd[0]["textureDims"][0]), int(self._rd[0]["textureDims"][1])
        out = np.zeros((h, w), abi.AOV)
        abi.check(abi.lib().okrt_read_aov(self._handle, abi.ptr(out)))
        return out

    def writeAccumulation(self, rgba32f, frames):
        a = np.ascontiguousarray(rgba32f, np.float32)
        abi.check(abi.lib().okrt_write_accumulation(self._handle, abi.ptr(a), frames))

    def resolve(self, totalFrames):
        abi.check(abi.lib().okrt_resolve(self._handle, totalFrames))

    def stats(self):
        s = np.zeros(1, abi.STATS)
        abi.check(abi.lib().okrt_get_stats(self._handle, abi.ptr(s)))
        return {k: (int(s[0][k]) if k != "lastRenderMs" else float(s[0][k])) for k in ("raySegments", "fallbackSegments", "pixelsTraced", "kernelLaunches", "lastRenderMs")}

    def resetStats(self):
        abi.check(abi.lib().okrt_reset_stats(self._handle))

    def accumulationDevicePtr(self):
        return abi.lib().okrt_accumulation_device_ptr(self._handle)

    def resultDevicePtr(self):
        return abi.lib().okrt_result_device_ptr(self._handle)

    @property
    def handle(self):
        return self._handle
