"""Procedural scenes for the BASELINE.json configurations.

The reference ships without its large assets (sponza.obj, dragon.obj, the HDR sky — see
/root/reference/.MISSING_LARGE_BLOBS) and the GPU box has no /root/reference at all, so every
benchmark scene here is a seeded, documented STAND-IN of the same triangle / instance / texture
budget, generated with numpy.  All builders return (ResourceManager, Scene, info) ready for
RayTracer.initiate / createOctTree; the same arrays feed the product and the CPU oracle.
"""
import numpy as np

from . import _abi as abi
from .raytracer import Mesh, ResourceManager, Scene


# ------------------------------------------------------------------ geometry helpers
def _tangent_frames(pos, uv):
    """Per-triangle tangent/bitangent from UV derivatives (what aiProcess_CalcTangentSpace provides)."""
    e1, e2 = pos[:, 1] - pos[:, 0], pos[:, 2] - pos[:, 0]
    d1, d2 = uv[:, 1] - uv[:, 0], uv[:, 2] - uv[:, 0]
    det = d1[:, 0] * d2[:, 1] - d2[:, 0] * d1[:, 1]
    det = np.where(np.abs(det) < 1e-12, 1.0, det)
    r = (1.0 / det)[:, None]
    t = (e1 * d2[:, 1:2] - e2 * d1[:, 1:2]) * r
    b = (e2 * d1[:, 0:1] - e1 * d2[:, 0:1]) * r

    def unit(v):
        n = np.linalg.norm(v, axis=1, keepdims=True)
        return np.where(n > 1e-20, v / np.maximum(n, 1e-20), np.array([[1.0, 0.0, 0.0]]))
    return unit(t), unit(b)


def make_mesh(pos, normals=None, uvs=None, name=""):
    """pos [n,3,3], normals [n,3,3] (default: face normals), uvs [n,3,2] (default 0) -> Mesh."""
    pos = np.asarray(pos, np.float64).reshape(-1, 3, 3)
    n = len(pos)
    if normals is None:
        fn = np.cross(pos[:, 1] - pos[:, 0], pos[:, 2] - pos[:, 0])
        ln = np.linalg.norm(fn, axis=1, keepdims=True)
        fn = np.where(ln > 1e-20, fn / np.maximum(ln, 1e-20), np.array([[0.0, 1.0, 0.0]]))
        normals = np.repeat(fn[:, None, :], 3, axis=1)
    if uvs is None:
        uvs = np.zeros((n, 3, 2))
    uvs = np.asarray(uvs, np.float64).reshape(n, 3, 2)
    t, b = _tangent_frames(pos, uvs)
    tp = np.zeros(n, abi.TRIANGLE)
    tp["position"] = pos.astype(np.float32)
    ti = np.zeros(n, abi.TRIANGLE_INFO)
    ti["vertexInfo"]["normal"] = np.asarray(normals, np.float32).reshape(n, 3, 3)
    ti["vertexInfo"]["uv"] = uvs.astype(np.float32)
    ti["vertexInfo"]["tangent"] = np.repeat(t[:, None, :], 3, axis=1).astype(np.float32)
    ti["vertexInfo"]["bitangent"] = np.repeat(b[:, None, :], 3, axis=1).astype(np.float32)
    return Mesh(tp, ti, name=name)


def grid_surface(fn, nu, nv, uv_scale=(1.0, 1.0), flip=False):
    """Triangulates p = fn(u, v), u,v in [0,1], on an nu x nv grid.  Normals by central differences."""
    u = np.linspace(0.0, 1.0, nu + 1)
    v = np.linspace(0.0, 1.0, nv + 1)
    U, V = np.meshgrid(u, v, indexing="ij")
    P = fn(U, V)                                            # [nu+1, nv+1, 3]
    h = 1e-4
    du = (fn(np.clip(U + h, 0, 1), V) - fn(np.clip(U - h, 0, 1), V))
    dv = (fn(U, np.clip(V + h, 0, 1)) - fn(U, np.clip(V - h, 0, 1)))
    N = np.cross(du, dv)
    ln = np.linalg.norm(N, axis=2, keepdims=True)
    N = np.where(ln > 1e-20, N / np.maximum(ln, 1e-20), np.array([0.0, 1.0, 0.0]))
    if flip:
        N = -N
    T = np.stack([U * uv_scale[0], V * uv_scale[1]], axis=2)
    i0, j0 = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    i0, j0 = i0.ravel(), j0.ravel()

    def corner(a, di, dj):
        return a[i0 + di, j0 + dj]
    order = ((0, 0), (1, 0), (1, 1)), ((0, 0), (1, 1), (0, 1))
    if flip:
        order = ((0, 0), (1, 1), (1, 0)), ((0, 0), (0, 1), (1, 1))
    pos, nor, uvs = [], [], []
    for tri in order:
        pos.append(np.stack([corner(P, *c) for c in tri], axis=1))
        nor.append(np.stack([corner(N, *c) for c in tri], axis=1))
        uvs.append(np.stack([corner(T, *c) for c in tri], axis=1))
    return np.concatenate(pos), np.concatenate(nor), np.concatenate(uvs)


def quad(p0, p1, p2, p3, uv_scale=1.0):
    """Two triangles p0 p1 p2 / p0 p2 p3."""
    p = np.array([[p0, p1, p2], [p0, p2, p3]], np.float64)
    uv = np.array([[[0, 0], [1, 0], [1, 1]], [[0, 0], [1, 1], [0, 1]]], np.float64) * uv_scale
    return p, None, uv


def box(lo, hi):
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    c = np.array([[lo[0], lo[1], lo[2]], [hi[0], lo[1], lo[2]], [hi[0], hi[1], lo[2]], [lo[0], hi[1], lo[2]],
                  [lo[0], lo[1], hi[2]], [hi[0], lo[1], hi[2]], [hi[0], hi[1], hi[2]], [lo[0], hi[1], hi[2]]])
    faces = [(0, 3, 2, 1), (4, 5, 6, 7), (0, 1, 5, 4), (2, 3, 7, 6), (1, 2, 6, 5), (0, 4, 7, 3)]
    ps, us = [], []
    for f in faces:
        p, _, uv = quad(*[c[i] for i in f])
        ps.append(p)
        us.append(uv)
    return np.concatenate(ps), None, np.concatenate(us)


def icosphere(subdivisions=3, radius=1.0, centre=(0, 0, 0)):
    """20 * 4^s triangles (s=3 -> 1280)."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], float)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                  [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]])
    tri = v[f]
    tri /= np.linalg.norm(tri, axis=2, keepdims=True)
    for _ in range(subdivisions):
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = a + b, b + c, c + a
        ab /= np.linalg.norm(ab, axis=1, keepdims=True)
        bc /= np.linalg.norm(bc, axis=1, keepdims=True)
        ca /= np.linalg.norm(ca, axis=1, keepdims=True)
        tri = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    nor = tri.copy()
    uv = np.stack([np.arctan2(nor[..., 0], nor[..., 2]) / (2 * np.pi) + 0.5, nor[..., 1] * 0.5 + 0.5], axis=2)
    return tri * radius + np.asarray(centre, float), nor, uv


# ------------------------------------------------------------------ textures / sky
def procedural_texture(kind, size, seed):
    """size x size x 4 uint8.  kinds: bricks, stone, fabric, checker, noise, spec."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:size, 0:size].astype(np.float32) / size
    base = rng.uniform(0.25, 0.9, 3).astype(np.float32)
    cells = 32 if size % 32 == 0 and size >= 32 else 1
    lowres = rng.uniform(0.0, 1.0, (32, 32)).astype(np.float32)[:cells, :cells]
    noise = np.kron(lowres, np.ones((size // cells, size // cells), np.float32))
    if kind == "bricks":
        row = np.floor(y * 16)
        bx = (x * 8 + 0.5 * (row % 2)) % 1.0
        by = (y * 16) % 1.0
        mortar = ((bx < 0.06) | (by < 0.1)).astype(np.float32)
        shade = 0.75 + 0.25 * noise
        img = base[None, None, :] * shade[..., None] * (1 - mortar[..., None]) + 0.55 * mortar[..., None]
    elif kind == "fabric":
        weave = 0.5 + 0.25 * np.sin(x * 400.0) * np.sin(y * 400.0)
        stripes = 0.8 + 0.2 * np.sign(np.sin(x * 40.0))
        img = base[None, None, :] * (weave * stripes)[..., None]
    elif kind == "checker":
        c = ((np.floor(x * 8) + np.floor(y * 8)) % 2).astype(np.float32)
        img = base[None, None, :] * (0.35 + 0.65 * c)[..., None]
    elif kind == "spec":
        g = 0.3 + 0.5 * noise
        img = np.repeat(g[..., None], 3, axis=2)
    else:  # stone / noise
        veins = 0.5 + 0.5 * np.sin((x + 0.3 * noise) * 30.0 + (y * 12.0))
        img = base[None, None, :] * (0.6 + 0.3 * noise + 0.1 * veins)[..., None]
    out = np.empty((size, size, 4), np.uint8)
    out[..., :3] = np.clip(img * 255.0 + 0.5, 0, 255).astype(np.uint8)
    out[..., 3] = 255
    return out


def procedural_sky_faces(size=512, sun_dir=(0.35, 0.85, 0.4), seed=7):
    """6 x size x size x 4 uint8 cube faces (+X,-X,+Y,-Y,+Z,-Z), D3D face conventions [Q13]."""
    sun = np.asarray(sun_dir, np.float64)
    sun /= np.linalg.norm(sun)
    t, s = np.mgrid[0:size, 0:size]
    sc = (s + 0.5) / size * 2.0 - 1.0
    tc = (t + 0.5) / size * 2.0 - 1.0
    one = np.ones_like(sc)
    dirs = [np.stack([one, -tc, -sc], -1), np.stack([-one, -tc, sc], -1), np.stack([sc, one, tc], -1),
            np.stack([sc, -one, -tc], -1), np.stack([sc, -tc, one], -1), np.stack([-sc, -tc, -one], -1)]
    faces = np.zeros((6, size, size, 4), np.uint8)
    for f, d in enumerate(dirs):
        d = d / np.linalg.norm(d, axis=-1, keepdims=True)
        up = d[..., 1]
        sky = np.array([0.40, 0.62, 0.95])[None, None, :] * (0.55 + 0.45 * np.clip(up, 0, 1))[..., None]
        horizon = np.array([0.92, 0.93, 0.95])
        k = np.clip(1.0 - np.abs(up) * 3.0, 0, 1)[..., None]
        col = sky * (1 - k) + horizon * k
        ground = np.array([0.36, 0.33, 0.30])
        col = np.where(up[..., None] < 0, ground * (0.6 + 0.4 * k), col)
        sunk = np.clip((np.sum(d * sun, -1) - 0.995) / 0.005, 0, 1)[..., None]
        col = col * (1 - sunk) + np.array([1.0, 0.98, 0.9]) * sunk
        faces[f, ..., :3] = np.clip(col * 255.0 + 0.5, 0, 255).astype(np.uint8)
        faces[f, ..., 3] = 255
    return faces


def faces_to_cross(faces):
    """Inverse of loadEnvironmentMap's cut (RayTracer.cpp:377-399): 4x3 horizontal cross image."""
    _, h, w, _ = faces.shape
    img = np.zeros((3 * h, 4 * w, 4), np.uint8)
    cell = [(2, 1), (0, 1), (1, 0), (1, 2), (1, 1), (3, 1)]
    for f, (cx, cy) in enumerate(cell):
        img[cy * h:(cy + 1) * h, cx * w:(cx + 1) * w] = faces[f]
    return img


# ------------------------------------------------------------------ materials
def material(**kw):
    m = abi.default_material().copy()
    for k, v in kw.items():
        m[k] = v
    return m


# ------------------------------------------------------------------ C1: TestScene
def test_scene(seed=1234, num_spheres=64, with_textures=False):
    """Config 1: 64 spheres U[0.5,3] on a 40x40 patch + 2 quads + one 1280-tri icosphere."""
    rng = np.random.default_rng(seed)
    rm, sc = ResourceManager(), Scene()
    if with_textures:
        for i, kind in enumerate(("checker", "bricks", "stone", "spec")):
            rm.addTexture(procedural_texture(kind, 128 if i % 2 else 256, seed + i))
    ground = rm.addMesh(make_mesh(*quad((-20, 0, -20), (20, 0, -20), (20, 0, 20), (-20, 0, 20), uv_scale=8.0), name="ground"))
    wall = rm.addMesh(make_mesh(*quad((-20, 0, 20), (20, 0, 20), (20, 25, 20), (-20, 25, 20), uv_scale=4.0), name="wall"))
    ico = rm.addMesh(make_mesh(*icosphere(3, 4.0, (0, 5.0, 4.0)), name="icosphere"))
    sc.addMeshEntity(ground, material(albedo=(0.7, 0.7, 0.65), albedoTextureIdx=0 if with_textures else abi.INVALID_UINT))
    sc.addMeshEntity(wall, material(albedo=(0.6, 0.3, 0.25), albedoTextureIdx=1 if with_textures else abi.INVALID_UINT))
    sc.addMeshEntity(ico, material(albedo=(0.9, 0.8, 0.3), metallic=1.0, roughness=0.15,
                                   specularTextureIdx=3 if with_textures else abi.INVALID_UINT))
    for i in range(num_spheres):
        r = float(rng.uniform(0.5, 3.0))
        x, z = rng.uniform(-20, 20, 2)
        kind = i % 4
        if kind == 0:
            m = material(albedo=tuple(rng.uniform(0.2, 0.9, 3)))
        elif kind == 1:
            m = material(albedo=tuple(rng.uniform(0.5, 1.0, 3)), metallic=1.0, roughness=float(rng.uniform(0.0, 0.4)))
        elif kind == 2:
            m = material(albedo=(1, 1, 1), transparency=1.0, indexOfRefraction=1.5, roughness=0.0)
        else:
            m = material(albedo=(0, 0, 0), emissionColour=tuple(rng.uniform(0.5, 1.0, 3)), emissionPower=float(rng.uniform(1, 4)))
        sc.addSphereEntity(r, m, (float(x), r, float(z)))
    sc.setCamera((0.0, 8.0, -28.0), (10.0, 0.0, 0.0), 90.0, 0.1)
    info = dict(name="TestScene", stand_in=False, octree=(2, 50), bvh=(30, 5))
    return rm, sc, info


# ------------------------------------------------------------------ C2: dragon stand-in
def knot_mesh(num_triangles=871_414, seed=42, scale=20.0):
    """Displaced torus-knot surface with ~num_triangles triangles (stand-in for dragon.obj, 871k tris)."""
    rng = np.random.default_rng(seed)
    nv = 96
    nu = max(8, int(round(num_triangles / (2 * nv))))
    ph = rng.uniform(0, 2 * np.pi, 6)

    def centre(t):
        a = 2 * np.pi * t
        r = 2.0 + np.cos(3 * a)
        return np.stack([r * np.cos(2 * a), 0.9 * np.sin(3 * a), r * np.sin(2 * a)], -1)

    def fn(U, V):
        c = centre(U)
        tangent = centre(U + 1e-4) - centre(U - 1e-4)
        tangent /= np.linalg.norm(tangent, axis=-1, keepdims=True)
        ref = np.array([0.0, 1.0, 0.0])
        n1 = np.cross(tangent, ref)
        n1 /= np.linalg.norm(n1, axis=-1, keepdims=True)
        n2 = np.cross(tangent, n1)
        a = 2 * np.pi * V
        bump = 1.0 + 0.12 * np.sin(40 * 2 * np.pi * U + ph[0]) * np.sin(6 * a + ph[1]) + 0.05 * np.sin(170 * 2 * np.pi * U + ph[2]) * np.cos(11 * a + ph[3])
        rad = 0.55 * bump
        return (c + (n1 * np.cos(a)[..., None] + n2 * np.sin(a)[..., None]) * rad[..., None]) * (scale / 6.0) + np.array([0.0, scale * 0.35, 0.0])
    return make_mesh(*grid_surface(fn, nu, nv, uv_scale=(64.0, 2.0)), name="knot")


def dragon_scene(num_triangles=871_414, seed=42):
    rm, sc = ResourceManager(), Scene()
    knot = rm.addMesh(knot_mesh(num_triangles, seed))
    ground = rm.addMesh(make_mesh(*quad((-60, 0, -60), (60, 0, -60), (60, 0, 60), (-60, 0, 60)), name="ground"))
    sc.addMeshEntity(knot, material(albedo=(0.8, 0.4, 0.1), metallic=1.0, roughness=0.2))     # Pics/dragon.png look
    sc.addMeshEntity(ground, material(albedo=(0.6, 0.6, 0.6)))
    sc.setCamera((0.0, 12.0, -26.0), (12.0, 0.0, 0.0), 90.0, 0.1)
    info = dict(name="dragon", stand_in=True, octree=(2, 50), bvh=(30, 5),
                note="displaced torus-knot stand-in: dragon.obj is listed in .MISSING_LARGE_BLOBS")
    return rm, sc, info


# ------------------------------------------------------------------ C3: glass stand-in
def lathe_mesh(profile, segments=64, rings=40, centre=(0, 0, 0), name="lathe"):
    """Surface of revolution of profile(v) -> (radius, height)."""
    def fn(U, V):
        r, h = profile(V)
        a = 2 * np.pi * U
        return np.stack([r * np.cos(a), h, r * np.sin(a)], -1) + np.asarray(centre, float)
    return make_mesh(*grid_surface(fn, segments, rings, uv_scale=(4.0, 1.0)), name=name)


def glass_scene(segments=64, rings=40):
    """Drinking glass with wall thickness (outer + inner surface, ~5k tris like ikea_glass.obj's 5310 faces)."""
    def profile(V):
        # outer wall up, over the rim, inner wall down, inner bottom
        v = np.asarray(V)
        outer_r, inner_r, height, base = 3.2, 2.9, 11.0, 0.8
        r = np.where(v < 0.45, 2.4 + (outer_r - 2.4) * (v / 0.45),
                     np.where(v < 0.5, outer_r - (outer_r - inner_r) * ((v - 0.45) / 0.05),
                              np.where(v < 0.95, inner_r - (inner_r - 2.2) * ((v - 0.5) / 0.45), 2.2 * (1 - (v - 0.95) / 0.05))))
        h = np.where(v < 0.45, height * (v / 0.45), np.where(v < 0.5, height, np.where(v < 0.95, height - (height - base) * ((v - 0.5) / 0.45), base)))
        return r, h
    rm, sc = ResourceManager(), Scene()
    glass = rm.addMesh(lathe_mesh(profile, segments, rings, name="glass"))
    ground = rm.addMesh(make_mesh(*quad((-40, 0, -40), (40, 0, -40), (40, 0, 40), (-40, 0, 40), uv_scale=10.0), name="ground"))
    sc.addMeshEntity(glass, material(albedo=(1, 1, 1), transparency=1.0, indexOfRefraction=1.5, roughness=0.0))
    sc.addMeshEntity(ground, material(albedo=(0.55, 0.5, 0.45)))
    sc.setCamera((0.0, 7.0, -12.0), (12.0, 0.0, 0.0), 90.0, 0.1)
    info = dict(name="glass", stand_in=True, octree=(2, 50), bvh=(30, 5),
                note="lathe-glass stand-in for ikea_glass.obj (not on the GPU box)")
    return rm, sc, info


def glass_scene_from_meshes(meshes, note):
    """BASELINE configs[2] as Application.cpp:42 would set it up: the cups of ikea_glass.obj (scale 10) as glass
    (transparency 1, IOR 1.5, roughness 0) on a ground quad, sky cube."""
    rm, sc = ResourceManager(), Scene()
    for m in meshes:
        sc.addMeshEntity(rm.addMesh(m), material(albedo=(1, 1, 1), transparency=1.0, indexOfRefraction=1.5, roughness=0.0))
    ground = rm.addMesh(make_mesh(*quad((-20, 0, -20), (20, 0, -20), (20, 0, 20), (-20, 0, 20), uv_scale=10.0), name="ground"))
    sc.addMeshEntity(ground, material(albedo=(0.55, 0.5, 0.45)))
    sc.setCamera((0.65, 1.45, -2.1), (22.0, 0.0, 0.0), 90.0, 0.1)
    info = dict(name="glass", stand_in=False, octree=(2, 50), bvh=(30, 5), note=note)
    return rm, sc, info


def glass_scene_from_fixture(path):
    """The real asset from tests/golden/ikea_glass_mesh.npz (written by tests/golden/make_golden.py from the reference's
    resources/meshes/ikea_glass.obj through okrt.loadObjects)."""
    z = np.load(path)
    meshes, start = [], 0
    for count, name in zip(z["counts"], z["names"]):
        end = start + int(count)
        info = np.zeros(end - start, dtype=abi.TRIANGLE_INFO)
        info["vertexInfo"]["normal"] = z["normals"][start:end]
        info["vertexInfo"]["uv"] = z["uvs"][start:end]
        tp = np.zeros(end - start, dtype=abi.TRIANGLE)
        tp["position"] = z["positions"][start:end]
        meshes.append(Mesh(tp, info, name=str(name)))
        start = end
    return glass_scene_from_meshes(meshes, "ikea_glass.obj via okrt_obj_load (fixture tests/golden/ikea_glass_mesh.npz)")


def glass_scene_from_obj(path, scale=10.0):
    from .raytracer import loadObjects
    return glass_scene_from_meshes([o.mesh for o in loadObjects(path, scale)], "ikea_glass.obj via okrt_obj_load")


# ------------------------------------------------------------------ C4/C5: Sponza stand-in
_SPONZA_MATERIALS = ["bricks", "stone", "stone", "fabric", "fabric", "fabric", "checker", "stone", "bricks", "stone", "noise",
                     "fabric", "stone", "noise", "stone", "bricks", "noise", "stone", "fabric", "fabric", "fabric", "stone",
                     "noise", "stone"]


def sponza_standin(target_triangles=262_267, seed=11, texture_size=1024):
    """Atrium with two arcade levels, columns, arches, curtains, vases and an open roof.

    Stand-in for resources/Sponza/sponza.obj (missing): ~262k triangles in ~390 mesh entities,
    24 diffuse + 3 specular texture layers (as sponza.mtl binds them through assimp: map_Kd ->
    albedo, map_Ks -> specular, map_bump ignored), default materials otherwise
    (roughness 1: diffuse), identity transforms, extents of the 0.07-scaled original
    (x -134..126, y -9..100, z -83..77), camera at (60,0,0) looking down -x (Application.cpp:67-70).
    """
    rng = np.random.default_rng(seed)
    rm, sc = ResourceManager(), Scene()
    for i, kind in enumerate(_SPONZA_MATERIALS):                       # 24 diffuse maps
        size = texture_size if i % 5 else texture_size // 2
        rm.addTexture(procedural_texture(kind, size, seed + i))
    spec_ids = [rm.addTexture(procedural_texture("spec", texture_size // 2, seed + 100 + i)) for i in range(3)]   # 3 specular maps

    def mat(diffuse, spec=None):
        return material(albedoTextureIdx=diffuse, specularTextureIdx=abi.INVALID_UINT if spec is None else spec_ids[spec])

    X0, X1, Y0, Y1, Z0, Z1 = -134.0, 126.0, -9.0, 100.0, -83.0, 77.0
    AX0, AX1, AZ0, AZ1 = -95.0, 87.0, -28.0, 22.0                      # open atrium footprint
    LEVELS = [Y0, 35.0, 70.0]
    parts = []                                                         # (pos, nor, uv, material, weight)

    def add(geom, m):
        parts.append((geom, m))

    # floor: 6 x 4 tiles
    for i in range(6):
        for j in range(4):
            x0, x1 = X0 + (X1 - X0) * i / 6, X0 + (X1 - X0) * (i + 1) / 6
            z0, z1 = Z0 + (Z1 - Z0) * j / 4, Z0 + (Z1 - Z0) * (j + 1) / 4
            fn = (lambda U, V, x0=x0, x1=x1, z0=z0, z1=z1: np.stack([x0 + (x1 - x0) * U, np.full_like(U, Y0) + 0.05 * np.sin(U * 37 + V * 23), z0 + (z1 - z0) * V], -1))
            add(("grid", fn, 14, 14, (6.0, 6.0), False), mat(6, 0))
    # outer walls, two levels x 5 segments per side
    for side in range(4):
        for lv in range(2):
            for seg in range(5):
                y0, y1 = (Y0, 45.0) if lv == 0 else (45.0, Y1)
                a, b = seg / 5.0, (seg + 1) / 5.0
                if side == 0:
                    fn = (lambda U, V, a=a, b=b, y0=y0, y1=y1: np.stack([X0 + (X1 - X0) * (a + (b - a) * U), y0 + (y1 - y0) * V, np.full_like(U, Z1) - 0.3 * np.sin(U * 31) * np.sin(V * 29)], -1))
                    flip = False
                elif side == 1:
                    fn = (lambda U, V, a=a, b=b, y0=y0, y1=y1: np.stack([X0 + (X1 - X0) * (a + (b - a) * U), y0 + (y1 - y0) * V, np.full_like(U, Z0) + 0.3 * np.sin(U * 31) * np.sin(V * 29)], -1))
                    flip = True
                elif side == 2:
                    fn = (lambda U, V, a=a, b=b, y0=y0, y1=y1: np.stack([np.full_like(U, X0) + 0.3 * np.sin(U * 31) * np.sin(V * 29), y0 + (y1 - y0) * V, Z0 + (Z1 - Z0) * (a + (b - a) * U)], -1))
                    flip = False
                else:
                    fn = (lambda U, V, a=a, b=b, y0=y0, y1=y1: np.stack([np.full_like(U, X1) - 0.3 * np.sin(U * 31) * np.sin(V * 29), y0 + (y1 - y0) * V, Z0 + (Z1 - Z0) * (a + (b - a) * U)], -1))
                    flip = True
                add(("grid", fn, 16, 16, (3.0, 3.0), flip), mat(0 if lv == 0 else 8, 1))
    # gallery slabs around the atrium at the two upper levels (ceiling of the arcades)
    for y in LEVELS[1:]:
        ring = [((X0, Z0), (X1, AZ0)), ((X0, AZ1), (X1, Z1)), ((X0, AZ0), (AX0, AZ1)), ((AX1, AZ0), (X1, AZ1))]
        for (xa, za), (xb, zb) in ring:
            for k in range(3):
                xs0, xs1 = xa + (xb - xa) * k / 3, xa + (xb - xa) * (k + 1) / 3
                add(("box", (xs0, y - 1.5, za), (xs1, y, zb)), mat(9))
    # columns + arches along the long sides, two levels
    ncol = 14
    col_x = np.linspace(AX0, AX1, ncol)

    def column_profile(V):
        v = np.asarray(V)
        r = 2.3 - 0.5 * np.clip((v - 0.05) * 8, 0, 1) + 0.5 * np.clip((v - 0.9) * 10, 0, 1) + 0.08 * np.sin(v * 60)
        return r, None
    for lv in range(2):
        y0, y1 = LEVELS[lv], LEVELS[lv + 1] - 9.0
        for z in (AZ0, AZ1):
            for cx in col_x:
                def fn(U, V, cx=cx, z=z, y0=y0, y1=y1):
                    r, _ = column_profile(V)
                    flute = 1.0 + 0.04 * np.cos(U * 2 * np.pi * 16)
                    a = 2 * np.pi * U
                    return np.stack([cx + r * flute * np.cos(a), y0 + (y1 - y0) * V, z + r * flute * np.sin(a)], -1)
                add(("grid", fn, 40, 18, (4.0, 6.0), False), mat(10 + lv, 2))
            for k in range(ncol - 1):
                xa, xb = col_x[k], col_x[k + 1]

                def arch(U, V, xa=xa, xb=xb, z=z, y1=y1):
                    a = np.pi * U
                    cxm, rx = 0.5 * (xa + xb), 0.5 * (xb - xa)
                    depth = (V - 0.5) * 5.0
                    return np.stack([cxm - rx * np.cos(a), y1 + 8.0 * np.sin(a) + 0.15 * np.sin(U * 50), z + depth], -1)
                add(("grid", arch, 48, 10, (4.0, 1.0), False), mat(1))
    # curtains hanging in the upper arcade openings
    for k in range(18):
        cx = col_x[k % (ncol - 1)] + 0.5 * (col_x[1] - col_x[0])
        z = AZ0 - 1.0 if k < 9 else AZ1 + 1.0
        ph = rng.uniform(0, 6.28, 2)

        def curtain(U, V, cx=cx, z=z, ph=ph):
            return np.stack([cx + (U - 0.5) * 11.0, LEVELS[2] - 10.0 - 22.0 * V, z + 0.9 * np.sin(U * 25 + ph[0]) * (0.3 + V) + 0.3 * np.sin(V * 9 + ph[1])], -1)
        add(("grid", curtain, 44, 44, (2.0, 2.0), False), mat(3 + k % 3))
    # vases + plants on the atrium floor
    for k in range(10):
        cx, cz = AX0 + 12.0 + (AX1 - AX0 - 24.0) * (k % 5) / 4.0, (AZ0 + 6.0) if k < 5 else (AZ1 - 6.0)

        def vase(U, V, cx=cx, cz=cz):
            r = 1.2 + 1.6 * np.sin(np.pi * np.clip(V, 0, 1)) ** 2 * (1 - 0.4 * V)
            a = 2 * np.pi * U
            return np.stack([cx + r * np.cos(a), Y0 + 7.0 * V, cz + r * np.sin(a)], -1)
        add(("grid", vase, 36, 28, (2.0, 2.0), False), mat(12, 2))
        leaves = []
        for _ in range(260):
            c = np.array([cx, Y0 + 7.0, cz]) + rng.normal(0, 1.0, 3) * np.array([1.6, 1.2, 1.6]) + np.array([0, 2.0, 0])
            d1, d2 = rng.normal(0, 0.5, 3), rng.normal(0, 0.5, 3)
            leaves.append(quad(c - d1 - d2, c + d1 - d2, c + d1 + d2, c - d1 + d2)[0])
        add(("tris", np.concatenate(leaves)), mat(13))
    # flagpoles + flags
    for k in range(12):
        cx = col_x[1 + k % 6 * 2]
        z = AZ0 + 2.0 if k < 6 else AZ1 - 2.0
        sgn = 1.0 if k < 6 else -1.0

        def pole(U, V, cx=cx, z=z, sgn=sgn):
            a = 2 * np.pi * U
            return np.stack([cx + 0.25 * np.cos(a), LEVELS[1] + 6.0 + 0.25 * np.sin(a), z + sgn * 14.0 * V], -1)
        add(("grid", pole, 12, 24, (1.0, 4.0), False), mat(14))

        def flag(U, V, cx=cx, z=z, sgn=sgn):
            return np.stack([cx + 0.4 * np.sin(V * 12 + U * 5), LEVELS[1] + 6.0 - 12.0 * V, z + sgn * (2.0 + 10.0 * U)], -1)
        add(("grid", flag, 30, 30, (1.0, 1.0), False), mat(18 + k % 3))
    # lion heads + shields on the end walls
    for k in range(2):
        cx = X0 + 3.0 if k == 0 else X1 - 3.0
        p, n, uv = icosphere(4, 1.0)
        bump = 1.0 + 0.15 * np.sin(p[..., 0] * 9) * np.sin(p[..., 1] * 7) * np.sin(p[..., 2] * 8)
        add(("tris", p * bump[..., None] * 5.0 + np.array([cx, 22.0, -3.0]), n, uv), mat(15))
    for k in range(8):
        cx = X0 + 1.0 if k < 4 else X1 - 1.0
        cz = -30.0 + 20.0 * (k % 4)
        p, n, uv = icosphere(3, 1.0)
        add(("tris", p * np.array([0.4, 3.0, 3.0]) + np.array([cx, 50.0, cz]), n, uv), mat(16))
    # chains, lamps and small details
    for k in range(40):
        c = np.array([rng.uniform(AX0, AX1), rng.uniform(LEVELS[1] + 3, LEVELS[2] - 6), rng.choice([AZ0 - 3.0, AZ1 + 3.0])])
        p, n, uv = icosphere(2, 1.0)
        add(("tris", p * rng.uniform(0.4, 1.1) + c, n, uv), mat(17))
    # roof beams across the top (sky visible between them) and side roofs
    for k in range(30):
        x = AX0 + (AX1 - AX0) * (k + 0.5) / 30.0
        add(("box", (x - 0.8, Y1 - 2.5, AZ0 - 2.0), (x + 0.8, Y1 - 0.5, AZ1 + 2.0)), mat(21))
    for (xa, za), (xb, zb) in [((X0, Z0), (X1, AZ0 - 2.0)), ((X0, AZ1 + 2.0), (X1, Z1)), ((X0, AZ0 - 2.0), (AX0, AZ1 + 2.0)), ((AX1, AZ0 - 2.0), (X1, AZ1 + 2.0))]:
        for k in range(4):
            xs0, xs1 = xa + (xb - xa) * k / 4, xa + (xb - xa) * (k + 1) / 4
            add(("box", (xs0, Y1 - 1.0, za), (xs1, Y1, zb)), mat(22))

    # scale grid resolutions so the total lands on target_triangles
    def count(g, s):
        if g[0] == "grid":
            return 2 * max(2, int(round(g[2] * s))) * max(2, int(round(g[3] * s)))
        if g[0] == "box":
            return 12
        return len(g[1])
    lo_s, hi_s = 0.2, 6.0
    for _ in range(40):
        mid = 0.5 * (lo_s + hi_s)
        if sum(count(g, mid) for g, _ in parts) < target_triangles:
            lo_s = mid
        else:
            hi_s = mid
    s = lo_s
    total = 0
    for idx, (g, m) in enumerate(parts):
        if g[0] == "grid":
            geo = grid_surface(g[1], max(2, int(round(g[2] * s))), max(2, int(round(g[3] * s))), uv_scale=g[4], flip=g[5])
        elif g[0] == "box":
            geo = box(g[1], g[2])
        else:
            geo = (g[1], g[2] if len(g) > 2 else None, g[3] if len(g) > 3 else None)
        mesh = make_mesh(*geo, name="sponza_part_%d" % idx)
        total += len(mesh.trianglesPos)
        sc.addMeshEntity(rm.addMesh(mesh), m)
    sc.setCamera((60.0, 0.0, 0.0), (0.0, -90.0, 0.0), 90.0, 0.1)
    info = dict(name="sponza", stand_in=True, octree=(2, 50), bvh=(30, 5), triangles=total, entities=len(parts),
                note="procedural Sponza-class stand-in: sponza.obj is listed in .MISSING_LARGE_BLOBS")
    return rm, sc, info


# ------------------------------------------------------------------ adversarial scenes (differential stress test)
def _flat_grid(p0, du, dv, nu, nv):
    """nu x nv grid of exactly planar quads spanned by du, dv from p0 (float64 lattice, then float32 at upload)."""
    p0, du, dv = np.asarray(p0, float), np.asarray(du, float), np.asarray(dv, float)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    i, j = i.ravel()[:, None], j.ravel()[:, None]
    a = p0 + du * (i / nu) + dv * (j / nv)
    b = p0 + du * ((i + 1) / nu) + dv * (j / nv)
    c = p0 + du * ((i + 1) / nu) + dv * ((j + 1) / nv)
    d = p0 + du * (i / nu) + dv * ((j + 1) / nv)
    return np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)])


def adversarial_scene(kind, offset=(0.0, 0.0, 0.0), seed=5):
    """Scenes built to break a near-first traversal that is not really order independent (tests/test_gpu_stress.py):

    flat_room  axis-aligned, exactly planar floors / walls / boxes cut into many triangles: every leaf box is flat and its
               triangles lie IN a box face; lots of corners, so bounce segments with t ~ 1e-3 next to large surfaces
    layers     stacks of coplanar sheets 1e-7 ... 1e-4 apart (decals), straight and tilted, in different meshes and leaves
    slivers    fans of long needle triangles (aspect 1e2 ... 1e5) in front of a wall
    contact    one 4000-unit ground quad (two huge triangles) touched by finely tessellated spheres: tiny triangles against
               huge ones along the contact curves
    `offset` translates the whole scene (vertices and camera): 1e3 / 1e4 leave 2^-14 / 2^-10 of absolute resolution."""
    rng = np.random.default_rng(seed)
    rm, sc = ResourceManager(), Scene()
    off = np.asarray(offset, float)

    def add(pos, m, name):
        pos = np.asarray(pos, float).reshape(-1, 3, 3) + off
        sc.addMeshEntity(rm.addMesh(make_mesh(pos, name=name)), m)

    grey = material(albedo=(0.75, 0.75, 0.7))
    lamp = material(albedo=(0, 0, 0), emissionColour=(1.0, 0.9, 0.8), emissionPower=3.0)
    mirror = material(albedo=(0.9, 0.9, 0.9), metallic=1.0, roughness=0.05)
    if kind == "flat_room":
        L = 16.0
        faces = [((-L, 0, -L), (2 * L, 0, 0), (0, 0, 2 * L)), ((-L, 12, -L), (0, 0, 2 * L), (2 * L, 0, 0)),
                 ((-L, 0, -L), (0, 12, 0), (2 * L, 0, 0)), ((-L, 0, L), (2 * L, 0, 0), (0, 12, 0)),
                 ((-L, 0, -L), (0, 0, 2 * L), (0, 12, 0)), ((L, 0, -L), (0, 12, 0), (0, 0, 2 * L))]
        for k, (p0, du, dv) in enumerate(faces):
            add(_flat_grid(p0, du, dv, 24, 24), lamp if k == 1 else grey, "room_face_%d" % k)
        for k in range(24):                                     # boxes standing on the floor, touching walls and each other
            w = rng.integers(1, 4, 3).astype(float)
            x, z = float(rng.integers(-14, 12)), float(rng.integers(-14, 12))
            lo, hi = np.array([x, 0.0, z]), np.array([x + w[0], w[1], z + w[2]])
            tris = []
            for p0, du, dv in (((lo[0], lo[1], lo[2]), (hi[0] - lo[0], 0, 0), (0, hi[1] - lo[1], 0)), ((lo[0], lo[1], hi[2]), (hi[0] - lo[0], 0, 0), (0, hi[1] - lo[1], 0)),
                               ((lo[0], lo[1], lo[2]), (0, 0, hi[2] - lo[2]), (0, hi[1] - lo[1], 0)), ((hi[0], lo[1], lo[2]), (0, 0, hi[2] - lo[2]), (0, hi[1] - lo[1], 0)),
                               ((lo[0], hi[1], lo[2]), (hi[0] - lo[0], 0, 0), (0, 0, hi[2] - lo[2])), ((lo[0], lo[1], lo[2]), (hi[0] - lo[0], 0, 0), (0, 0, hi[2] - lo[2]))):
                tris.append(_flat_grid(p0, du, dv, 3, 3))
            add(np.concatenate(tris), mirror if k % 5 == 0 else grey, "box_%d" % k)
        sc.setCamera(tuple(np.array([0.5, 5.0, -14.5]) + off), (8.0, 10.0, 0.0), 90.0, 0.1)
    elif kind == "layers":
        add(_flat_grid((-30, -10, 30), (60, 0, 0), (0, 40, 0), 8, 8), lamp, "back_wall")
        add(_flat_grid((-30, -10, -30), (60, 0, 0), (0, 0, 60), 8, 8), grey, "floor")
        k = 0
        for gap in (1e-7, 1e-6, 1e-5, 1e-4):
            for tilt in (0.0, 0.37):
                c, s = np.cos(tilt), np.sin(tilt)
                x0 = -24.0 + 6.0 * k
                for layer in range(4):                          # four sheets `gap` apart along their common normal
                    n = np.array([s, 0.0, -c]) * gap * layer
                    p0 = np.array([x0, -6.0, 10.0 + 0.5 * k]) + n
                    m = material(albedo=tuple(rng.uniform(0.2, 0.9, 3)), roughness=float(rng.uniform(0.1, 1.0)), metallic=float(layer % 2))
                    add(_flat_grid(p0, (5.0 * c, 0, 5.0 * s), (0, 14.0, 0), 3 + layer, 5 + layer), m, "sheet_%d_%d" % (k, layer))
                k += 1
        sc.setCamera(tuple(np.array([0.0, 2.0, -14.0]) + off), (0.0, 0.0, 0.0), 90.0, 0.1)
    elif kind == "slivers":
        add(_flat_grid((-40, -20, 25), (80, 0, 0), (0, 60, 0), 4, 4), lamp, "back_wall")
        add(_flat_grid((-40, -8, -30), (80, 0, 0), (0, 0, 60), 4, 4), grey, "floor")
        tris = []
        for k in range(6000):
            c = np.array([rng.uniform(-18, 18), rng.uniform(-6, 14), rng.uniform(4, 18)])
            d = rng.normal(0, 1, 3)
            d /= np.linalg.norm(d)
            e = np.cross(d, rng.normal(0, 1, 3))
            e /= np.linalg.norm(e)
            length, width = rng.uniform(4, 40), 10.0 ** rng.uniform(-4, -0.5)
            tris.append([c - d * length * 0.5, c + d * length * 0.5, c + e * width])
        add(np.array(tris), mirror, "needles")
        sc.setCamera(tuple(np.array([0.0, 3.0, -16.0]) + off), (0.0, 0.0, 0.0), 90.0, 0.1)
    elif kind == "contact":
        add(_flat_grid((-2000, 0, -2000), (4000, 0, 0), (0, 0, 4000), 1, 1), grey, "ground")
        for k in range(9):
            r = 1.0 + 0.5 * (k % 3)
            c = np.array([-9.0 + 4.5 * (k % 5), r - (0.0, 1e-3, 0.05)[k % 3], 2.0 + 5.0 * (k // 5)])
            p, _, _ = icosphere(5, r, c)
            add(p, mirror if k % 2 else material(albedo=tuple(rng.uniform(0.3, 0.9, 3))), "ball_%d" % k)
        add(_flat_grid((-30, 30, -30), (0, 0, 60), (60, 0, 0), 2, 2), lamp, "sky_panel")
        sc.setCamera(tuple(np.array([0.0, 1.2, -9.0]) + off), (5.0, 0.0, 0.0), 90.0, 0.1)
    else:
        raise ValueError(kind)
    info = dict(name="adversarial_" + kind, stand_in=False, octree=(2, 50), bvh=(30, 5), offset=tuple(offset))
    return rm, sc, info


def moved_props_scene(seed=17):
    """24 props whose LOCAL boxes lie behind a wall (z 14..26) while two thirds of them are translated in front of it
    (z -16..-4): createOctTree sorts occupants by their untransformed boxes (RayTracer.cpp:562), so the oct-tree nodes that
    list the moved props do not contain them, and the reference culls those nodes against the wall."""
    rng = np.random.default_rng(seed)
    rm, sc = ResourceManager(), Scene()
    for k in range(24):
        c = np.array([rng.uniform(-10, 10), rng.uniform(-10, 10), rng.uniform(15, 25)])
        if k % 2:
            mesh = make_mesh(*box(c - 1.0, c + 1.0))
        else:
            p, n, uv = icosphere(2, 1.2, tuple(c))
            mesh = make_mesh(p, n, uv)
        sc.addMeshEntity(rm.addMesh(mesh), material(albedo=tuple(rng.uniform(0.3, 0.9, 3)), roughness=float(rng.uniform(0.1, 1.0)), metallic=float(k % 2)),
                         position=(float(rng.uniform(-3, 3)), float(rng.uniform(-3, 3)), -30.0) if k % 3 else (0, 0, 0))
    wall = make_mesh(*quad((-40, -40, 10), (40, -40, 10), (40, 40, 10), (-40, 40, 10)))
    sc.addMeshEntity(rm.addMesh(wall), material(albedo=(0.9, 0.9, 0.9)))
    sc.setCamera((0, 0, -30), (0, 0, 0), 80.0, 0.1)
    return rm, sc, dict(name="moved_props", stand_in=False, bvh=(30, 5), octree=(3, 3))
