"""TEST INFRASTRUCTURE — not part of the product.  Only tests/ may import this module.

Independent restatement (numpy float32, plain Python loops) of what the reference obtains from
Importer::loadObjects (source/Graphics/Importer.cpp:77-120) + aiMeshToMeshData (:15-56) for a Wavefront OBJ:
assimp 5.x's ObjFileParser / ObjFileImporter followed by the post-process steps the reference requests
(aiProcess_Triangulate | aiProcess_ConvertToLeftHanded | aiProcess_CalcTangentSpace), in assimp's fixed step order
MakeLeftHanded, FlipUVs, FlipWindingOrder, Triangulate, CalcTangents.

Parity unpinned: assimp is a prebuilt Windows library in the reference (deps/lib/assimp-vc143-mt.lib, no source in
the tree, not runnable here) and the reference holds no test or golden mesh for it.  This module restates the
published algorithm; tests/test_cpu_obj.py checks the product's C++ loader (csrc/okrt_obj.cpp) against it bit for bit.
"""
import math

import numpy as np

F = np.float32
TEX_SLOTS = ("albedo", "roughness", "metallic", "specular", "normal")


def _f(x):
    return F(x)


def parse_mtl(text):
    """newmtl blocks -> {name: {slot: path, 'displacement': path}}; options before the file name are skipped."""
    mats, cur = {}, None
    keymap = {"map_kd": "albedo", "map_pr": "roughness", "map_pm": "metallic", "map_ks": "specular",
              "norm": "normal", "map_kn": "normal", "disp": "displacement", "map_disp": "displacement"}
    for line in text.splitlines():
        toks = line.split()
        if not toks:
            continue
        key = toks[0].lower()
        if key == "newmtl":
            cur = mats.setdefault(" ".join(toks[1:]), {})
        elif cur is not None and key in keymap:
            rest, i = toks[1:], 0
            while i < len(rest) and len(rest[i]) > 1 and rest[i][0] == "-":
                i += 1 + (3 if rest[i] in ("-o", "-s", "-t") else 2 if rest[i] == "-mm" else 1)
            cur[keymap[key]] = " ".join(rest[i:]).replace("\\", "/")
    return mats


def parse_obj(text, materials):
    """ObjFileParser: returns (positions, uvs, normals, objects) with objects = [[mesh, ...], ...] in file order and
    mesh = dict(name, material, faces=[[(v, vt, vn), ...], ...])."""
    pos, uvs, nors, objects = [], [], [], []
    state = dict(obj=None, mesh=None, mat=None, group="")

    def create_mesh(name):
        m = dict(name=name, material=state["mat"], faces=[])
        if state["obj"] is None:
            objects.append([])
            state["obj"] = objects[-1]
        state["obj"].append(m)
        state["mesh"] = m

    def create_object(name):
        objects.append([])
        state["obj"] = objects[-1]
        create_mesh(name)

    def idx(tok, count):
        if not tok:
            return -1
        i = int(tok)
        return i - 1 if i > 0 else count + i if i < 0 else -1

    for line in text.splitlines():
        toks = line.split()
        if not toks:
            continue
        key = toks[0]
        if key in ("v", "vn", "vt"):
            vals = [F(t) for t in toks[1:4]] + [F(0)] * 3
            (pos if key == "v" else nors if key == "vn" else uvs).append(vals[:3])
        elif key == "f":
            corners = []
            for t in toks[1:]:
                parts = (t.split("/") + ["", ""])[:3]
                v, vt, vn = idx(parts[0], len(pos)), idx(parts[1], len(uvs)), idx(parts[2], len(nors))
                if 0 <= v < len(pos):
                    corners.append((v, vt if vt < len(uvs) else -1, vn if vn < len(nors) else -1))
            if corners:
                if state["mesh"] is None:
                    create_object("defaultobject") if state["obj"] is None else create_mesh("defaultobject")
                if state["mesh"]["material"] is None:
                    state["mesh"]["material"] = state["mat"]
                state["mesh"]["faces"].append(corners)
        elif key == "o":
            name = line.split(None, 1)[1].strip() if len(toks) > 1 else ""
            if name:
                create_object(name)
        elif key == "g":
            name = line.split(None, 1)[1].strip() if len(toks) > 1 else ""
            if name != state["group"]:
                create_object(name)
                state["group"] = name
        elif key == "usemtl":
            name = line.split(None, 1)[1].strip() if len(toks) > 1 else ""
            materials.setdefault(name, {})
            m = state["mesh"]
            if m is None or (m["material"] is not None and m["material"] != name and m["faces"]):
                state["mat"] = name
                create_mesh(name)
            state["mat"] = name
            state["mesh"]["material"] = name
    return pos, uvs, nors, objects


def _sub(a, b):
    return [F(a[0] - b[0]), F(a[1] - b[1]), F(a[2] - b[2])]


def _dot(a, b):
    return F(F(F(a[0] * b[0]) + F(a[1] * b[1])) + F(a[2] * b[2]))


def _scale(a, s):
    return [F(a[0] * s), F(a[1] * s), F(a[2] * s)]


def _length(a):
    return F(np.sqrt(_dot(a, a)))


def _normalize(a):
    return _scale(a, F(F(1) / _length(a)))


def _normalize_safe(a):
    l = _length(a)
    return _scale(a, F(F(1) / l)) if l > F(0) else list(a)


def _cross(a, b):
    return [F(F(a[1] * b[2]) - F(a[2] * b[1])), F(F(a[2] * b[0]) - F(a[0] * b[2])), F(F(a[0] * b[1]) - F(a[1] * b[0]))]


def _special(v):
    return any(not np.isfinite(c) for c in v)


def triangulate(verts, idx):
    """TriangulateProcess for one face (corner indices): triangles pass, quads fan from the concave corner,
    larger polygons fan from corner 0."""
    n = len(idx)
    if n < 3:
        return []
    if n == 3:
        return list(idx)
    if n == 4:
        start = 0
        for i in range(4):
            v0, v1, v2, v = verts[idx[(i + 3) % 4]], verts[idx[(i + 2) % 4]], verts[idx[(i + 1) % 4]], verts[idx[i]]
            left, diag, right = _normalize(_sub(v0, v)), _normalize(_sub(v1, v)), _normalize(_sub(v2, v))
            angle = F(F(np.arccos(_dot(left, diag))) + F(np.arccos(_dot(right, diag))))
            if angle > F(math.pi):
                start = i
                break
        t = idx
        return [t[start], t[(start + 1) % 4], t[(start + 2) % 4], t[start], t[(start + 2) % 4], t[(start + 3) % 4]]
    out = []
    for k in range(1, n - 1):
        out += [idx[0], idx[k], idx[k + 1]]
    return out


def calc_tangents(pos, nor, uv, tris):
    """CalcTangentsProcess::ProcessMesh: per-face frames, then the smoothing pass over coincident vertices."""
    nv = len(pos)
    nan = F(np.nan)
    tang = [[nan, nan, nan] for _ in range(nv)]
    bit = [[nan, nan, nan] for _ in range(nv)]
    with np.errstate(all="ignore"):
        for f in range(0, len(tris) - 2, 3):
            p0, p1, p2 = tris[f], tris[f + 1], tris[f + 2]
            v, w = _sub(pos[p1], pos[p0]), _sub(pos[p2], pos[p0])
            sx, sy = F(uv[p1][0] - uv[p0][0]), F(uv[p1][1] - uv[p0][1])
            tx, ty = F(uv[p2][0] - uv[p0][0]), F(uv[p2][1] - uv[p0][1])
            dirc = F(-1) if F(F(tx * sy) - F(ty * sx)) < F(0) else F(1)
            if F(sx * ty) == F(sy * tx):
                sx, sy, tx, ty = F(0), F(1), F(1), F(0)
            tangent = [F(F(F(w[k] * sy) - F(v[k] * ty)) * dirc) for k in range(3)]
            bitangent = [F(F(F(F(-w[k]) * sx) + F(v[k] * tx)) * dirc) for k in range(3)]
            for p in (p0, p1, p2):
                lt = _sub(tangent, _scale(nor[p], _dot(tangent, nor[p])))
                lb = _sub(_sub(bitangent, _scale(nor[p], _dot(bitangent, nor[p]))), _scale(lt, _dot(bitangent, lt)))
                lt, lb = _normalize_safe(lt), _normalize_safe(lb)
                bad_t, bad_b = _special(lt), _special(lb)
                if bad_t != bad_b:
                    if bad_t:
                        lt = _normalize_safe(_cross(nor[p], lb))
                    else:
                        lb = _normalize_safe(_cross(lt, nor[p]))
                tang[p], bit[p] = lt, lb
        P = np.array(pos, dtype=np.float32).reshape(-1, 3)
        lo, hi = P.min(axis=0), P.max(axis=0)
        eps = F(_length(_sub(list(hi), list(lo))) * F(1e-4))
        angle_eps, f_limit = F(0.9999), F(np.cos(F(F(45.0) * F(math.pi)) / F(180.0)))
        order = sorted(range(nv), key=lambda i: (P[i, 0], i))
        xs = P[order, 0]
        done = [False] * nv
        for a in range(nv):
            if done[a]:
                continue
            o_pos, o_nor, o_t, o_b = pos[a], nor[a], tang[a], bit[a]
            lo_k = int(np.searchsorted(xs, F(o_pos[0]) - eps - F(1e-30), side="left"))
            hi_k = int(np.searchsorted(xs, F(o_pos[0]) + eps + F(1e-30), side="right"))
            found = []
            for k in range(max(0, lo_k - 1), min(nv, hi_k + 1)):
                i = order[k]
                if abs(F(P[i, 0] - o_pos[0])) > eps:
                    continue
                if i == a or _length(_sub(pos[i], o_pos)) < eps:
                    found.append(i)
            close = []
            for i in sorted(found):
                if done[i]:
                    continue
                if i != a and (_dot(nor[i], o_nor) < angle_eps or _dot(tang[i], o_t) < f_limit or _dot(bit[i], o_b) < f_limit):
                    continue
                close.append(i)
                done[i] = True
            st, sb = [F(0)] * 3, [F(0)] * 3
            for i in close:
                st = [F(st[k] + tang[i][k]) for k in range(3)]
                sb = [F(sb[k] + bit[i][k]) for k in range(3)]
            st, sb = _normalize(st), _normalize(sb)
            for i in close:
                tang[i], bit[i] = st, sb
    return tang, bit


def load_objects(obj_text, mtl_text="", scale=1.0):
    """Returns a list of dicts(name, positions (T,3,3), normals (T,3,3), uvs (T,3,2), tangents, bitangents,
    box_min, box_max, textures {slot: path}) in the order the reference's loop over aiScene::mMeshes sees them."""
    materials = parse_mtl(mtl_text) if mtl_text else {}
    positions, uvs, normals, objects = parse_obj(obj_text, materials)
    scale = F(scale)
    out = []
    for obj in objects:
        for m in obj:
            if not m["faces"]:
                continue
            pos, nor, uv, faces = [], [], [], []
            has_n = any(c[2] >= 0 for f in m["faces"] for c in f)
            has_uv = any(c[1] >= 0 for f in m["faces"] for c in f)
            if not has_n:
                raise ValueError("mesh '%s' has no normals" % m["name"])
            for f in m["faces"]:
                idx = []
                for (v, vt, vn) in f:
                    idx.append(len(pos))
                    pos.append(list(positions[v]))
                    nor.append(list(normals[vn]) if vn >= 0 else [F(0)] * 3)
                    uv.append(list(uvs[vt]) if vt >= 0 else [F(0)] * 3)
                faces.append(idx)
            for p in pos:                                   # MakeLeftHanded
                p[2] = F(-p[2])
            for n in nor:
                n[2] = F(-n[2])
            if has_uv:                                      # FlipUVs
                for t in uv:
                    t[1] = F(F(1) - t[1])
            faces = [f[::-1] for f in faces]                # FlipWindingOrder
            tris = []
            for f in faces:                                 # Triangulate
                tris += triangulate(pos, f)
            has_t = has_uv and len(tris) > 0
            if has_t:
                tang, bit = calc_tangents(pos, nor, uv, tris)
            nt = len(tris) // 3
            P = np.zeros((nt, 3, 3), np.float32); N = np.zeros((nt, 3, 3), np.float32); U = np.zeros((nt, 3, 2), np.float32)
            T = np.zeros((nt, 3, 3), np.float32); B = np.zeros((nt, 3, 3), np.float32)
            for t in range(nt):
                for j in range(3):
                    i = tris[t * 3 + j]
                    P[t, j] = [F(pos[i][k] * scale) for k in range(3)]
                    N[t, j] = nor[i]
                    if has_uv:
                        U[t, j] = uv[i][:2]
                    if has_t:
                        T[t, j] = tang[i]
                        B[t, j] = bit[i]
            mat = materials.get(m["material"], {}) if m["material"] is not None else {}
            tex = {s: mat.get(s, "") for s in TEX_SLOTS}
            if not tex["normal"]:
                tex["normal"] = mat.get("displacement", "")
            flat = P.reshape(-1, 3)
            out.append(dict(name=m["name"], positions=P, normals=N, uvs=U, tangents=T, bitangents=B,
                            box_min=flat.min(axis=0), box_max=flat.max(axis=0), textures=tex))
    return out
