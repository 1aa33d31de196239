"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the debug box overlay, row f4.

Follows DebugRenderer::renderBvhNodeBBs / renderOctTreeNodeBBs (source/Graphics/DebugRenderer.cpp:264-303, 335-364),
executeDrawMode (:447-506), drawNodeBoundingBox / drawOctTreeNodeBoundingBox (:407-445), updateCameraData (:366-380)
and resources/shaders/DebugBBVS.hlsl:13-33 / DebugBBPS.hlsl.  The reference rasterises the line list with D3D11
(unavailable on this box), so the pixel rule is this repository's own (include/okrt.h, okrt_draw_overlay) restated:
parity of the overlay is unpinned against the reference, pinned only between the two implementations here.

Host part (segment list): float64 numpy with numpy's trigonometry, compared with a tolerance.
Device part (rasterisation of given float32 segments): float32 numpy, compared bit for bit."""
import numpy as np

INVALID = 0xFFFFFFFF
BVH, OCTREE = 0, 1
NONE, SINGLE, WITH_CHILDREN, WITH_DESCENDANTS = 0, 1, 2, 3
BVH_NODE_COLOUR = (0.9, 0.7, 0.5)          # DebugRenderer.cpp:10
OCT_NODE_COLOUR = (0.4, 0.7, 0.4)          # DebugRenderer.cpp:12
EDGES = ((0, 1), (2, 3), (4, 5), (6, 7), (0, 2), (1, 3), (4, 6), (5, 7), (0, 4), (1, 5), (2, 6), (3, 7))


def draw_order(kind, nodes, node_idx, mode):
    """executeDrawMode, DebugRenderer.cpp:447-506."""
    def children(i):
        first = int(nodes[i]["firstChildIdx"])
        if first == INVALID:
            return []
        count = 2 if kind == BVH else int(nodes[i]["numChildren"])
        return [first + c for c in range(count)]
    if mode == NONE or node_idx >= len(nodes):
        return []
    if mode == SINGLE:
        return [node_idx]
    if mode == WITH_CHILDREN:
        return [node_idx] + children(node_idx)
    order, stack = [], [node_idx]
    while stack:
        cur = stack.pop()
        stack.extend(children(cur))
        order.append(cur)
    return order


def _rotation(rotation_deg):
    """glm::toMat3(glm::quat(glm::radians(rotation))) columns, float64."""
    ex, ey, ez = np.radians(np.asarray(rotation_deg, np.float64))
    cx, cy, cz = np.cos(ex / 2), np.cos(ey / 2), np.cos(ez / 2)
    sx, sy, sz = np.sin(ex / 2), np.sin(ey / 2), np.sin(ez / 2)
    w = cx * cy * cz + sx * sy * sz
    x = sx * cy * cz - cx * sy * sz
    y = cx * sy * cz + sx * cy * sz
    z = cx * cy * sz - sx * sy * cz
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y)],
                     [2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x)],
                     [2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y)]])      # rows = columns of the matrix


def segments(kind, nodes, node_idx, mode, object_matrix, camera, width, height):
    """Returns (segments Nx4 float64 pixels, keys N uint32, colours per draw as uint32 RGBA8).
    object_matrix: 4x4 row-major world matrix (okrt_transform_matrix) or None; camera: abi.CAMERA record."""
    order = draw_order(kind, nodes, node_idx, mode)
    pos = np.asarray(camera["position"], np.float64)
    f = _rotation(camera["rotation"])[2]
    f = f / np.linalg.norm(f)
    s = np.cross((0.0, 1.0, 0.0), f)
    s = s / np.linalg.norm(s)
    u = np.cross(f, s)
    rad = np.radians(float(camera["fov"]))
    ph = np.cos(rad / 2) / np.sin(rad / 2)
    pw = ph * height / width
    zn, zf = float(camera["nearZ"]), float(camera["farZ"])
    M = np.eye(4) if object_matrix is None else np.asarray(object_matrix, np.float64).reshape(4, 4)
    segs, keys, colours = [], [], []
    for k, n in enumerate(order):
        strength = 1.0 - n / len(nodes)
        rgb = [c * strength for c in BVH_NODE_COLOUR] if kind == BVH else list(OCT_NODE_COLOUR)
        q = [int(np.float32(min(max(np.float32(c), 0), 1)) * np.float32(255) + np.float32(0.5)) for c in rgb]
        colours.append(q[0] | (q[1] << 8) | (q[2] << 16) | (0xFF << 24))
        lo = np.asarray(nodes[n]["boundingBox"]["min"], np.float64)
        hi = np.asarray(nodes[n]["boundingBox"]["max"], np.float64)
        view = []
        for c in range(8):
            unit = np.array([1.0 if (c >> a) & 1 else -1.0 for a in range(3)])
            local = unit * ((hi - lo) * 0.5) + (hi + lo) * 0.5                        # DebugBBVS.hlsl:27-30
            world = M[:3, :3] @ local + M[:3, 3]
            rel = world - pos
            view.append(np.array([s @ rel, u @ rel, f @ rel]))
        for a_i, c_i in EDGES:
            a, c = view[a_i], view[c_i]
            t0, t1, dz, visible = 0.0, 1.0, c[2] - a[2], True
            for plane, sign in ((zn, 1.0), (zf, -1.0)):
                da, dc = sign * (a[2] - plane), sign * (c[2] - plane)
                if da < 0 and dc < 0:
                    visible = False
                elif da < 0:
                    t0 = max(t0, (plane - a[2]) / dz)
                elif dc < 0:
                    t1 = min(t1, (plane - a[2]) / dz)
            if not visible or t0 > t1:
                continue
            seg = []
            for t in (t0, t1):
                p = a + (c - a) * t
                seg += [((pw * p[0] / p[2]) * 0.5 + 0.5) * width, (1.0 - ((ph * p[1] / p[2]) * 0.5 + 0.5)) * height]
            segs.append(seg)
            keys.append(k + 1)
    return np.array(segs, np.float64).reshape(-1, 4), np.array(keys, np.uint32), np.array(colours, np.uint32)


def rasterize(segs_f32, keys, colours, image):
    """okrt_draw_overlay's pixel rule on float32 segments; `image` (H x W x 4 uint8) is modified in place."""
    H, W = image.shape[:2]
    f = np.float32
    pixel_key = np.zeros((H, W), np.uint32)
    for (x0, y0, x1, y1), key in zip(np.asarray(segs_f32, np.float32), keys):
        dx, dy = f(x1 - x0), f(y1 - y0)
        x_major = abs(dx) >= abs(dy)
        a0, a1, b0, db, da = (x0, x1, y0, dy, dx) if x_major else (y0, y1, x0, dx, dy)
        if da == 0 or not all(np.isfinite(v) for v in (a0, a1, b0, db)):
            continue
        lo, hi = min(a0, a1), max(a0, a1)
        major_limit, minor_limit = (W, H) if x_major else (H, W)
        first = max(np.ceil(f(lo - f(0.5))), f(0))
        last = min(np.ceil(f(hi - f(0.5))), f(major_limit))
        if not last > first:
            continue
        c = np.arange(int(first), int(last)).astype(np.float32)
        t = ((c + f(0.5)) - a0) / da
        b = np.floor(b0 + t * db)
        ok = (b >= 0) & (b < minor_limit)
        ci, bi = c[ok].astype(np.int64), b[ok].astype(np.int64)
        ys, xs = (bi, ci) if x_major else (ci, bi)
        np.maximum.at(pixel_key, (ys, xs), key)
    hit = pixel_key > 0
    col = np.asarray(colours, np.uint32)[pixel_key[hit] - 1]
    image[hit] = np.stack([col & 0xFF, (col >> 8) & 0xFF, (col >> 16) & 0xFF, col >> 24], axis=-1).astype(np.uint8)
    return image
