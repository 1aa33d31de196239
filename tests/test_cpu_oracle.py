"""CPU tests of the oracle: known-answer vectors (the reference ships none, SURVEY.md §4 — these pin OUR restatement
against regressions and against independent implementations: the host libm, the reference's compiled BvhBuilder/glm)."""
import ctypes
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _f(*v):
    return np.array(v, np.float32)


def test_kat1_pcg_hash_sequence(oracle):
    """KAT-1: the buggy PCG (GPU-Utilities.hlsli:18-25, F7): multiplier is 747796405+2891336453, no increment."""
    g = np.load(os.path.join(GOLD, "kat_rng_math.npz"))
    seed = ctypes.c_uint32(1)
    got = [oracle.lib.oracle_pcg_hash(ctypes.byref(seed)) for _ in range(16)]
    assert got == list(g["hashes"])
    # independent restatement in Python integers
    s, exp = 1, []
    for _ in range(16):
        s = (s * 3639132858) & 0xFFFFFFFF
        s = ((((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xFFFFFFFF)
        s = ((s >> 22) ^ s) & 0xFFFFFFFF
        exp.append(s)
    assert got == exp
    seed = ctypes.c_uint32(12345)
    fl = np.array([oracle.lib.oracle_random_float(ctypes.byref(seed)) for _ in range(16)], np.float32)
    assert np.array_equal(fl, g["floats"])
    assert np.all((fl >= 0) & (fl <= 1))


def test_kat2_ray_triangle_edges(oracle):
    """KAT-2: Moeller-Trumbore with E1=p0-p2, E2=p1-p2 (GPU-Utilities.hlsli:117-155): u->p0, v->p1, eps slack."""
    L = oracle.lib
    p0, p1, p2 = _f(1, 0, 5), _f(0, 1, 5), _f(0, 0, 5)
    uv = np.zeros(2, np.float32)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    d = _f(0, 0, 1)
    t = L.oracle_ray_triangle(P(_f(0.25, 0.5, 0)), P(d), P(p0), P(p1), P(p2), P(uv))
    assert t == 5.0 and uv[0] == 0.25 and uv[1] == 0.5
    # exactly on the vertices / edges: accepted
    for o in (_f(0, 0, 0), _f(1, 0, 0), _f(0, 1, 0), _f(0.5, 0.5, 0), _f(0, 0.5, 0)):
        assert L.oracle_ray_triangle(P(o), P(d), P(p0), P(p1), P(p2), P(uv)) == 5.0
    # just outside by more than the 1e-6 slack: rejected; inside the slack: accepted
    assert L.oracle_ray_triangle(P(_f(-1e-5, 0.5, 0)), P(d), P(p0), P(p1), P(p2), P(uv)) == -1.0
    assert L.oracle_ray_triangle(P(_f(-5e-7, 0.5, 0)), P(d), P(p0), P(p1), P(p2), P(uv)) == 5.0
    assert L.oracle_ray_triangle(P(_f(0.6, 0.6, 0)), P(d), P(p0), P(p1), P(p2), P(uv)) == -1.0        # u+v > 1
    # parallel ray, |det| < 1e-6
    assert L.oracle_ray_triangle(P(_f(0, 0, 0)), P(_f(1, 0, 0)), P(p0), P(p1), P(p2), P(uv)) == -1.0
    # behind the origin
    assert L.oracle_ray_triangle(P(_f(0.2, 0.2, 10)), P(d), P(p0), P(p1), P(p2), P(uv)) == -1.0
    # two-sided
    assert L.oracle_ray_triangle(P(_f(0.2, 0.2, 10)), P(_f(0, 0, -1)), P(p0), P(p1), P(p2), P(uv)) == 5.0


def test_kat2_ray_box_and_sphere(oracle):
    L = oracle.lib
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    FLT_MAX = np.float32(3.402823466e+38)
    bmin, bmax = _f(-1, -1, -1), _f(1, 1, 1)
    assert L.oracle_ray_aabb_dist(P(_f(0, 0, -5)), P(_f(0, 0, 1)), P(bmin), P(bmax)) == 4.0
    assert L.oracle_ray_aabb_dist(P(_f(0, 0, 0)), P(_f(0, 0, 1)), P(bmin), P(bmax)) == -1.0          # inside: negative near
    assert L.oracle_ray_aabb_dist(P(_f(0, 0, 5)), P(_f(0, 0, 1)), P(bmin), P(bmax)) == FLT_MAX       # behind
    assert L.oracle_ray_aabb_dist(P(_f(3, 0, -5)), P(_f(0, 0, 1)), P(bmin), P(bmax)) == FLT_MAX      # miss
    # flat box (floor quad): near == far == hit distance
    assert L.oracle_ray_aabb_dist(P(_f(0, 2, 0)), P(_f(0, -1, 0)), P(_f(-1, 0, -1)), P(_f(1, 0, 1))) == 2.0
    # sphere: outside -> near root, inside -> far root, miss -> -1 [Q15]
    assert L.oracle_ray_sphere(P(_f(0, 0, -5)), P(_f(0, 0, 1)), P(_f(0, 0, 0)), 1.0) == 4.0
    assert L.oracle_ray_sphere(P(_f(0, 0, 0)), P(_f(0, 0, 1)), P(_f(0, 0, 0)), 1.0) == 1.0
    assert L.oracle_ray_sphere(P(_f(0, 3, -5)), P(_f(0, 0, 1)), P(_f(0, 0, 0)), 1.0) == -1.0
    assert L.oracle_ray_sphere(P(_f(0, 0, 5)), P(_f(0, 0, 1)), P(_f(0, 0, 0)), 1.0) == -1.0


def test_canonical_math_against_libm(oracle):
    """The canonical float32 transcendentals (oracle_math.h) stay within a few ulp of the host's double-precision libm."""
    L = oracle.lib
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)

    def ulps(got, exact):
        exact32 = exact.astype(np.float32)
        sp = np.spacing(np.abs(exact32)).astype(np.float64)
        return np.abs(got.astype(np.float64) - exact) / np.maximum(sp, 1e-45)
    rng = np.random.default_rng(3)
    x = np.concatenate([rng.uniform(2.0 ** -32, 1.0, 20000), rng.uniform(1, 1e6, 2000)]).astype(np.float32)
    o = np.zeros_like(x); L.oracle_math_eval(0, P(x), P(x), P(o), len(x))
    ok = np.abs(np.log(x.astype(np.float64))) > 1e-3
    assert ulps(o, np.log(x.astype(np.float64)))[ok].max() <= 2.0
    a = rng.uniform(0, 2 * np.pi, 20000).astype(np.float32)
    for which, fn in ((1, np.cos), (2, np.sin)):
        o = np.zeros_like(a); L.oracle_math_eval(which, P(a), P(a), P(o), len(a))
        assert np.abs(o.astype(np.float64) - fn(a.astype(np.float64))).max() <= 2.5e-7       # D3D11 allows 8e-4
    e = rng.uniform(-80, 80, 20000).astype(np.float32)
    o = np.zeros_like(e); L.oracle_math_eval(3, P(e), P(e), P(o), len(e))
    assert ulps(o, np.exp(e.astype(np.float64))).max() <= 2.0
    b = rng.uniform(0, 1, 20000).astype(np.float32); five = np.full_like(b, 5)
    o = np.zeros_like(b); L.oracle_math_eval(5, P(b), P(five), P(o), len(b))
    exact = b.astype(np.float64) ** 5
    big = exact > 1e-30
    assert (np.abs(o[big] - exact[big]) / exact[big]).max() <= 2e-5                          # exp(5*log x): D3D's own definition
    yy = rng.uniform(-1, 1, 20000).astype(np.float32); xx = rng.uniform(-1, 1, 20000).astype(np.float32)
    o = np.zeros_like(yy); L.oracle_math_eval(4, P(yy), P(xx), P(o), len(yy))
    assert np.abs(o.astype(np.float64) - np.arctan2(yy.astype(np.float64), xx.astype(np.float64))).max() <= 1e-6
    # special values
    sp = _f(0.0, -1.0, np.inf, np.nan)
    o = np.zeros_like(sp); L.oracle_math_eval(0, P(sp), P(sp), P(o), 4)
    assert o[0] == -np.inf and np.isnan(o[1]) and o[2] == np.inf and np.isnan(o[3])


def test_canonical_math_golden(oracle):
    """Bit-level regression pin of the canonical math (the GPU restates the same algorithms)."""
    g = np.load(os.path.join(GOLD, "kat_rng_math.npz"))
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    for which, name in ((0, "log"), (1, "cos"), (2, "sin"), (3, "exp")):
        x = np.ascontiguousarray(g[name + "_x"]); o = np.zeros_like(x)
        oracle.lib.oracle_math_eval(which, P(x), P(x), P(o), len(x))
        assert np.array_equal(o.view(np.uint32), g[name + "_y"].view(np.uint32)), name
    x = np.ascontiguousarray(g["pow5_x"]); five = np.full_like(x, 5.0); o = np.zeros_like(x)
    oracle.lib.oracle_math_eval(5, P(x), P(five), P(o), len(x))
    assert np.array_equal(o.view(np.uint32), g["pow5_y"].view(np.uint32))


@pytest.mark.parametrize("name,build,spheres", [("testscene", "test", True), ("sponza20k", "sponza", False)])
def test_oracle_render_golden(okrt, oracle, name, build, spheres):
    """The oracle reproduces its committed AOV/radiance fixtures (where the numpy scene generator is bit-reproducible)."""
    from tests import helpers
    g = np.load(os.path.join(GOLD, "render_%s.npz" % name))
    W, H, bounces = int(g["W"]), int(g["H"]), int(g["bounces"])
    rm, sc, info = okrt.scenes.test_scene(with_textures=True) if build == "test" else okrt.scenes.sponza_standin(20000, texture_size=32)
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=spheres)
    env = okrt.scenes.procedural_sky_faces(16)
    if helpers.scene_digest(host, env) != str(g["scene_digest"]):
        pytest.skip("numpy generated a scene that differs in the last bit on this CPU; fixture not comparable")
    rd = host["rd"].copy(); rd[0]["numAccumulationFrames"] = 1
    o = oracle.render(helpers.oracle_scene(oracle, host, envFaces=env), rd, bounces, aov_dtype=okrt.abi.AOV)
    assert o["segments"] == int(g["segments"])
    assert np.array_equal(o["aov"]["triIdx"], g["aov"]["triIdx"])
    assert np.array_equal(o["accum"].view(np.uint32), g["accum"].view(np.uint32))
    assert np.array_equal(o["result"], g["result"])


def test_oracle_libm_mode_is_statistically_equal(okrt, oracle):
    """math=libm (independent transcendentals) diverges bit-wise after the first bounce but not statistically:
    primary hits are identical and the mean radiance agrees."""
    from tests import helpers, oracle_binding
    rm, sc, info = okrt.scenes.test_scene()
    W, H = 96, 54
    host = helpers.build_host_side(okrt, rm, sc, info, W, H, spheres=True)
    env = okrt.scenes.procedural_sky_faces(16)
    s = helpers.oracle_scene(oracle, host, envFaces=env)
    acc = [np.zeros((H, W, 4), np.float32), np.zeros((H, W, 4), np.float32)]
    rd = host["rd"].copy()
    for f in range(1, 33):
        rd[0]["numAccumulationFrames"] = f
        a = oracle.render(s, rd, 4, accum=acc[0], aov_dtype=okrt.abi.AOV, math=oracle_binding.MATH_CANONICAL)
        b = oracle.render(s, rd, 4, accum=acc[1], aov_dtype=okrt.abi.AOV, math=oracle_binding.MATH_LIBM)
    assert np.array_equal(a["aov"]["triIdx"], b["aov"]["triIdx"])
    ma, mb = acc[0][..., :3].mean() / 32, acc[1][..., :3].mean() / 32
    assert abs(ma - mb) / max(ma, 1e-6) < 0.03


def test_unorm8_division_by_refinement():
    """csrc/okrt_kernels.cuh unorm8ToFloat: q = fma(fma(-255, q0, k), r, q0) with q0 = RN(k * r), r = RN(1/255), equals the
    correctly rounded k / 255 (what the oracle's and the shader's division gives) for all 256 texel values —
    exact rational arithmetic, one rounding per fused operation."""
    from fractions import Fraction

    def rn(x):
        if x == 0:
            return Fraction(0)
        sign = -1 if x < 0 else 1
        x = abs(x)
        e = 0
        while Fraction(2) ** e > x:
            e -= 1
        while Fraction(2) ** (e + 1) <= x:
            e += 1
        ulp = Fraction(2) ** (e - 23)
        q = x / ulp
        fl = q.numerator // q.denominator
        rem = q - fl
        if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and fl % 2 == 1):
            fl += 1
        return sign * fl * ulp

    r = rn(Fraction(1, 255))
    assert r == Fraction(int(np.float32(np.uint32(0x3B808081).view(np.float32)) * 2 ** 40), 2 ** 40) or float(r) == float(np.uint32(0x3B808081).view(np.float32))
    for k in range(256):
        q0 = rn(Fraction(k) * r)
        q = rn(q0 + rn(Fraction(k) - 255 * q0) * r)
        assert q == rn(Fraction(k, 255)), k
        assert float(q) == float(np.float32(k) / np.float32(255.0))
