"""The tree's slab test is FP32 and must be conservative: a ray the FP64 sphere test hits may
never be turned away by the box of any node above that sphere, and the entry distance the
traversal orders and culls by may never exceed the hit distance.  This restates
csrc/rt_accel.cuh (ray_slabs, slab_enter) in NumPy float32 and checks both properties on the
boxes rt_accel_build_host makes, over scenes from millimetres to 1e5 units, grazing rays,
rays parallel to an axis and origins inside or on the boxes.  CPU only."""
import ctypes

import numpy as np
import pytest

from python_ray_tracer_b200.infrastructure.cuda import abi
from python_ray_tracer_b200.infrastructure.flatten import FlatScene

from test_host_logic import _parse_accel

f32 = np.float32


def _build(centres, radii):
    n = len(radii)
    table = np.zeros((12, n))
    table[0:3] = centres.T
    table[3] = radii
    table[4:8] = 1.0
    flat = FlatScene(np.ascontiguousarray(table), np.array([[5.0, 10.0, -10.0]]), (0.0, 0.2, -2.0), [0.1], (1.0, 1.0, 1.0))
    lib = abi.load()
    env = abi.make_env(flat.eye, 1, flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST)
    cap = int(lib.rt_accel_bytes(n))
    buf = np.zeros(cap, dtype=np.uint8)
    used = ctypes.c_int64(0)
    assert lib.rt_accel_build_host(flat.spheres.ctypes.data, n, ctypes.byref(env), buf.ctypes.data, cap, ctypes.byref(used)) == 0
    return _parse_accel(buf[: used.value].tobytes())


def _fma32(a, b, c):
    """float32 fma: the product of two float32 is exact in float64; one rounding to float64 of
    the sum, then to float32 (double rounding moves the result by at most one ulp, rarely)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def _slab_enter(o, d, lo, hi, extent, limit, inv_nudge):
    """rt_accel.cuh: ray_slabs + slab_enter for rays [k,3] against boxes [k,3] (float32 throughout)."""
    with np.errstate(all="ignore"):
        of = o.astype(f32)
        reach = np.abs(of).max(axis=1) + f32(extent)
        eps = (reach * f32(2.0 ** -20) + f32(1e-30)).astype(f32)[:, None]
        inv = (f32(1.0) / d.astype(f32)).astype(f32)
        inv = np.where(np.isfinite(inv), np.nextafter(inv, inv * f32(inv_nudge) * f32(np.inf)), inv)  # rcp.approx: +-1 ulp
        inv = np.fmin(np.fmax(inv, f32(-2.0 ** 64)), f32(2.0 ** 64))
        c_lo = (-(of + eps) * inv).astype(f32)
        c_hi = (-(of - eps) * inv).astype(f32)
        a = _fma32(lo, inv, c_lo)
        b = _fma32(hi, inv, c_hi)
        t_near = np.fmax(np.fmax.reduce(np.fmin(a, b), axis=1), f32(0))
        t_far = np.fmin(np.fmin.reduce(np.fmax(a, b), axis=1), limit)
    return t_near, ~(t_near > t_far)


def _sphere_t(o, d, c, r):
    """FP64 distance of the fast policy (half-b quadratic); nan where the ray misses."""
    L = o - c
    b = np.einsum("ij,ij->i", d, L)
    cc = np.einsum("ij,ij->i", L, L) - r * r
    disc = b * b - cc
    with np.errstate(invalid="ignore"):
        sq = np.sqrt(disc)
    t = np.where(-b - sq > 0, -b - sq, -b + sq)
    return np.where((disc > 0) & (t > 0), t, np.nan)


def _members(a):
    """(node index, shape index) for every shape below every node."""
    out = []

    def walk(node):
        ref = int(a["nodes"][node]["ref"])
        first, count = ref >> 3, ref & 7
        shapes = a["order"][first: first + count].tolist() if count else walk(first) + walk(first + 1)
        out.extend((node, s) for s in shapes)
        return shapes

    if len(a["nodes"]):
        walk(0)
    return np.array(out, dtype=np.int64)


@pytest.mark.parametrize("scale, spread, seed", [(1.0, 24.0, 1), (1e-3, 0.05, 2), (1e3, 3e4, 3), (1e5, 1e5, 4), (0.3, 1e4, 5)])
def test_fp32_slab_test_never_rejects_a_ray_that_hits(scale, spread, seed):
    rng = np.random.default_rng(seed)
    n = 96
    centres = rng.uniform(-spread, spread, (n, 3))
    radii = scale * rng.uniform(0.05, 1.0, n)
    centres[1] = centres[0]                      # coincident shapes
    centres[2] = centres[3] + [radii[3], 0, 0]   # overlapping
    a = _build(centres, radii)
    pairs = _members(a)
    assert len(pairs) > n  # every shape sits below several nodes
    node, shape = np.repeat(pairs[:, 0], 40), np.repeat(pairs[:, 1], 40)
    k = len(node)
    c, r = centres[shape], radii[shape]
    lo, hi = a["nodes"]["lo"][node], a["nodes"]["hi"][node]
    # aim at a point of the sphere: anywhere inside, or all but grazing it
    u = rng.normal(size=(k, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    frac = np.where(rng.random(k) < 0.5, rng.random(k), 1.0 - 10.0 ** rng.uniform(-12, -3, k))
    target = c + u * (r * frac)[:, None]
    # origins: far away, near, inside the node's box, exactly on one of its planes
    kind = rng.integers(0, 4, k)
    far = target + rng.normal(size=(k, 3)) * (spread * 3 + scale)
    near = target + rng.normal(size=(k, 3)) * (r * 3)[:, None]
    inside = lo + (hi - lo) * rng.random((k, 3))
    on_plane = inside.copy()
    axis = rng.integers(0, 3, k)
    on_plane[np.arange(k), axis] = np.where((rng.random(k) < 0.5)[:, None], lo, hi)[np.arange(k), axis]
    o = np.select([(kind == 0)[:, None], (kind == 1)[:, None], (kind == 2)[:, None]], [far, near, inside.astype(np.float64)], on_plane.astype(np.float64))
    # a fifth of the rays run exactly along an axis through the target
    par = rng.random(k) < 0.2
    pa = rng.integers(0, 3, k)
    o_par = target.copy()
    o_par[np.arange(k), pa] += rng.choice([-1.0, 1.0], k) * (r * rng.uniform(1.5, 50.0, k))
    o = np.where(par[:, None], o_par, o)
    d = target - o
    d[par] = np.where(np.arange(3)[None, :] == pa[par][:, None], np.sign(d[par]), 0.0)
    norm = np.linalg.norm(d, axis=1, keepdims=True)
    ok = norm[:, 0] > 0
    d = d / np.where(norm > 0, norm, 1.0)
    t = _sphere_t(o, d, c, r)
    hits = ok & np.isfinite(t)
    assert hits.mean() > 0.5  # the construction does produce hits
    o, d, lo, hi, t = o[hits], d[hits], lo[hits], hi[hits], t[hits]
    with np.errstate(over="ignore"):
        limit = np.nextafter(t.astype(f32), f32(np.inf))  # __double2float_ru at worst
    for nudge in (-1.0, 1.0):
        t_near, accepted = _slab_enter(o, d, lo, hi, a["extent"], limit, nudge)
        assert accepted.all(), f"{(~accepted).sum()} hitting rays rejected by a box above their sphere"
        assert (t_near <= limit).all()  # the entry distance never exceeds the hit distance
    # with no closer hit known (limit = +inf) the same boxes are of course accepted
    _, accepted = _slab_enter(o, d, lo, hi, a["extent"], f32(np.inf), 1.0)
    assert accepted.all()


def test_slab_test_does_reject_boxes_that_are_missed():
    """...and it is not vacuous: rays pointing away from a far box are turned away."""
    rng = np.random.default_rng(9)
    centres = rng.uniform(-20, 20, (64, 3))
    a = _build(centres, rng.uniform(0.05, 0.3, 64))
    leaves = np.flatnonzero(a["nodes"]["ref"] & 7)
    lo, hi = a["nodes"]["lo"][leaves], a["nodes"]["hi"][leaves]
    mid = 0.5 * (lo + hi)
    o = np.tile(np.array([[100.0, 100.0, 100.0]]), (len(leaves), 1))
    d = o - mid  # straight away from the box
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    _, accepted = _slab_enter(o, d, lo, hi, a["extent"], f32(np.inf), 1.0)
    assert not accepted.any()
