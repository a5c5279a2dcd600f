"""Parity of the CUDA path with the reference (golden fixtures) and the oracle.

Every test calls through the C ABI of include/rt_abi.h (CudaRenderer binds it with
ctypes; some tests call the host-buffer entry points directly).  Tolerances are the
north_star's: linear RGB within 1e-4 max-abs, 8-bit image within 1 LSB on >= 99.9 % of
pixels; integer/byte work (quantisation, counters in exact mode) is bit-exact.
"""

import ctypes as C
import os
import socket

import numpy as np
import pytest

from _cases import (FULL_FRAME_CASES, GOLDEN, SUBSET_CASES, assert_linear_parity, assert_png_parity, load_case,
                    manifest)

pytestmark = pytest.mark.gpu

PRECISIONS = ["exact", "fast"]
# exact mode evaluates the reference's expressions in its order without FMA: only pow/sin
# may differ in the last place
EXACT_TOL = 1e-10


@pytest.fixture(scope="module")
def rt():
    import torch

    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from python_ray_tracer_b200.infrastructure import CudaRenderer
    from python_ray_tracer_b200.infrastructure.cuda import abi

    abi.load()  # fails loudly if librt_b200.so is missing

    def make(precision, **kw):
        return CudaRenderer(precision=precision, collect_stats=True, **kw)

    return make


def _camera(case):
    from python_ray_tracer_b200 import Camera
    from python_ray_tracer_b200.infrastructure import NumpyVector3D

    return Camera(NumpyVector3D(*case.flat.eye), case.width, case.height)


def _render(renderer, case, rows=None):
    cam = _camera(case)
    row0, row1 = rows or (0, case.height)
    return renderer.render_rows(case.flat, cam, row0, row1, want_rgba8=True)


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", FULL_FRAME_CASES)
def test_full_frames_match_reference(rt, c_oracle, name, precision):
    case = load_case(name)
    renderer = rt(precision, max_depth=case.max_depth or None)
    color = _render(renderer, case)
    got = color.numpy()
    tol = EXACT_TOL if precision == "exact" else 1e-4
    assert_linear_parity(got, case.rgb, tol, f"{name}/{precision} vs reference fixture")
    want, stats = c_oracle.render(case.flat, case.width, case.height, max_depth=case.max_depth)
    assert_linear_parity(got, want, tol, f"{name}/{precision} vs oracle")
    counters = color.counters()
    if precision == "exact":
        assert counters["rays_traced"] == case.traced == int(stats[0])
        assert counters["generations"] == case.deepest
        assert counters["hits_shaded"] == int(stats[1])
    # the fused 8-bit output equals the truncating quantisation of the linear values
    got8 = color.rgba8.cpu().numpy()
    assert np.array_equal(got8[:, :3], c_oracle.quantize(got)) and (got8[:, 3] == 255).all()
    assert_png_parity(got8[:, :3], c_oracle.quantize(case.rgb), name)


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", SUBSET_CASES)
def test_subsets_match_reference(rt, c_oracle, name, precision):
    """Explicit-ray entry point (rt_trace_rays_dev) on the pixel subsets the reference
    rendered at 960x540 / 1080p / 4K (1025 shapes, depth 8) / 8K."""
    if not (GOLDEN / f"{name}.npz").exists():
        pytest.skip(f"{name} not generated")
    from python_ray_tracer_b200.infrastructure import NumpyVector3D

    case = load_case(name)
    renderer = rt(precision, max_depth=case.max_depth or None)
    d = c_oracle.ray_directions(case.flat.eye, case.width, case.height, case.pixels)
    color = renderer.trace_rays(case.flat, NumpyVector3D(*case.flat.eye), NumpyVector3D(d[0], d[1], d[2]))
    tol = EXACT_TOL if precision == "exact" else 1e-4
    assert_linear_parity(color.numpy(), case.rgb, tol, f"{name}/{precision}")


@pytest.mark.parametrize("precision", PRECISIONS)
def test_golden_png_960x540(rt, precision):
    """G0 (the reference's committed render.png) and C1 through render_image_pipeline."""
    import tempfile
    from pathlib import Path

    from python_ray_tracer_b200 import render_image_pipeline, scenes
    from python_ray_tracer_b200.infrastructure.png import read_png

    for builder, png, gains in ((scenes.main_scene, "g0_main_960x540.png", {}),
                                (scenes.readme_scene, "c1_readme_960x540.png", scenes.README_GAINS)):
        from python_ray_tracer_b200.infrastructure import CudaRenderer

        renderer = CudaRenderer(precision=precision, **gains)
        with tempfile.TemporaryDirectory() as tmp:
            path = Path(tmp) / "out.png"
            render_image_pipeline(builder(960, 540), path, renderer)
            got = read_png(path)
        want = read_png(GOLDEN / png)
        frac = assert_png_parity(got, want, png)
        if precision == "exact":
            assert np.count_nonzero(got != want) <= 8, np.count_nonzero(got != want)
        assert frac >= 0.9999


@pytest.mark.parametrize("precision", PRECISIONS)
def test_readme_1080p_and_8k_full_size(rt, c_oracle, precision):
    """BASELINE.json's sizes: full frames on the GPU, checked on the reference's pixel
    subsets and through size-independent properties (band == full frame, idempotence,
    traced-ray totals)."""
    man = manifest()
    for name in ("c2_readme_1080p_subset", "c4_readme_8k_subset"):
        case = load_case(name)
        renderer = rt(precision)
        color = _render(renderer, case)
        got = color.numpy()
        tol = EXACT_TOL if precision == "exact" else 1e-4
        assert_linear_parity(got[:, case.pixels], case.rgb, tol, f"{name}/{precision}")
        again = _render(renderer, case).numpy()
        assert np.array_equal(got, again)  # deterministic
        # a row band is bit-identical to the same rows of the full frame
        h, w = case.height, case.width
        for row0, row1 in ((0, 7), (h // 2 - 3, h // 2 + 10), (h - 5, h)):
            band = _render(renderer, case, rows=(row0, row1)).numpy()
            assert np.array_equal(band, got[:, row0 * w:row1 * w])
        assert np.isfinite(got).all() and got.min() >= 0.0
        if name.startswith("c2") and precision == "exact":
            assert color.counters()["rays_traced"] == 4244448  # SURVEY.md 8c pin (reference, 1080p)
    assert man["c1_readme_960x540_traced"] == 1060815


def test_exact_and_fast_agree_at_1080p(rt):
    case = load_case("c2_readme_1080p_subset")
    a = _render(rt("exact"), case).numpy()
    b = _render(rt("fast"), case).numpy()
    rep = assert_linear_parity(b, a, 1e-4, "fast vs exact 1080p")
    a8 = (255 * np.clip(a, 0, 1)).astype(np.uint8).T
    b8 = (255 * np.clip(b, 0, 1)).astype(np.uint8).T
    assert_png_parity(b8, a8, "fast vs exact 1080p")
    assert rep["max_abs"] < 1e-5  # the fast policy shades in single precision (values up to ~1.7)


def test_full_1080p_frame_against_the_oracle(rt, c_oracle):
    """BASELINE configs[1] at size, EVERY pixel: the README frame at 1920x1080 in both policies
    against oracle/librt_oracle.so (the C restatement, OpenMP over the box's cores), linear values
    and the fused 8-bit image; exact policy: identical ray / hit / depth counters."""
    case = load_case("c2_readme_1080p_subset")
    want, counters = c_oracle.render(case.flat, case.width, case.height)
    assert_linear_parity(want[:, case.pixels], case.rgb, 1e-12, "oracle vs the reference's own 1080p pixels")
    want8 = c_oracle.quantize(want)
    for precision in PRECISIONS:
        color = _render(rt(precision), case)
        got = color.numpy()
        rep = assert_linear_parity(got, want, EXACT_TOL if precision == "exact" else 1e-4, f"full 1080p frame/{precision}")
        assert rep["max_abs"] < (1e-12 if precision == "exact" else 1e-5), rep
        got8 = color.rgba8.cpu().numpy()[:, :3]
        assert_png_parity(got8, want8, f"full 1080p frame/{precision}")
        off = int((got8.astype(np.int16) != want8.astype(np.int16)).any(axis=1).sum())
        assert off <= (0 if precision == "exact" else 6_000), off  # bytes next to a rounding boundary (of 2 073 600 pixels)
        if precision == "exact":
            k = color.counters()
            assert (k["rays_traced"], k["hits_shaded"], k["generations"]) == (int(counters[0]), int(counters[1]), int(counters[2]))


def test_c3_1024_spheres_4k_band(rt, c_oracle):
    """Config 3 at size: a 4K band of the 1025-shape scene (depth 8) against the oracle."""
    case = load_case("c3_1024_4k_subset_depth8")
    rows = (1400, 1404)
    want, _ = c_oracle.render(case.flat, case.width, case.height, first=rows[0] * case.width,
                              count=(rows[1] - rows[0]) * case.width, max_depth=8)
    for precision in PRECISIONS:
        got = _render(rt(precision, max_depth=8), case, rows=rows).numpy()
        assert_linear_parity(got, want, EXACT_TOL if precision == "exact" else 1e-4, f"c3 band/{precision}")


def test_host_buffer_entry_points(c_oracle):
    """rt_render_frame_host / rt_trace_rays_host straight through ctypes with NumPy buffers."""
    from python_ray_tracer_b200.infrastructure.cuda import abi

    lib = abi.load()
    case = load_case("readme_160x90")
    flat, w, h = case.flat, case.width, case.height
    n = w * h
    env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_EXACT)
    cam = abi.make_camera(flat.eye, w, h)
    rgb = np.empty((3, n), dtype=np.float64)
    rgb8 = np.empty((n, 3), dtype=np.uint8)
    stats = np.zeros(4, dtype=np.uint64)
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib.rt_render_frame_host(0, ptr(flat.spheres), flat.n_spheres, ptr(flat.point_lights), C.byref(env),
                                  C.byref(cam), 0, h, ptr(rgb), abi.RT_DTYPE_F64, ptr(rgb8), ptr(stats))
    abi.check(rc, "rt_render_frame_host")
    assert_linear_parity(rgb, case.rgb, EXACT_TOL, "host entry")
    assert np.array_equal(rgb8, c_oracle.quantize(rgb)) and int(stats[0]) == case.traced
    # the 8-bit image alone (no linear planes are rendered), and beside float32 planes: same bytes
    only8 = np.zeros((n, 3), dtype=np.uint8)
    rc = lib.rt_render_frame_host(0, ptr(flat.spheres), flat.n_spheres, ptr(flat.point_lights), C.byref(env),
                                  C.byref(cam), 0, h, None, abi.RT_DTYPE_F32, ptr(only8), None)
    abi.check(rc, "rt_render_frame_host rgb8 only")
    assert np.array_equal(only8, rgb8)
    band8, band32 = np.zeros((5 * w, 3), dtype=np.uint8), np.empty((3, 5 * w), dtype=np.float32)
    rc = lib.rt_render_frame_host(0, ptr(flat.spheres), flat.n_spheres, ptr(flat.point_lights), C.byref(env),
                                  C.byref(cam), 41, 46, ptr(band32), abi.RT_DTYPE_F32, ptr(band8), None)
    abi.check(rc, "rt_render_frame_host band f32 + rgb8")
    assert np.array_equal(band8, rgb8[41 * w:46 * w]) and np.array_equal(band32, rgb[:, 41 * w:46 * w].astype(np.float32))
    # float32 output plane
    rgb32 = np.empty((3, n), dtype=np.float32)
    rc = lib.rt_render_frame_host(0, ptr(flat.spheres), flat.n_spheres, ptr(flat.point_lights), C.byref(env),
                                  C.byref(cam), 0, h, ptr(rgb32), abi.RT_DTYPE_F32, None, None)
    abi.check(rc, "rt_render_frame_host f32")
    assert np.array_equal(rgb32, rgb.astype(np.float32))
    # explicit rays with per-ray origins: bounce rays of the oracle's own first generation
    d = c_oracle.ray_directions(flat.eye, w, h)
    o = np.repeat(np.asarray(flat.eye)[:, None], n, axis=1).copy()
    o[0] += np.linspace(-0.1, 0.1, n)
    want, _ = c_oracle.trace_rays(flat, o, d)
    out = np.empty((3, n), dtype=np.float64)
    rc = lib.rt_trace_rays_host(0, ptr(flat.spheres), flat.n_spheres, ptr(flat.point_lights), C.byref(env), ptr(o),
                                None, ptr(d), n, ptr(out), abi.RT_DTYPE_F64, None)
    abi.check(rc, "rt_trace_rays_host")
    assert_linear_parity(out, want, EXACT_TOL, "host rays")
    # scenes of RT_ACCEL_AUTO_SPHERES shapes or more: the entry points build the tree themselves
    big = load_case("c3_1024_4k_subset_depth8")
    assert big.flat.n_spheres >= abi.RT_ACCEL_AUTO_SPHERES
    bcam = abi.make_camera(big.flat.eye, big.width, big.height)
    outs = []
    for flags in (abi.RT_FLAG_NONE, abi.RT_FLAG_NO_ACCEL):
        benv = abi.make_env(big.flat.eye, len(big.flat.point_lights), big.flat.dome_intensity, big.flat.dome_color,
                            max_depth=8, mode=abi.RT_MODE_FAST, flags=flags, cutoff=1e-7)
        band = np.empty((3, 8 * big.width), dtype=np.float64)
        st = np.zeros(4, dtype=np.uint64)
        abi.check(lib.rt_render_frame_host(0, ptr(big.flat.spheres), big.flat.n_spheres, ptr(big.flat.point_lights),
                                           C.byref(benv), C.byref(bcam), 1400, 1408, ptr(band), abi.RT_DTYPE_F64, None,
                                           ptr(st)), "rt_render_frame_host (auto tree)")
        db = c_oracle.ray_directions(big.flat.eye, big.width, big.height, big.pixels)
        rays = np.empty((3, db.shape[1]), dtype=np.float64)
        eye = (C.c_double * 3)(*big.flat.eye)
        abi.check(lib.rt_trace_rays_host(0, ptr(big.flat.spheres), big.flat.n_spheres, ptr(big.flat.point_lights),
                                         C.byref(benv), None, eye, ptr(db), db.shape[1], ptr(rays), abi.RT_DTYPE_F64,
                                         None), "rt_trace_rays_host (auto tree)")
        assert_linear_parity(rays, big.rgb, 1e-4, "host rays, 1025 shapes")
        outs.append((band, rays, st.copy()))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert tuple(outs[0][2][:3]) == tuple(outs[1][2][:3]) and outs[0][2][3] * 8 < outs[1][2][3]
    # bad arguments fail loudly
    assert lib.rt_render_frame_host(0, None, 3, None, C.byref(env), C.byref(cam), 0, h, None, 0, None, None) == -1
    assert lib.rt_render_frame_host(99, ptr(flat.spheres), flat.n_spheres, ptr(flat.point_lights), C.byref(env),
                                    C.byref(cam), 0, h, ptr(rgb), abi.RT_DTYPE_F64, None, None) == -3
    lib.rt_release_cache()


def test_intersect_and_ray_directions_bit_exact(rt):
    """Integer-free but IEEE-basic arithmetic: exact mode reproduces the reference bit for bit."""
    from python_ray_tracer_b200 import Camera
    from python_ray_tracer_b200.infrastructure import NumpySphere, NumpyVector3D
    from python_ray_tracer_b200.infrastructure.cuda.renderer import sphere_distance

    z = np.load(GOLDEN / "intersect_vectors.npz")
    o, d = z["origins"], z["directions"]
    for centre, radius, want in zip(z["centres"], z["radii"], z["distances"]):
        sphere = NumpySphere(NumpyVector3D(*centre), float(radius), None)
        got = sphere.intersect(NumpyVector3D(o[0], o[1], o[2]), NumpyVector3D(d[0], d[1], d[2]))
        assert np.array_equal(got, want)
        fast = sphere_distance(sphere, NumpyVector3D(o[0], o[1], o[2]), NumpyVector3D(d[0], d[1], d[2]), precision="fast")
        hit = want < 1e38
        assert np.array_equal(fast < 1e38, hit) or np.count_nonzero((fast < 1e38) != hit) <= 2
        both = hit & (fast < 1e38)
        assert np.abs(fast[both] - want[both]).max() <= 1e-9 * max(1.0, float(radius)) 
    unit = NumpySphere(NumpyVector3D(0, 0, 0), 1.0, None)
    assert unit.intersect(NumpyVector3D(0.0, 0.0, -3.0), NumpyVector3D(np.array([0.0]), np.array([0.0]), np.array([1.0])))[0] == 2.0
    g = np.load(GOLDEN / "ray_directions.npz")
    renderer = rt("exact")
    for key in g.files:
        w, h = (int(v) for v in key[2:].split("x"))
        rays = renderer.get_ray_directions(Camera(NumpyVector3D(0.0, 0.2, -2.0), w, h))
        got = np.stack([rays.x, rays.y, rays.z])
        if w * h > 1024:
            got = got[:, ::97]
        assert np.array_equal(got, g[key]), key


def test_edge_cases(rt, c_oracle):
    from python_ray_tracer_b200 import Camera, PointLight, Scene3D
    from python_ray_tracer_b200.infrastructure import NumpyRGBColor, NumpySphere, NumpyVector3D

    renderer = rt("exact")
    # a frame in which nothing is hit renders black (the reference raises in save_image)
    behind = Scene3D([NumpySphere(NumpyVector3D(0, 0, -50), 1.0, NumpyRGBColor(1, 1, 1))],
                     [PointLight(NumpyVector3D(0, 5, 0))], Camera(NumpyVector3D(0, 0.2, -2), 33, 17))
    color = renderer.raytrace_scene(behind.camera.position, renderer.get_ray_directions(behind.camera), behind)
    assert color.numpy().shape == (3, 33 * 17) and not color.numpy().any()
    assert color.counters()["hits_shaded"] == 0 and color.counters()["rays_traced"] == 33 * 17
    # empty ray batch
    empty = renderer.trace_rays(renderer.flatten(behind), NumpyVector3D(0.0, 0.0, 0.0),
                                NumpyVector3D(np.zeros(0), np.zeros(0), np.zeros(0)))
    assert empty.numpy().shape == (3, 0)
    # zero rows
    assert renderer.render_rows(renderer.flatten(behind), behind.camera, 5, 5).numpy().shape == (3, 0)
    # exact tie between two identical spheres: the reference adds both contributions (and recurses on both,
    # base.py:102-119).  The exact policy does the same (a lone sphere: twice the colour); the fast policy
    # shades the lowest index once -- documented deviation, see test_a_shape_listed_twice
    one = Scene3D([NumpySphere(NumpyVector3D(0, 0, 3), 1.0, NumpyRGBColor(1, 0, 0))],
                  [PointLight(NumpyVector3D(5, 10, -10))], Camera(NumpyVector3D(0, 0.2, -2), 16, 9))
    two = Scene3D(list(one.shapes) * 2, one.lights, one.camera)
    for precision in PRECISIONS:
        r = rt(precision)
        a = r.raytrace_scene(one.camera.position, r.get_ray_directions(one.camera), one).numpy()
        b = r.raytrace_scene(two.camera.position, r.get_ray_directions(two.camera), two).numpy()
        assert np.array_equal(b, a + a if precision == "exact" else a) and a.max() > 0
    # all-point-lights extension against its own CPU restatement (not the reference)
    case = load_case("c3_small_48_192x108_depth8")
    ext = rt("exact", max_depth=8, all_point_lights=True)
    got = _render(ext, case).numpy()
    want, _ = c_oracle.render(case.flat, case.width, case.height, max_depth=8, all_point_lights=True)
    assert_linear_parity(got, want, EXACT_TOL, "all point lights")
    assert np.abs(got - case.rgb).max() > 1e-3  # and it does change the picture
    # too many spheres for one block's shared memory and no tree -> loud error, no fallback
    from python_ray_tracer_b200.infrastructure.flatten import FlatScene

    big = FlatScene(np.ones((12, 20000)), case.flat.point_lights, case.flat.eye, [0.1], (1.0, 1.0, 1.0))
    with pytest.raises(RuntimeError):
        rt("exact", accel=False).render_rows(big, _camera(case), 0, 1)


def test_quantize_kernel_bit_exact(rt, c_oracle):
    rng = np.random.default_rng(3)
    v = np.concatenate([rng.uniform(-0.2, 1.8, (3, 5000)), np.array([[0.0, 1.0, 0.5, 1 / 255, 254.999 / 255]] * 3)], axis=1)
    from python_ray_tracer_b200.infrastructure import NumpyRGBColor

    renderer = rt("exact")
    got = renderer.quantize(NumpyRGBColor(v[0], v[1], v[2]), 3).cpu().numpy()
    assert np.array_equal(got, c_oracle.quantize(v))
    got4 = renderer.quantize(NumpyRGBColor(v[0], v[1], v[2]), 4).cpu().numpy()
    assert np.array_equal(got4[:, :3], got) and (got4[:, 3] == 255).all()


def test_strip_shards_reassemble_to_the_full_frame(rt):
    """Multi-GPU partition on one GPU: every rank's interleaved shard (rt_render_strips_dev)
    is rendered in turn and permuted back with the host's image_order(); the result is
    bit-identical to the single-launch frame (pixels are independent)."""
    from python_ray_tracer_b200.infrastructure.cuda import sharding

    for name, world, strip_rows in (("readme_160x90", 3, 4), ("c2_readme_1080p_subset", 8, 16),
                                    ("readme_37x23", 2, 16), ("readme_1x7", 4, 4)):
        case = load_case(name)
        for precision in PRECISIONS:
            renderer = rt(precision)
            cam = _camera(case)
            full = _render(renderer, case)
            want, want8 = full.numpy(), full.rgba8.cpu().numpy()
            w, h = case.width, case.height
            shards = sharding.strip_shards(h, world, strip_rows)
            max_rows = max(s.n_rows for s in shards)
            stacked = np.zeros((world * max_rows, w, 3))
            stacked8 = np.zeros((world * max_rows, w, 4), dtype=np.uint8)
            traced = 0
            for s in shards:
                color = renderer.render_strips(case.flat, cam, s, want_rgba8=True)
                stacked[s.rank * max_rows:s.rank * max_rows + s.n_rows] = color.numpy().T.reshape(s.n_rows, w, 3)
                stacked8[s.rank * max_rows:s.rank * max_rows + s.n_rows] = color.rgba8.cpu().numpy().reshape(s.n_rows, w, 4)
                traced += color.counters()["rays_traced"]
            order = sharding.image_order(h, world, strip_rows)
            assert np.array_equal(stacked[order].reshape(-1, 3).T, want), (name, precision)
            assert np.array_equal(stacked8[order].reshape(-1, 4), want8)
            assert traced == full.counters()["rays_traced"]


def test_image_order_shards_fill_one_framebuffer(rt):
    """RT_FLAG_IMAGE_ORDER: every rank's strips are stored at their image positions of ONE
    whole-frame RGBA8 buffer (what the peer-memory exchange does across GPUs); after all
    shards ran the buffer equals the single-launch frame, byte for byte."""
    import torch

    from python_ray_tracer_b200.infrastructure.cuda import sharding

    for name, world, strip_rows in (("readme_160x90", 3, 4), ("readme_37x23", 2, 16), ("c2_readme_1080p_subset", 8, 16)):
        case = load_case(name)
        renderer = rt("fast")
        cam = _camera(case)
        want8 = _render(renderer, case).rgba8.cpu().numpy().reshape(case.height, case.width, 4)
        frame = torch.zeros((case.height, case.width, 4), dtype=torch.uint8, device="cuda")
        for s in sharding.strip_shards(case.height, world, strip_rows):
            renderer.render_strips_into(case.flat, cam, s, frame.data_ptr())
        assert np.array_equal(frame.cpu().numpy(), want8), name


def test_wide_stores_and_rgb8_keep_every_pixel(rt):
    """RT_FLAG_WIDE_STORES (32x1 patches, 128-byte stores) and RT_FLAG_RGB8 (3-byte pixels,
    packed into words by the warp where the width allows, byte stores elsewhere) change how a
    frame is stored, never what is stored: linear planes bit-identical to the default launch,
    8-bit pixels equal to its RGBA8 without the A; RT_FLAG_EVEN_STORES only reorders the patches.  Widths that are and are not multiples of
    4 and 32, bands, strips, image order, with and without the sphere tree."""
    import torch

    from python_ray_tracer_b200.infrastructure.cuda import abi, sharding

    wide, rgb8, even = abi.RT_FLAG_WIDE_STORES, abi.RT_FLAG_RGB8, abi.RT_FLAG_EVEN_STORES
    for name in ("readme_160x90", "readme_37x23", "readme_1x7", "readme_9x1", "c2_readme_1080p_subset", "c3_small_48_192x108_depth8"):
        case = load_case(name)
        cam = _camera(case)
        w, h = case.width, case.height
        for precision in PRECISIONS:
            renderer = rt(precision)
            full = _render(renderer, case)
            want, want8 = full.numpy(), full.rgba8.cpu().numpy()
            for flags in (wide, rgb8, wide | rgb8, wide | even, wide | rgb8 | even):
                got = renderer.render_rows(case.flat, cam, 0, h, want_rgba8=True, flags=flags)
                assert np.array_equal(got.numpy(), want), (name, precision, flags)
                got8 = got.rgba8.cpu().numpy()
                assert got8.shape == (w * h, 3 if flags & rgb8 else 4)
                assert np.array_equal(got8[:, :3], want8[:, :3]), (name, precision, flags)
                assert got.counters() == full.counters()
            if h >= 8:  # a band that starts and ends off the patch grid
                r0, r1 = 3, h - 2
                band = renderer.render_rows(case.flat, cam, r0, r1, want_rgba8=True, flags=wide | rgb8)
                assert np.array_equal(band.numpy(), want[:, r0 * w:r1 * w])
                assert np.array_equal(band.rgba8.cpu().numpy(), want8[r0 * w:r1 * w, :3])
        # image order, 3 shards into one RGB8 frame whose bytes around it must stay untouched
        renderer = rt("fast")
        want8 = _render(renderer, case).rgba8.cpu().numpy().reshape(h, w, 4)[:, :, :3]
        guard = 64
        buf = torch.full((guard + h * w * 3 + guard,), 0xA5, dtype=torch.uint8, device="cuda")
        for s in sharding.strip_shards(h, 3, 4):
            renderer.render_strips_into(case.flat, cam, s, buf.data_ptr() + guard, flags=wide | rgb8)
        host = buf.cpu().numpy()
        assert (host[:guard] == 0xA5).all() and (host[-guard:] == 0xA5).all()
        assert np.array_equal(host[guard:-guard].reshape(h, w, 3), want8), name
    # the tree kernels have their own instantiations of the wide patch
    case = load_case("readme_160x90")
    tree = rt("fast", accel=True)
    plain = _render(tree, case)
    got = tree.render_rows(case.flat, _camera(case), 0, case.height, want_rgba8=True, flags=wide | rgb8)
    assert np.array_equal(got.numpy(), plain.numpy())
    assert np.array_equal(got.rgba8.cpu().numpy(), plain.rgba8.cpu().numpy()[:, :3])


def _peer_writer(rank, world, port, name, strip_rows, out):
    import torch
    import torch.distributed as dist

    from python_ray_tracer_b200.infrastructure import CudaRenderer
    from python_ray_tracer_b200.infrastructure.cuda import sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = load_case(name)
        renderer = CudaRenderer(precision="fast")
        cam = _camera(case)
        peer = sharding.PeerFrame(case.height, case.width, rank, world, 0, dst=0)
        shard = sharding.StripShard(case.height, world, rank, strip_rows)
        renderer.render_strips_into(case.flat, cam, shard, peer.ptr)
        frame = peer.complete()
        if rank == 0:
            out.put(frame.cpu().numpy())
        peer.close()
    finally:
        dist.destroy_process_group()


def test_peer_framebuffer_two_processes(rt):
    """The exchange step of config 4 with one process per rank: rank 0 owns the framebuffer,
    rank 1 maps it through the CUDA IPC handle and its kernel stores into it.  (Both
    processes use cuda:0 here; across GPUs the same mapping goes over NVLink.)"""
    import torch.multiprocessing as mp

    name, world, strip_rows = "readme_160x90", 2, 8
    case = load_case(name)
    want8 = _render(rt("fast"), case).rgba8.cpu().numpy().reshape(case.height, case.width, 4)
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=_peer_writer, args=(r, world, port, name, strip_rows, out)) for r in range(world)]
    for p in procs:
        p.start()
    frame = out.get(timeout=240)
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    assert np.array_equal(frame, want8)


@pytest.mark.parametrize("encoder", ["device", "host"])
def test_turntable_batch_writes_reference_frames(rt, c_oracle, tmp_path, encoder):
    """Config 5 at small size: the batch driver (PNGs made on the device, or render || copy || parallel host
    encode) writes the frames this rank owns; frame 37 decodes to the reference's fixture within 1 LSB, a frame
    left half-written by a killed run is rendered again, a second call resumes (writes nothing)."""
    from python_ray_tracer_b200 import scenes, turntable
    from python_ray_tracer_b200.infrastructure import CudaRenderer
    from python_ray_tracer_b200.infrastructure.png import read_png

    case = load_case("turntable37_160x90")
    renderer = CudaRenderer(**scenes.README_GAINS, precision="fast")
    turntable.frame_path(tmp_path, 101).with_suffix(".png.tmp").write_bytes(b"\x89PNG cut short")
    stats = turntable.render_turntable(renderer, 256, tmp_path, 160, 90, rank=37, world=64, encoder_threads=4, encoder=encoder)
    assert stats["frames"] == 4 and stats["skipped"] == 0 and not list(tmp_path.glob("*.tmp"))
    assert sorted(p.name for p in tmp_path.glob("*.png")) == [f"frame_{k:04d}.png" for k in (37, 101, 165, 229)]
    got = read_png(turntable.frame_path(tmp_path, 37))
    assert got.shape == (90, 160, 3)
    assert_png_parity(got, c_oracle.quantize(case.rgb), "turntable frame 37")
    again = turntable.render_turntable(renderer, 256, tmp_path, 160, 90, rank=37, world=64, encoder_threads=4, encoder=encoder)
    assert again["frames"] == 0 and again["skipped"] == 4


@pytest.mark.parametrize("precision", PRECISIONS)
def test_bvh_gives_the_same_picture_as_testing_every_sphere(precision, c_oracle):
    """rt_env.accel_dev: the tree only skips sphere tests that miss, so frames, 8-bit images
    and ray / hit / depth counters are identical to the exhaustive kernels', bit for bit --
    on the 49- and 1025-shape scenes, on every stress scene (forced on tiny scenes too), with
    explicit rays, strips and the all-lights extension -- and equal to the reference fixtures."""
    from python_ray_tracer_b200.infrastructure import CudaRenderer, NumpyVector3D
    from python_ray_tracer_b200.infrastructure.cuda import sharding

    def pair(**kw):
        return (CudaRenderer(precision=precision, collect_stats=True, accel=False, **kw),
                CudaRenderer(precision=precision, collect_stats=True, accel=True, **kw))

    tol = EXACT_TOL if precision == "exact" else 1e-4
    names = ["c3_small_48_192x108", "c3_small_48_192x108_depth8", "c3_small_48_192x108_depth1", "readme_160x90",
             "main_160x90", "readme_1x7"] + [f"{n}_48x27" for n in __import__("_cases").STRESS_CASES]
    for name in names:
        case = load_case(name)
        plain, tree = pair(max_depth=case.max_depth or None)
        a, b = _render(plain, case), _render(tree, case)
        assert np.array_equal(a.numpy(), b.numpy()), name
        assert np.array_equal(a.rgba8.cpu().numpy(), b.rgba8.cpu().numpy()), name
        ca, cb = a.counters(), b.counters()
        assert (ca["rays_traced"], ca["hits_shaded"], ca["generations"]) == (cb["rays_traced"], cb["hits_shaded"], cb["generations"]), name
        assert cb["sphere_tests"] <= ca["sphere_tests"] + ca["rays_traced"] + ca["hits_shaded"], name
        assert_linear_parity(b.numpy(), case.rgb, tol, f"{name}/{precision} with the tree vs reference fixture")
    # config 3 at size: 1025 shapes, 4K, depth 8 -- a band, the reference's pixel subset, strips
    case = load_case("c3_1024_4k_subset_depth8")
    plain, tree = pair(max_depth=8)
    a, b = _render(plain, case, rows=(1400, 1408)), _render(tree, case, rows=(1400, 1408))
    assert np.array_equal(a.numpy(), b.numpy())
    assert b.counters()["sphere_tests"] * 8 < a.counters()["sphere_tests"]  # and it does skip most of them
    d = c_oracle.ray_directions(case.flat.eye, case.width, case.height, case.pixels)
    rays = tree.trace_rays(case.flat, NumpyVector3D(*case.flat.eye), NumpyVector3D(d[0], d[1], d[2]))
    assert_linear_parity(rays.numpy(), case.rgb, tol, f"c3 subset/{precision} with the tree")
    shard = sharding.StripShard(case.height, 64, 5, 16)
    assert np.array_equal(plain.render_strips(case.flat, _camera(case), shard).numpy(),
                          tree.render_strips(case.flat, _camera(case), shard).numpy())
    # rays exactly parallel to an axis (1/d = inf on two axes), from outside and inside the boxes,
    # through and beside every sphere: the degenerate inputs of the slab test
    case = load_case("c3_small_48_192x108_depth8")
    centres, radii = case.flat.spheres[0:3, :-1].T, case.flat.spheres[3, :-1]
    origins, directions = [], []
    for axis in range(3):
        for sign in (-1.0, 1.0):
            for shift, aside in ((30.0, 0.0), (30.0, 0.999), (30.0, 1.5), (0.5, 0.3)):
                o = centres.copy()
                o[:, axis] -= sign * shift * np.maximum(radii, 0.2)
                o[:, (axis + 1) % 3] += aside * radii
                d = np.zeros_like(o)
                d[:, axis] = sign
                origins.append(o)
                directions.append(d)
    o, d = np.concatenate(origins).T.copy(), np.concatenate(directions).T.copy()
    plain, tree = pair(max_depth=8)
    a = plain.trace_rays(case.flat, NumpyVector3D(o[0], o[1], o[2]), NumpyVector3D(d[0], d[1], d[2]))
    b = tree.trace_rays(case.flat, NumpyVector3D(o[0], o[1], o[2]), NumpyVector3D(d[0], d[1], d[2]))
    assert np.array_equal(a.numpy(), b.numpy()) and a.counters()["hits_shaded"] == b.counters()["hits_shaded"] > o.shape[1] // 2
    # the all-lights extension walks the tree once per light
    plain, tree = pair(max_depth=8, all_point_lights=True)
    assert np.array_equal(_render(plain, case).numpy(), _render(tree, case).numpy())


def test_a_shape_listed_twice(rt, c_oracle):
    """Ties at the nearest distance, pinned by a reference-rendered fixture (tests/golden/make_stress.py:
    tie_case): the reference lets every shape at exactly the nearest distance contribute and recurse
    (base.py:102-119), so a duplicated sphere comes out twice as bright.  The exact policy (and the C oracle)
    does the same, with and without the sphere tree, ray totals included.  The fast policy shades the lowest
    index once -- the one documented deviation: it renders the reference's picture of the scene WITHOUT the
    duplicate."""
    z = np.load(GOLDEN / "tie_duplicate_48x27.npz")
    case = load_case("tie_duplicate_48x27")
    seen_twice = (np.abs(z["rgb"] - z["rgb_single"]) > 1e-12).any(axis=0)
    assert 50 < int(seen_twice.sum()) < case.width * case.height // 2
    assert (z["rgb"][:, seen_twice].sum(axis=0) > z["rgb_single"][:, seen_twice].sum(axis=0)).all()  # the reference adds both
    want, counters = c_oracle.render(case.flat, case.width, case.height, max_depth=4)
    assert_linear_parity(want, z["rgb"], 1e-12, "oracle vs reference on the duplicated scene")
    for accel in (False, True):
        color = _render(rt("exact", max_depth=4, accel=accel), case)
        assert_linear_parity(color.numpy(), z["rgb"], EXACT_TOL, f"duplicate sphere/exact/tree={accel}")
        k = color.counters()
        assert (k["rays_traced"], k["hits_shaded"]) == (case.traced, int(counters[1])), (k, case.traced)
        fast = _render(rt("fast", max_depth=4, accel=accel), case).numpy()
        assert_linear_parity(fast, z["rgb_single"], 1e-4, f"duplicate sphere/fast/tree={accel}")
        assert_linear_parity(fast[:, ~seen_twice], z["rgb"][:, ~seen_twice], 1e-4, "pixels without a tie")


def test_camera_rays_that_do_not_start_at_the_scene_camera(rt, c_oracle):
    """``raytrace_scene`` handed ``CameraRays`` of a camera that is NOT ``scene.camera``: the reference shades
    "towards scene.camera.position" at every generation (shader.py:76) while the rays leave the other camera.
    The camera kernel takes that direction from the ray, so such calls are traced as explicit rays instead
    (the C entry point refuses them): fast and exact agree with each other and with the oracle."""
    from python_ray_tracer_b200 import Camera, scenes
    from python_ray_tracer_b200.infrastructure import NumpyVector3D
    from python_ray_tracer_b200.infrastructure.cuda import abi

    scene = scenes.readme_scene(96, 54)
    other = Camera(NumpyVector3D(0.4, 0.5, -1.5), 96, 54)
    got = {}
    for precision in PRECISIONS:
        renderer = rt(precision)
        rays = renderer.get_ray_directions(other)
        got[precision] = renderer.raytrace_scene(other.position, rays, scene).numpy()
    flat = rt("exact").flatten(scene)
    d = c_oracle.ray_directions((0.4, 0.5, -1.5), 96, 54)
    want, _ = c_oracle.trace_rays(flat, (0.4, 0.5, -1.5), d)
    assert_linear_parity(got["exact"], want, EXACT_TOL, "rays from another camera / exact")
    assert_linear_parity(got["fast"], want, 1e-4, "rays from another camera / fast")
    # the C entry point: camera rays must start at env->eye
    lib = abi.load()
    env = rt("fast")._plain_env(flat)
    cam = abi.make_camera((0.4, 0.5, -1.5), 96, 54)
    assert lib.rt_render_rows_dev(0, 8, flat.n_spheres, 8, C.byref(env), C.byref(cam), 0, 54, None, 0, 0, None, None, None) == -1
    assert b"env->eye" in lib.rt_last_error()


def test_a_blob_of_another_scene_fails_loudly(rt):
    """rt_env.accel_dev must be the blob of THIS sphere table: a launch given another scene's tree renders
    nothing and says so -- through rt_device_status() and through the next device call -- instead of
    quietly mixing two scenes (materials from the table, geometry from the blob)."""
    import torch

    from python_ray_tracer_b200.infrastructure.cuda import abi

    lib = abi.load()
    case, other = load_case("c3_small_48_192x108"), load_case("stress_random20_48x27")
    renderer = rt("fast", accel=True)
    good = _render(renderer, case).numpy()  # builds and caches the tree of `case`
    assert lib.rt_device_status(0) == 0
    spheres, lights = (torch.from_numpy(a).cuda() for a in (other.flat.spheres, other.flat.point_lights))
    env = renderer._plain_env(other.flat)
    env.accel_dev, env.accel_bytes = renderer._accel_dev.data_ptr(), renderer._accel_dev.numel()  # the WRONG blob
    cam = abi.make_camera(other.flat.eye, other.width, other.height)
    n = other.width * other.height
    planes = torch.full((3, n), -1.0, dtype=torch.float64, device="cuda")
    call = lambda: lib.rt_render_rows_dev(0, spheres.data_ptr(), other.flat.n_spheres, lights.data_ptr(), C.byref(env), C.byref(cam),  # noqa: E731
                                          0, other.height, planes.data_ptr(), abi.RT_DTYPE_F64, n, None, None, None)
    assert call() == 0  # enqueued: the mismatch is only visible on the device
    torch.cuda.synchronize()
    assert (planes == -1.0).all()  # nothing was rendered
    assert lib.rt_device_status(0) == -4 and b"not a blob" in lib.rt_last_error()
    assert lib.rt_device_status(0) == 0  # reading it clears it
    assert call() == 0
    torch.cuda.synchronize()
    assert call() == -4 and b"not a blob" in lib.rt_last_error()  # ... and the next device call refuses
    assert np.array_equal(_render(renderer, case).numpy(), good)


def test_bvh_lifts_the_shared_memory_limit(rt, c_oracle):
    """4 000 spheres do not fit one block's shared memory: without a tree the call is refused,
    with one (the default for large scenes) it renders and matches the oracle."""
    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure import CudaRenderer
    from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene

    scene = scenes.random_spheres_scene(96, 54, n_spheres=4000, seed=5)
    flat = flatten_scene(scene, DefaultMaterial(**scenes.README_GAINS))
    with pytest.raises(RuntimeError):
        CudaRenderer(precision="fast", accel=False, max_depth=4).render_rows(flat, scene.camera, 0, 54)
    got = CudaRenderer(precision="fast", max_depth=4).render_rows(flat, scene.camera, 0, 54).numpy()
    want, _ = c_oracle.render(flat, 96, 54, max_depth=4)
    assert_linear_parity(got, want, 1e-4, "4001 shapes through the tree")


def test_cli_renders_a_scene_file_like_the_programmatic_scene(tmp_path):
    """N4: `python -m python_ray_tracer_b200 render` on an exported scene file writes the PNG
    that render_image_pipeline writes for the scene built in Python, byte for byte in pixels."""
    from python_ray_tracer_b200 import render_image_pipeline, scene_io, scenes
    from python_ray_tracer_b200.__main__ import main
    from python_ray_tracer_b200.infrastructure import CudaRenderer
    from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial
    from python_ray_tracer_b200.infrastructure.png import read_png

    scene = scenes.random_spheres_scene(96, 54, n_spheres=40, seed=9)
    settings = scene_io.RenderSettings(max_specular_depth=4, output_path=str(tmp_path / "from_file.png"))
    scene_io.save_scene(scene, tmp_path / "scene.json", tmp_path / "settings.json", settings,
                        default=DefaultMaterial(**scenes.README_GAINS))
    assert main(["render", "--scene", str(tmp_path / "scene.json"), "--settings", str(tmp_path / "settings.json")]) == 0
    render_image_pipeline(scene, tmp_path / "direct.png", CudaRenderer(**scenes.README_GAINS, max_depth=4))
    a, b = read_png(tmp_path / "from_file.png"), read_png(tmp_path / "direct.png")
    assert a.shape == (54, 96, 3) and np.array_equal(a, b) and a.max() > 0


@pytest.mark.parametrize("scale, spread, n_spheres, seed", [(1.0, 12.0, 33, 1), (1e-3, 0.03, 200, 2), (50.0, 4e4, 700, 3),
                                                             (0.2, 3.0, 1300, 4), (1e4, 1e7, 64, 5)])
def test_bvh_equals_exhaustive_on_random_scenes_and_rays(scale, spread, n_spheres, seed):
    """Random scenes from millimetres to 1e7 units (coincident, nested and touching spheres
    included) and rays chosen to stress the FP32 slab test -- origins inside spheres, on
    surfaces and far away; directions with components of 0, 1e-20, 1e-45 (below FP32) and
    1e-300 -- give bit-identical colours and equal counters with and without the tree."""
    from python_ray_tracer_b200.infrastructure import CudaRenderer, NumpyVector3D
    from python_ray_tracer_b200.infrastructure.flatten import FlatScene

    rng = np.random.default_rng(seed)
    centres = rng.uniform(-spread, spread, (n_spheres, 3))
    radii = scale * rng.uniform(0.05, 1.0, n_spheres)
    centres[1], radii[1] = centres[0], radii[0]                    # the same sphere twice (tie -> lowest index)
    centres[2], radii[2] = centres[3], 0.5 * radii[3]              # nested
    centres[4] = centres[5] + [radii[4] + radii[5], 0.0, 0.0]      # touching
    table = np.zeros((12, n_spheres))
    table[0:3], table[3] = centres.T, radii
    table[4], table[5], table[6], table[7] = 1.0, rng.uniform(0.2, 1.0, n_spheres), 0.5, 0.05
    table[8:11] = rng.uniform(0.1, 1.0, (3, n_spheres))
    table[11, ::7] = 1.0
    flat = FlatScene(np.ascontiguousarray(table), np.array([[0.3 * spread, 2.0 * spread, -1.5 * spread]]),
                     (0.0, 0.0, -2.0 * spread), [0.1], (1.0, 1.0, 1.0))
    k = 20000
    target = centres[rng.integers(0, n_spheres, k)] + rng.normal(size=(k, 3)) * (radii[rng.integers(0, n_spheres, k)] * 0.7)[:, None]
    kind = rng.integers(0, 4, k)
    u = rng.normal(size=(k, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    which = rng.integers(0, n_spheres, k)
    o = np.select([(kind == 0)[:, None], (kind == 1)[:, None], (kind == 2)[:, None]],
                  [target + rng.normal(size=(k, 3)) * spread * 3,               # far away
                   centres[which] + u * (radii[which] * rng.random(k))[:, None],  # inside a sphere
                   centres[which] + u * radii[which][:, None]],                # on a surface
                  target + rng.normal(size=(k, 3)) * scale * 2)                # close by
    d = target - o
    tiny = rng.random(k) < 0.4  # squash one or two components to (almost) nothing
    comp = rng.integers(0, 3, k)
    value = rng.choice([0.0, 1e-20, -1e-20, 1e-45, -1e-45, 1e-300], k)
    d[tiny, comp[tiny]] = value[tiny] * np.abs(d[tiny]).max(axis=1)
    two = tiny & (rng.random(k) < 0.5)
    d[two, (comp[two] + 1) % 3] = 0.0
    norm = np.linalg.norm(d, axis=1, keepdims=True)
    d = d / np.where(norm > 0, norm, 1.0)
    d[norm[:, 0] == 0] = [0.0, 0.0, 1.0]
    for precision in PRECISIONS:
        plain = CudaRenderer(precision=precision, accel=False, max_depth=3, collect_stats=True)
        tree = CudaRenderer(precision=precision, accel=True, max_depth=3, collect_stats=True)
        if n_spheres > 1000 and precision == "exact":
            continue  # (time: the exhaustive exact kernel on 1 300 spheres x 20 000 rays x 3 generations)
        a = plain.trace_rays(flat, NumpyVector3D(o[:, 0], o[:, 1], o[:, 2]), NumpyVector3D(d[:, 0], d[:, 1], d[:, 2]))
        b = tree.trace_rays(flat, NumpyVector3D(o[:, 0], o[:, 1], o[:, 2]), NumpyVector3D(d[:, 0], d[:, 1], d[:, 2]))
        A, B = a.numpy(), b.numpy()
        same = (A == B) | (np.isnan(A) & np.isnan(B))
        assert same.all(), f"{precision}: {(~same.all(axis=0)).sum()} of {k} rays differ"
        ca, cb = a.counters(), b.counters()
        assert (ca["rays_traced"], ca["hits_shaded"]) == (cb["rays_traced"], cb["hits_shaded"]) and ca["hits_shaded"] > k // 4
        assert cb["sphere_tests"] < ca["sphere_tests"]
