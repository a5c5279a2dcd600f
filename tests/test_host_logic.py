"""Host side of the boundary: the Renderer port mirror, both constructor surfaces, the
scene flattener, the PNG container and the C-ABI library's symbol table (no compute)."""

import ctypes
import re
import sys
from pathlib import Path

import numpy as np
import pytest

from python_ray_tracer_b200 import (Camera, DomeLight, PointLight, Renderer, Scene3D, render_image_pipeline, scenes)
from python_ray_tracer_b200.infrastructure import (FARAWAY, CheckeredSphere, NumpyRGBColor, NumpyShader, NumpySphere,
                                                   NumpyVector3D, Texture, TextureChecker)
from python_ray_tracer_b200.infrastructure.cuda import abi
from python_ray_tracer_b200.infrastructure.flatten import (F_COLOR_R, F_CX, F_DIFFUSE_GAIN, F_RADIUS, F_SPECULAR_GAIN,
                                                           F_SPECULAR_ROUGHNESS, F_TEXTURE, DefaultMaterial,
                                                           flatten_scene)
from python_ray_tracer_b200.infrastructure.png import encode_png, read_png, write_png

ROOT = Path(__file__).resolve().parents[1]


def test_port_has_the_reference_methods():
    # application.py:7-32: exactly these three abstract methods
    assert Renderer.__abstractmethods__ == {"raytrace_scene", "get_ray_directions", "save_image"}


def test_pipeline_calls_in_reference_order(tmp_path):
    calls = []

    class Spy(Renderer):
        def get_ray_directions(self, camera):
            calls.append(("dirs", camera.width))
            return "D"

        def raytrace_scene(self, origin, directions, scene):
            calls.append(("trace", origin, directions))
            return "C"

        def save_image(self, color, camera, path):
            calls.append(("save", color, path))

    scene = scenes.readme_scene(8, 4)
    render_image_pipeline(scene, tmp_path / "x.png", Spy())
    assert calls == [("dirs", 8), ("trace", scene.camera.position, "D"), ("save", "C", tmp_path / "x.png")]


def test_readme_and_head_surfaces_flatten_identically():
    default = DefaultMaterial(**scenes.README_GAINS)
    a = flatten_scene(scenes.readme_scene(64, 36), default)
    b = flatten_scene(scenes.readme_scene_head_api(64, 36))
    assert np.array_equal(a.spheres, b.spheres)
    assert np.array_equal(a.point_lights, b.point_lights)
    assert a.dome_intensity == b.dome_intensity == [0.1] and a.dome_color == (1.0, 1.0, 1.0)
    assert a.spheres[F_TEXTURE].tolist() == [0.0, 0.0, 1.0]
    assert a.spheres[F_CX, 0] == 0.55 and a.spheres[F_RADIUS, 2] == 99999.0
    assert a.spheres[F_SPECULAR_ROUGHNESS].tolist() == [0.5, 0.5, 0.5]
    assert a.eye == (0.0, 0.2, -2.0)


def test_renderer_gains_are_the_default_material_only_for_colour_shapes():
    shader = NumpyShader(0.0, 0.25, 0.1, 0.0, 0.5, Texture(NumpyRGBColor(1, 0, 0)))
    scene = Scene3D(
        [NumpySphere(NumpyVector3D(0, 0, 3), 1.0, shader), NumpySphere(NumpyVector3D(1, 0, 3), 1.0, NumpyRGBColor(0, 1, 0))],
        [PointLight(NumpyVector3D(0, 5, 0))],
        Camera(NumpyVector3D(0, 0, -1), 4, 4),
    )
    flat = flatten_scene(scene, DefaultMaterial(specular_gain=0.75, diffuse_gain=0.3))
    assert flat.spheres[F_SPECULAR_GAIN].tolist() == [0.25, 0.75]
    assert flat.spheres[F_DIFFUSE_GAIN].tolist() == [0.5, 0.3]
    assert flat.spheres[F_COLOR_R].tolist() == [1.0, 0.0]
    assert flat.dome_intensity == [] and flat.dome_color == (1.0, 1.0, 1.0)


def test_checker_ignores_its_colour_and_extra_argument():
    s = CheckeredSphere(NumpyVector3D(0, -99999.5, 0), 99999, NumpyRGBColor(0.2, 0.3, 0.4), 1.0)
    scene = Scene3D([s], [PointLight(NumpyVector3D(0, 5, 0))], Camera(NumpyVector3D(0, 0, -1), 4, 4))
    flat = flatten_scene(scene)
    assert flat.spheres[F_TEXTURE, 0] == 1.0 and flat.spheres[F_COLOR_R, 0] == 1.0
    assert isinstance(TextureChecker().color, NumpyRGBColor)


def test_lights_rules():
    sphere = NumpySphere(NumpyVector3D(0, 0, 3), 1.0, NumpyRGBColor(1, 1, 1))
    cam = Camera(NumpyVector3D(0, 0, -1), 4, 4)
    dome = DomeLight(0.1, NumpyRGBColor(1, 1, 1))
    with pytest.raises(ValueError):  # shader.py:75 raises AttributeError in the reference
        flatten_scene(Scene3D([sphere], [dome, PointLight(NumpyVector3D(0, 5, 0))], cam))
    with pytest.raises(ValueError):
        flatten_scene(Scene3D([sphere], [], cam))
    with pytest.raises(ValueError):
        flatten_scene(Scene3D([], [PointLight(NumpyVector3D(0, 5, 0))], cam))
    many = [PointLight(NumpyVector3D(0, 5, 0))] + [DomeLight(0.01 * i, NumpyRGBColor(i, 0, 0)) for i in range(1, 12)]
    flat = flatten_scene(Scene3D([sphere], many, cam))
    assert len(flat.dome_intensity) == 8 and flat.dome_color == (11.0, 0.0, 0.0)  # colour of the LAST dome
    assert abs(sum(flat.dome_intensity) - sum(0.01 * i for i in range(1, 12))) < 1e-15
    flat = flatten_scene(scenes.random_spheres_scene(16, 9, n_spheres=5))
    assert flat.point_lights.shape == (4, 3) and flat.n_spheres == 6


def test_zero_radius_and_non_sphere_rejected():
    cam = Camera(NumpyVector3D(0, 0, -1), 4, 4)
    light = [PointLight(NumpyVector3D(0, 5, 0))]
    with pytest.raises(ValueError):
        flatten_scene(Scene3D([NumpySphere(NumpyVector3D(0, 0, 3), 0.0, NumpyRGBColor(1, 1, 1))], light, cam))
    with pytest.raises(ValueError):
        flatten_scene(Scene3D([object()], light, cam))


def test_vector_value_types_match_reference_semantics():
    a, b = NumpyVector3D(1.0, 2.0, 3.0), NumpyVector3D(4.0, 5.0, 6.0)
    assert a.dot(b) == 32.0 and abs(a) == 14.0  # abs is the SQUARED length (base.py:37-38)
    assert (a + b).components() == (5.0, 7.0, 9.0) and (b - a).components() == (3.0, 3.0, 3.0)
    assert (a * 2).components() == (2.0, 4.0, 6.0) and (a / 2).components() == (0.5, 1.0, 1.5)
    n = NumpyVector3D(3.0, 0.0, 4.0).norm()
    assert np.allclose(n.components(), (0.6, 0.0, 0.8))
    assert NumpyVector3D(0.0, 0.0, 0.0).norm().components() == (0.0, 0.0, 0.0)  # zero-safe (base.py:61-64)
    v = NumpyVector3D(np.arange(4.0), np.arange(4.0) * 2, 7.0)
    mask = np.array([True, False, True, False])
    e = v.extract(mask)
    assert e.x.tolist() == [0.0, 2.0] and e.z == 7.0
    p = e.place(mask)
    assert p.x.tolist() == [0.0, 0.0, 2.0, 0.0] and p.z.tolist() == [7.0, 0.0, 7.0, 0.0]
    assert FARAWAY == 1.0e39


def test_png_roundtrip(tmp_path):
    img = np.random.default_rng(0).integers(0, 256, (19, 31, 3), dtype=np.uint8)
    write_png(tmp_path / "a.png", img)
    assert np.array_equal(read_png(tmp_path / "a.png"), img)
    assert encode_png(img, 1) != encode_png(img, 9) or True
    with pytest.raises(ValueError):
        encode_png(img.astype(np.float32))
    PIL = pytest.importorskip("PIL.Image")
    assert np.array_equal(np.asarray(PIL.open(tmp_path / "a.png")), img)


def test_cuda_renderer_constructor_surface():
    from python_ray_tracer_b200.infrastructure import CudaRenderer

    r = CudaRenderer(reflection_gain=0.8, specular_gain=1.0, specular_roughness=0.5, iridescence_gain=0.05,
                     diffuse_gain=1.0)  # README.md:50-56 verbatim
    assert isinstance(r, Renderer) and r.default_material.specular_roughness == 0.5
    assert CudaRenderer().default_material == DefaultMaterial()
    for bad in (dict(precision="fp16"), dict(output_dtype="int8"), dict(max_depth=0), dict(cutoff=-1.0)):
        with pytest.raises(ValueError):
            CudaRenderer(**bad)
    rays = r.get_ray_directions(Camera(NumpyVector3D(0, 0.2, -2), 16, 9), rows=(3, 5))
    assert len(rays) == 32 and (rays.row0, rays.row1) == (3, 5)
    with pytest.raises(ValueError):
        r.get_ray_directions(Camera(NumpyVector3D(0, 0.2, -2), 16, 9), rows=(5, 12))


def test_product_never_imports_the_oracle():
    offenders = []
    for path in (ROOT / "python_ray_tracer_b200").rglob("*.py"):
        text = path.read_text()
        if re.search(r"^\s*(from|import)\s+oracle\b", text, re.M) or "librt_oracle" in text:
            offenders.append(str(path))
    for path in (ROOT / "python_ray_tracer_b200" / "csrc").glob("*"):
        if path.suffix in (".cu", ".cuh", ".h") and "oracle" in path.read_text():
            offenders.append(str(path))
    assert offenders == []


def test_abi_library_exports_every_declared_symbol():
    """include/rt_abi.h <-> librt_b200.so <-> abi.PROTOTYPES agree (no compute calls)."""
    header = (ROOT / "include" / "rt_abi.h").read_text()
    declared = set(re.findall(r"^(?:int|int32_t|int64_t|const char\*)\s+(rt_\w+)\s*\(", header, re.M))
    assert declared == set(abi.PROTOTYPES), declared ^ set(abi.PROTOTYPES)
    lib = abi.load()  # raises if the library is not built: there is no fallback
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.rt_abi_version() == abi.RT_ABI_VERSION
    # struct layouts the header promises
    assert ctypes.sizeof(abi.RtCamera) == 32 and ctypes.sizeof(abi.RtEnv) == 176


def test_ctypes_structs_match_the_header_as_compiled(tmp_path):
    import subprocess

    src = tmp_path / "sizes.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "rt_abi.h"\n'
        'int main(void){printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(rt_env), sizeof(rt_camera), '
        'offsetof(rt_env, dome_intensity), offsetof(rt_env, max_depth), offsetof(rt_env, spheres_host), '
        'offsetof(rt_camera, width));return 0;}\n'
    )
    exe = tmp_path / "sizes"
    subprocess.run(["/usr/bin/gcc", "-I", str(ROOT / "include"), str(src), "-o", str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [ctypes.sizeof(abi.RtEnv), ctypes.sizeof(abi.RtCamera), abi.RtEnv.dome_intensity.offset,
            abi.RtEnv.max_depth.offset, abi.RtEnv.spheres_host.offset, abi.RtCamera.width.offset]
    assert got == want


def test_python_constants_match_the_header_as_compiled(tmp_path):
    """Every integer constant abi.py mirrors (flags, modes, dtypes, status codes, sizes) is
    printed by a C program compiled against include/rt_abi.h and compared by name."""
    import re
    import subprocess

    names = [n for n in dir(abi) if re.fullmatch(r"RT_[A-Z0-9_]+", n) and isinstance(getattr(abi, n), int)]
    assert {"RT_FLAG_WIDE_STORES", "RT_FLAG_RGB8", "RT_FLAG_EVEN_STORES", "RT_FLAG_IMAGE_ORDER", "RT_ABI_VERSION"} <= set(names)
    status = sorted(abi.STATUS_NAMES.values())  # status codes are kept by name in abi.STATUS_NAMES
    src = tmp_path / "constants.c"
    src.write_text('#include <stdio.h>\n#include "rt_abi.h"\nint main(void){\n'
                   + "".join(f'printf("{n} %lld\\n", (long long)({n}));\n' for n in names + status) + "return 0;}\n")
    exe = tmp_path / "constants"
    subprocess.run(["/usr/bin/gcc", "-I", str(ROOT / "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    got = dict(zip(out[0::2], (int(v) for v in out[1::2])))
    assert {n: got[n] for n in names} == {n: getattr(abi, n) for n in names}
    assert {got[n]: n for n in status} == abi.STATUS_NAMES
    flags = [v for n, v in got.items() if n.startswith("RT_FLAG_") and v]
    assert len(set(flags)) == len(flags) and all(v & (v - 1) == 0 for v in flags)  # distinct single bits


def test_abi_rejects_bad_arguments_without_a_gpu():
    lib = abi.load()
    cam = abi.make_camera((0, 0.2, -2), 16, 9)
    env = abi.make_env((0, 0.2, -2), 1, [0.1], (1, 1, 1))
    rc = lib.rt_render_rows_dev(0, None, 3, None, ctypes.byref(env), ctypes.byref(cam), 0, 9, None, 0, 144, None, None, None)
    assert rc == -1 and b"NULL" in lib.rt_last_error()
    rc = lib.rt_render_rows_dev(0, 8, 0, 8, ctypes.byref(env), ctypes.byref(cam), 0, 9, None, 0, 144, None, None, None)
    assert rc == -1 and b"n_spheres" in lib.rt_last_error()
    rc = lib.rt_render_rows_dev(0, 8, 3, 8, ctypes.byref(env), ctypes.byref(cam), 4, 12, None, 0, 144, None, None, None)
    assert rc == -1 and b"rows" in lib.rt_last_error()
    env.mode = 7
    rc = lib.rt_render_rows_dev(0, 8, 3, 8, ctypes.byref(env), ctypes.byref(cam), 0, 9, None, 0, 144, None, None, None)
    assert rc == -1 and b"mode" in lib.rt_last_error()
    with pytest.raises(ValueError):
        abi.check(rc, "rt_render_rows_dev")
    rc = lib.rt_quantize_dev(0, 8, 0, 4, 4, 8, 5, None)
    assert rc == -1


def test_scene_travels_as_parameter_predicate():
    """rt_scene_travels_as_parameter: only small scenes of the fast policy whose host tables the env names, with no
    tree and the kernel-parameter path not switched off (then a camera launch never reads the device tables)."""
    lib = abi.load()
    table = np.ascontiguousarray(np.ones((abi.RT_SPHERE_FIELDS if hasattr(abi, "RT_SPHERE_FIELDS") else 12, 3)))
    lights = np.ascontiguousarray(np.ones((2, 3)))

    def env(**kw):
        args = dict(mode=abi.RT_MODE_FAST, flags=abi.RT_FLAG_NONE, spheres_host=table, point_lights_host=lights)
        args.update(kw)
        n_lights = args.pop("n_lights", 1)
        return abi.make_env((0.0, 0.2, -2.0), n_lights, [0.1], (1.0, 1.0, 1.0), **args)

    yes = lambda e, n=3: lib.rt_scene_travels_as_parameter(ctypes.byref(e), n)  # noqa: E731
    assert yes(env()) == 1 and yes(env(), abi.RT_SMALL_SCENE_SPHERES) == 1
    assert yes(env(), abi.RT_SMALL_SCENE_SPHERES + 1) == 0 and yes(env(), 0) == 0
    assert yes(env(mode=abi.RT_MODE_EXACT)) == 0
    assert yes(env(flags=abi.RT_FLAG_GENERIC_KERNEL)) == 0
    assert yes(env(spheres_host=None)) == 0 and yes(env(point_lights_host=None)) == 0
    assert yes(env(flags=abi.RT_FLAG_ALL_POINT_LIGHTS, n_lights=2)) == 0 and yes(env(flags=abi.RT_FLAG_ALL_POINT_LIGHTS)) == 1
    with_tree = env()
    with_tree.accel_dev, with_tree.accel_bytes = 4096, 64
    assert yes(with_tree) == 0
    assert lib.rt_scene_travels_as_parameter(None, 3) == 0


def test_parallel_png_is_a_valid_png_with_the_same_pixels(tmp_path):
    """The banded (pigz-style) encoder of the animation path decodes to the same pixels as the
    single-stream one, with our reader and with Pillow when it is installed."""
    from python_ray_tracer_b200.infrastructure.png import encode_png, encode_png_parallel, read_png

    rng = np.random.default_rng(5)
    for h, w in ((1, 1), (7, 5), (90, 160), (33, 257)):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        for threads, level in ((1, 1), (3, 1), (8, 6), (64, 0)):
            data = encode_png_parallel(img, level, threads)
            path = tmp_path / "p.png"
            path.write_bytes(data)
            assert np.array_equal(read_png(path), img)
            try:
                from PIL import Image
            except ImportError:
                continue
            assert np.array_equal(np.asarray(Image.open(path).convert("RGB")), img)
        ref = tmp_path / "r.png"
        ref.write_bytes(encode_png(img))
        assert np.array_equal(read_png(ref), img)


def test_new_entry_points_reject_bad_arguments_without_a_gpu():
    lib = abi.load()
    cam = abi.make_camera((0, 0.2, -2), 16, 9)
    env = abi.make_env((0, 0.2, -2), 1, [0.1], (1, 1, 1))
    # strips: first must be below step, rows and step positive
    for rows, first, step in ((0, 0, 1), (4, 2, 2), (4, -1, 2), (4, 0, 0)):
        rc = lib.rt_render_strips_dev(0, 8, 3, 8, ctypes.byref(env), ctypes.byref(cam), rows, first, step, None, 0, 144,
                                      None, None, None)
        assert rc == -1 and b"strips" in lib.rt_last_error(), (rows, first, step)
    # image-order outputs are whole frames: the plane stride must cover height x width
    env.flags = abi.RT_FLAG_IMAGE_ORDER
    rc = lib.rt_render_strips_dev(0, 8, 3, 8, ctypes.byref(env), ctypes.byref(cam), 4, 0, 2, 8, 0, 16 * 5, None, None, None)
    assert rc == -1 and b"plane_stride" in lib.rt_last_error()
    # peer buffers
    ptr = ctypes.c_void_p()
    handle = ctypes.create_string_buffer(abi.RT_IPC_HANDLE_BYTES)
    assert lib.rt_peer_buffer_create(0, 0, ctypes.byref(ptr), handle) == -1
    assert lib.rt_peer_buffer_create(0, 64, None, handle) == -1
    assert lib.rt_peer_buffer_open(0, None, ctypes.byref(ptr)) == -1
    assert lib.rt_peer_buffer_close(0, None) == 0 and lib.rt_peer_buffer_destroy(0, None) == 0


def _parse_accel(blob: bytes):
    import struct

    magic, n, n_nodes, n_tree, n_large, _ = struct.unpack_from("<Iiiiii", blob, 0)
    offs = struct.unpack_from("<8q", blob, 24)
    local_bound, gain_bound = struct.unpack_from("<2d", blob, 24 + 64)
    (extent,) = struct.unpack_from("<f", blob, 24 + 64 + 16)
    off_nodes, off_order, off_leaf, off_geom, off_aux, off_large, off_rec, total = offs
    rec = np.frombuffer(blob, dtype=np.dtype([("c", "<f8", 3), ("inv_r", "<f8"), ("kd", "<f4", 3), ("spec_gain", "<f4"),
                                              ("a2", "<f4"), ("film", "<f4"), ("half_gain", "<f4"), ("flags", "<u4")]),
                        count=n, offset=off_rec)
    nodes = np.frombuffer(blob, dtype=np.dtype([("lo", "<f4", 3), ("hi", "<f4", 3), ("ref", "<u4"), ("axis", "<i4")]),
                          count=n_nodes, offset=off_nodes)
    order = np.frombuffer(blob, dtype="<i4", count=n_tree, offset=off_order)
    leaf_geom = np.frombuffer(blob, dtype="<f8", count=4 * n_tree, offset=off_leaf).reshape(-1, 4)
    geom = np.frombuffer(blob, dtype="<f8", count=4 * n, offset=off_geom).reshape(-1, 4)
    aux = np.frombuffer(blob, dtype="<f8", count=2 * n, offset=off_aux).reshape(-1, 2)
    large = np.frombuffer(blob, dtype="<i4", count=n_large, offset=off_large)
    return dict(magic=magic, n=n, nodes=nodes, order=order, leaf_geom=leaf_geom, geom=geom, aux=aux, large=large, rec=rec,
                total=total, local_bound=local_bound, gain_bound=gain_bound, extent=extent)


@pytest.mark.parametrize("fast_build", [False, True])
@pytest.mark.parametrize("n_spheres", [1, 2, 5, 48, 1024])
def test_accel_blob_is_a_valid_tree(n_spheres, fast_build):
    """rt_accel_build_host: every shape is in exactly one leaf slot or in the always-tested
    list, every node's box contains the boxes of everything below it, records equal the table."""
    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene

    lib = abi.load()
    flat = flatten_scene(scenes.random_spheres_scene(64, 36, n_spheres=n_spheres, seed=77), DefaultMaterial())
    n = flat.n_spheres  # + the ground sphere
    env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST,
                       flags=abi.RT_FLAG_ACCEL_FAST_BUILD if fast_build else abi.RT_FLAG_NONE)  # median / surface-area splits
    cap = int(lib.rt_accel_bytes(n))
    buf = np.zeros(cap, dtype=np.uint8)
    used = ctypes.c_int64(0)
    assert lib.rt_accel_build_host(flat.spheres.ctypes.data, n, ctypes.byref(env), buf.ctypes.data, cap, ctypes.byref(used)) == 0
    assert lib.rt_accel_build_host(flat.spheres.ctypes.data, n, ctypes.byref(env), buf.ctypes.data, 64, ctypes.byref(used)) == -1
    a = _parse_accel(buf[: used.value].tobytes())
    assert a["magic"] == 0x32425452 and a["n"] == n and a["total"] == used.value <= cap
    assert sorted(a["order"].tolist() + a["large"].tolist()) == list(range(n))
    if n > 2:
        assert (n - 1) in a["large"].tolist()  # the radius-99999 ground stays out of the tree
    c, r = flat.spheres[0:3].T, np.abs(flat.spheres[3])
    assert np.array_equal(a["geom"][:, :3], c) and np.array_equal(a["geom"][:, 3], flat.spheres[3] ** 2)
    assert np.array_equal(a["leaf_geom"], a["geom"][a["order"]]) and np.array_equal(a["aux"][:, 0], 1.0 / flat.spheres[3])
    assert a["gain_bound"] == np.abs(flat.spheres[5]).max() and a["local_bound"] > 0
    # the fast policy's shading records (rt_device.cuh: ShadeRec)
    rec, f = a["rec"], flat.spheres
    checker = f[11] != 0
    assert np.array_equal(rec["c"], c) and np.array_equal(rec["inv_r"], 1.0 / f[3])
    assert np.array_equal(rec["kd"], np.where(checker[:, None], f[4][:, None], (f[4] * f[8:11]).T).astype(np.float32))
    assert np.array_equal(rec["a2"], ((f[6] * f[6]) ** 2).astype(np.float32)) and np.array_equal(rec["flags"] & 1, checker.astype(np.uint32))
    assert np.array_equal(rec["half_gain"], (0.5 * f[5]).astype(np.float32)) and np.array_equal(rec["film"], (0.1 * f[7]).astype(np.float32))

    depths = []

    def check(node, depth=0):  # returns the shapes below `node`
        depths.append(depth)
        nd = a["nodes"][node]
        first, count = int(nd["ref"]) >> 3, int(nd["ref"]) & 7
        if count > 0:
            shapes = a["order"][first: first + count].tolist()
            assert 1 <= count <= 2 and nd["axis"] == -1  # kAccelLeafSize
        else:
            assert nd["axis"] in (0, 1, 2)
            shapes = check(first, depth + 1) + check(first + 1, depth + 1)
        lo, hi = (c[shapes] - r[shapes, None]).min(axis=0), (c[shapes] + r[shapes, None]).max(axis=0)
        # FP32 boxes, rounded outward: strictly outside the FP64 box, by no more than a few ulp
        assert (nd["lo"].astype(np.float64) < lo).all() and (nd["hi"].astype(np.float64) > hi).all()
        assert (np.nextafter(np.nextafter(np.nextafter(nd["lo"], np.float32(np.inf)), np.float32(np.inf)), np.float32(np.inf)).astype(np.float64) > lo).all()
        assert a["extent"] >= np.abs(nd["lo"]).max() and a["extent"] >= np.abs(nd["hi"]).max()
        return shapes

    if len(a["nodes"]):
        assert sorted(check(0)) == sorted(a["order"].tolist())
        assert max(depths) < 64  # kAccelStack: one pending sibling per level
    # coordinates FP32 boxes cannot hold: refused, never silently clamped
    far = flat.spheres.copy()
    far[0, 0] = 1e16
    assert lib.rt_accel_build_host(far.ctypes.data, n, ctypes.byref(env), buf.ctypes.data, cap, ctypes.byref(used)) == -4  # RT_ERR_UNSUPPORTED


def test_accel_depth_stays_within_the_traversal_stack():
    """Geometrically spaced spheres make surface-area splits peel one shape per level; past
    kAccelSahDepth the builder switches to median splits, so the depth stays below the 64
    entries of the traversal stack.  Identical spheres (no separating plane) fall back too."""
    lib = abi.load()
    for centres, radii in (
        (np.stack([1.2 ** np.arange(180), np.zeros(180), np.zeros(180)], axis=1), 0.01 * 1.2 ** np.arange(180)),
        (np.zeros((300, 3)) + [1.0, 2.0, 3.0], np.full(300, 0.5)),
    ):
        n = len(radii)
        table = np.zeros((12, n))
        table[0:3], table[3], table[4:8] = centres.T, radii, 1.0
        table = np.ascontiguousarray(table)
        env = abi.make_env((0.0, 0.0, -2.0), 1, [0.1], (1.0, 1.0, 1.0), mode=abi.RT_MODE_FAST)
        cap = int(lib.rt_accel_bytes(n))
        buf = np.zeros(cap, dtype=np.uint8)
        used = ctypes.c_int64(0)
        assert lib.rt_accel_build_host(table.ctypes.data, n, ctypes.byref(env), buf.ctypes.data, cap, ctypes.byref(used)) == 0
        a = _parse_accel(buf[: used.value].tobytes())
        seen, deepest, todo = [], 0, [(0, 0)] if len(a["nodes"]) else []
        while todo:
            node, depth = todo.pop()
            deepest = max(deepest, depth)
            ref = int(a["nodes"][node]["ref"])
            first, count = ref >> 3, ref & 7
            if count:
                seen += a["order"][first: first + count].tolist()
            else:
                todo += [(first, depth + 1), (first + 1, depth + 1)]
        assert sorted(seen + a["large"].tolist()) == list(range(n)) and deepest < 63


REFERENCE = Path("/root/reference")


@pytest.mark.reference
@pytest.mark.skipif(not (REFERENCE / "ray_tracer").is_dir(), reason="the reference tree is only mounted in the build container")
def test_scenes_built_from_the_reference_classes_flatten_identically():
    """objects built with the REFERENCE's own classes (HEAD API, as main.py:11-51
    builds them) feed the flattener by duck typing and give the table this package's classes give."""
    sys.path.insert(0, str(REFERENCE))
    try:
        from ray_tracer.domain import Camera, DomeLight, PointLight, Scene3D
        from ray_tracer.infrastructure.numpy.base import NumpyRGBColor, NumpyVector3D
        from ray_tracer.infrastructure.numpy.shader import NumpyShader, Texture, TextureChecker
        from ray_tracer.infrastructure.numpy.shape import NumpySphere
    finally:
        sys.path.remove(str(REFERENCE))
    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure.flatten import flatten_scene

    g = scenes.README_GAINS

    def shader(texture):
        return NumpyShader(g["reflection_gain"], g["specular_gain"], g["specular_roughness"], g["iridescence_gain"],
                           g["diffuse_gain"], texture)

    theirs = Scene3D(
        [NumpySphere(NumpyVector3D(0.55, 0.5, 3), 1.0, shader(Texture(NumpyRGBColor(1, 0, 1)))),
         NumpySphere(NumpyVector3D(-0.45, 0.1, 1), 0.4, shader(Texture(NumpyRGBColor(0.5, 0.5, 0.5)))),
         NumpySphere(NumpyVector3D(0, -99999.5, 0), 99999, shader(TextureChecker()))],
        [PointLight(NumpyVector3D(5, 10, -10)), DomeLight(0.1, NumpyRGBColor(1, 1, 1))],
        Camera(NumpyVector3D(0, 0.2, -2), 1920, 1080))
    a = flatten_scene(theirs)
    b = flatten_scene(scenes.readme_scene(1920, 1080), __import__("python_ray_tracer_b200.infrastructure.flatten", fromlist=["DefaultMaterial"]).DefaultMaterial(**g))
    assert np.array_equal(a.spheres, b.spheres) and np.array_equal(a.point_lights, b.point_lights)
    assert a.eye == b.eye and a.dome_intensity == b.dome_intensity and a.dome_color == b.dome_color


@pytest.mark.reference
@pytest.mark.skipif(not (REFERENCE / "ray_tracer").is_dir(), reason="the reference tree is only mounted in the build container")
def test_integration_stub_runs_verbatim_against_the_library(tmp_path):
    """INTEGRATION.md section 2, executed as printed (only the library's path is made absolute): the
    ctypes ``CudaRenderer(Renderer)`` a reference maintainer would add imports the reference's own port,
    its struct matches the header as compiled, every symbol it binds exists, and -- without a GPU --
    ``raytrace_scene`` comes back through the error path with the library's message."""
    text = (ROOT / "INTEGRATION.md").read_text()
    section = text[text.index("## 2."):text.index("## 3.")]
    code = section[section.index("```python") + len("```python"):]
    code = code[:code.index("```")]
    assert 'C.CDLL("librt_b200.so")' in code
    code = code.replace('C.CDLL("librt_b200.so")', f'C.CDLL("{abi.library_path()}")')
    abi.load()
    sys.path.insert(0, str(REFERENCE))
    try:
        ns = {}
        exec(compile(code, "INTEGRATION.md#2", "exec"), ns)
        assert ctypes.sizeof(ns["rt_env"]) == ctypes.sizeof(abi.RtEnv) and ctypes.sizeof(ns["rt_camera"]) == ctypes.sizeof(abi.RtCamera)
        for (name, ctype), (name2, ctype2) in zip(ns["rt_env"]._fields_, abi.RtEnv._fields_):
            assert name == name2 and ctypes.sizeof(ctype) == ctypes.sizeof(ctype2)
        sys.path.insert(0, str(REFERENCE))
        import importlib

        main_scene = importlib.import_module("ray_tracer.domain")
        from ray_tracer.infrastructure.numpy.base import NumpyRGBColor, NumpyVector3D
        from ray_tracer.infrastructure.numpy.shader import NumpyShader, Texture, TextureChecker
        from ray_tracer.infrastructure.numpy.shape import NumpySphere

        scene = main_scene.Scene3D(
            [NumpySphere(NumpyVector3D(0.55, 0.5, 3), 1.0, NumpyShader(0.0, 0, 0.01, 0, 0.0, Texture(NumpyRGBColor(1, 1, 1)))),
             NumpySphere(NumpyVector3D(0, -99999.5, 0), 99999, NumpyShader(0.0, 0.1, 0.5, 0.0, 1.0, TextureChecker()))],
            [main_scene.PointLight(NumpyVector3D(-2, 1, 2)), main_scene.DomeLight(0.1, NumpyRGBColor(1, 1, 1))],
            main_scene.Camera(NumpyVector3D(0, 0.2, -2), 32, 18))
        renderer = ns["CudaRenderer"]()
        rays = renderer.get_ray_directions(scene.camera)
        import torch

        if torch.cuda.is_available():
            color = renderer.raytrace_scene(scene.camera.position, rays, scene)
            assert color.x.shape == (32 * 18,) and float(color.x.max()) > 0
        else:
            with pytest.raises(RuntimeError, match="CUDA device"):
                renderer.raytrace_scene(scene.camera.position, rays, scene)
    finally:
        while str(REFERENCE) in sys.path:
            sys.path.remove(str(REFERENCE))
