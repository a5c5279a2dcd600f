"""The device PNG encoder (csrc/rt_png_dev.cu: rt_png_encode_bands_dev + rt_png_pack_dev, rt_render_png_host):
its files decode -- with zlib through this repo's reader, and with Pillow where installed -- to exactly the
pixels that went in, whatever the content (rendered frames, flat colour whose runs span every thread of a band,
noise that has to be stored, gradients), size (1x1 up to 8K rows, widths off every alignment) and pixel layout
(RGB or RGBA, whole frames or the interleaved strips of the multi-GPU partition)."""

import ctypes as C
import io
import zlib

import numpy as np
import pytest

from _cases import load_case

pytestmark = pytest.mark.gpu


def _decode(data: bytes, tmp_path) -> np.ndarray:
    from python_ray_tracer_b200.infrastructure.png import read_png

    path = tmp_path / "device.png"
    path.write_bytes(bytes(data))
    ours = read_png(path)
    try:
        from PIL import Image

        theirs = np.asarray(Image.open(io.BytesIO(bytes(data))).convert("RGB"))
        assert np.array_equal(ours, theirs)
    except ImportError:
        pass
    return ours


def _chunks(data: bytes):
    pos, out = 8, []
    while pos < len(data):
        n = int.from_bytes(data[pos:pos + 4], "big")
        tag, body, crc = data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n], data[pos + 8 + n:pos + 12 + n]
        assert zlib.crc32(tag + body) == int.from_bytes(crc, "big"), tag
        out.append((tag, body))
        pos += 12 + n
    return out


def _images(rng):
    yield "1x1", np.array([[[7, 200, 33]]], dtype=np.uint8)
    yield "3x2", rng.integers(0, 256, (2, 3, 3), dtype=np.uint8)
    yield "flat 700x300 (one run through every band)", np.full((300, 700, 3), 0, dtype=np.uint8)
    yield "flat colour 257x65", np.broadcast_to(np.array([10, 20, 30], dtype=np.uint8), (65, 257, 3)).copy()
    yield "noise 333x97 (stored bands)", rng.integers(0, 256, (97, 333, 3), dtype=np.uint8)
    g = np.add.outer(np.arange(211), np.arange(1021))[:, :, None] // np.array([1, 2, 3])
    yield "gradient 1021x211", (g % 256).astype(np.uint8)
    mix = np.zeros((540, 960, 3), dtype=np.uint8)
    mix[100:300, 200:700] = rng.integers(0, 256, (200, 500, 3))
    mix[300:, :, 1] = (np.arange(960) // 7 % 256).astype(np.uint8)
    yield "mixed 960x540", mix
    runs = np.zeros((40, 3000, 3), dtype=np.uint8)  # runs whose lengths sit around the 258 cut and the minimum of 3
    x = 0
    for k, length in enumerate([1, 2, 3, 4, 85, 86, 87, 171, 172, 173, 257, 258, 259, 260, 516, 517]):
        runs[:, x:x + length] = (37 * k + 11) % 256
        x += length + 1
        runs[:, x - 1] = rng.integers(0, 256, (40, 3))
    yield "runs around 3 and 258", runs


def test_device_png_decodes_to_its_input(tmp_path):
    import torch

    from python_ray_tracer_b200.infrastructure.png import DevicePngEncoder

    rng = np.random.default_rng(99)
    for name, img in _images(rng):
        h, w, _ = img.shape
        enc = DevicePngEncoder(w, h)
        data = enc.encode(torch.from_numpy(img).cuda())
        assert np.array_equal(_decode(data, tmp_path), img), name
        tags = [t for t, _ in _chunks(bytes(data))]
        assert tags[0] == b"IHDR" and tags[-1] == b"IEND" and set(tags[1:-1]) == {b"IDAT"}, name
        # RGBA input: alpha is skipped
        rgba = np.concatenate([img, np.full((h, w, 1), 255, dtype=np.uint8)], axis=2)
        assert bytes(enc.encode(torch.from_numpy(rgba).cuda())) == bytes(data), name


def test_device_png_of_rendered_frames(rt_fast, tmp_path):
    """A 1080p README frame and the 960x540 main.py frame: the file decodes to the kernel's 8-bit image, is a
    tenth of it, and rt_render_png_host hands back the same file through host buffers."""
    import torch

    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure.cuda import abi
    from python_ray_tracer_b200.infrastructure.png import DevicePngEncoder

    lib = abi.load()
    for scene in (scenes.readme_scene(1920, 1080), scenes.main_scene(960, 540)):
        w, h = int(scene.camera.width), int(scene.camera.height)
        flat = rt_fast.flatten(scene)
        color = rt_fast.render_rows(flat, scene.camera, 0, h, want_rgba8=True, flags=abi.RT_FLAG_RGB8, no_planes=True)
        want = color.rgba8.cpu().numpy().reshape(h, w, 3)
        data = bytes(DevicePngEncoder(w, h).encode(color.rgba8))
        assert np.array_equal(_decode(data, tmp_path), want)
        assert len(data) < want.size // 6, (len(data), want.size)
        env = rt_fast._plain_env(flat)
        cam = abi.make_camera(flat.eye, w, h)
        cap = int(lib.rt_png_file_bound(w, h))
        for pinned in (True, False):  # the device writes pinned memory directly; anything else goes through a staging buffer
            out = torch.empty(cap, dtype=torch.uint8)
            out = out.pin_memory() if pinned else out
            used = C.c_int64(0)
            abi.check(lib.rt_render_png_host(0, flat.spheres.ctypes.data, flat.n_spheres, flat.point_lights.ctypes.data, C.byref(env),
                                             C.byref(cam), out.data_ptr(), cap, C.byref(used)), "rt_render_png_host")
            assert bytes(out.numpy()[: used.value]) == data
        small = torch.empty(4096, dtype=torch.uint8).pin_memory()
        assert lib.rt_render_png_host(0, flat.spheres.ctypes.data, flat.n_spheres, flat.point_lights.ctypes.data, C.byref(env),
                                      C.byref(cam), small.data_ptr(), 4096, C.byref(used)) == -1 and b"too small" in lib.rt_last_error()


def test_device_png_from_interleaved_strips(rt_fast, tmp_path):
    """The multi-GPU form on one device: three "ranks" each encode the bands of their own packed strips into
    the one scratch area, rank 0 packs: the file equals the single-call file."""
    import torch

    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure.cuda import abi, sharding
    from python_ray_tracer_b200.infrastructure.png import DevicePngEncoder

    lib = abi.load()
    w, h, world = 1000, 612, 3
    scene = scenes.readme_scene(w, h)
    flat = rt_fast.flatten(scene)
    whole = rt_fast.render_rows(flat, scene.camera, 0, h, want_rgba8=True, flags=abi.RT_FLAG_RGB8, no_planes=True).rgba8
    enc = DevicePngEncoder(w, h)
    want = bytes(enc.encode(whole))
    assert sharding.DEFAULT_STRIP_ROWS % enc.rows_per_band == 0
    stream = int(torch.cuda.current_stream(0).cuda_stream)
    enc.scratch.zero_()
    for rank in range(world):
        shard = sharding.StripShard(h, world, rank)
        packed = rt_fast.render_strips(flat, scene.camera, shard, want_rgba8=True, flags=abi.RT_FLAG_RGB8).rgba8
        abi.check(lib.rt_png_encode_bands_dev(0, packed.data_ptr(), 3, w * 3, shard.n_rows, w, h, rank * shard.strip_rows,
                                              shard.strip_rows, world, enc.scratch.data_ptr(), enc.scratch_bytes, stream), "bands")
    abi.check(lib.rt_png_pack_dev(0, w, h, enc.scratch.data_ptr(), enc.file.data_ptr(), enc.capacity, enc.size.data_ptr(), stream), "pack")
    torch.cuda.synchronize()
    assert bytes(enc.finish()) == want
    assert np.array_equal(_decode(want, tmp_path), whole.cpu().numpy().reshape(h, w, 3))


def test_pipeline_writes_the_device_png(rt_fast, tmp_path, c_oracle):
    """render_image_pipeline with the default (device) encoder and with the host encoder: same decoded pixels,
    the reference's within 1 LSB."""
    from _cases import assert_png_parity
    from python_ray_tracer_b200 import render_image_pipeline, scenes
    from python_ray_tracer_b200.infrastructure import CudaRenderer
    from python_ray_tracer_b200.infrastructure.png import read_png

    case = load_case("readme_160x90")
    scene = scenes.readme_scene(160, 90)
    render_image_pipeline(scene, tmp_path / "dev.png", CudaRenderer(**scenes.README_GAINS))
    render_image_pipeline(scene, tmp_path / "host.png", CudaRenderer(**scenes.README_GAINS, png_encoder="host"))
    a, b = read_png(tmp_path / "dev.png"), read_png(tmp_path / "host.png")
    assert np.array_equal(a, b)
    assert_png_parity(a.reshape(-1, 3), c_oracle.quantize(case.rgb), "pipeline PNG vs the reference")


@pytest.fixture(scope="module")
def rt_fast():
    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure import CudaRenderer

    return CudaRenderer(**scenes.README_GAINS, precision="fast")
