"""The device PNG encoder (csrc/rt_png_dev.cu: rt_png_encode_file_dev, rt_png_encode_bands_dev + rt_png_pack_dev, rt_render_png_host):
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


def test_device_png_of_random_sizes_and_contents(tmp_path):
    """Sizes off every alignment the kernel cares about (4 bytes a compare, 32 bytes a run-bit word, 16 bytes a packed
    store, bands cut short by the image's end), contents from one run to all literals, and the widest row a band can
    hold: the file always decodes to its input."""
    import torch

    from python_ray_tracer_b200.infrastructure.png import DevicePngEncoder

    rng = np.random.default_rng(20260611)

    def content(kind, h, w):
        if kind == 0:
            return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)                              # noise: stored bands
        if kind == 1:
            return np.broadcast_to(rng.integers(0, 256, 3, dtype=np.uint8), (h, w, 3)).copy()   # one colour: one run per row
        if kind == 2:
            img = np.zeros((h, w, 3), dtype=np.uint8)                                           # flat with sparse specks: short and long runs
            k = max(1, h * w // 17)
            img[rng.integers(0, h, k), rng.integers(0, w, k)] = rng.integers(0, 256, (k, 3))
            return img
        if kind == 3:
            g = np.add.outer(np.arange(h) * 3, np.arange(w))[:, :, None] // np.array([1, 2, 5])  # gradients: constant residuals
            return (g % 256).astype(np.uint8)
        img = np.repeat(rng.integers(0, 256, (h, (w + 6) // 7, 3), dtype=np.uint8), 7, axis=1)[:, :w]  # steps of 7 pixels
        return np.ascontiguousarray(img)

    sizes = [(1, 1), (1, 7), (2, 1), (3, 3), (5, 1081), (86, 1), (87, 2), (11, 10), (12, 21), (13, 43), (64, 85), (65, 86), (31, 341),
             (33, 342), (80, 599), (7, 2731), (3, 21666)]
    sizes += [(int(rng.integers(1, 90)), int(rng.integers(1, 700))) for _ in range(14)]
    for n, (h, w) in enumerate(sizes):
        img = content(n % 5, h, w)
        data = DevicePngEncoder(w, h).encode(torch.from_numpy(img).cuda())
        assert np.array_equal(_decode(data, tmp_path), img), (h, w, n % 5)
    with pytest.raises(ValueError):
        DevicePngEncoder(21667, 2)  # a row that no stored block can hold


def test_single_launch_file_equals_bands_plus_pack(tmp_path):
    """rt_png_encode_file_dev (one launch: chunk positions by a single-pass prefix sum, the last block writes the tail)
    makes byte for byte the file of rt_png_encode_bands_dev + rt_png_pack_dev, for every content and size, again and
    again on the same scratch area; a buffer too small for the file is reported as 0 bytes, never overrun."""
    import torch

    from python_ray_tracer_b200.infrastructure.cuda import abi

    lib = abi.load()
    rng = np.random.default_rng(7)
    stream = int(torch.cuda.current_stream(0).cuda_stream)
    big = np.zeros((2200, 7680, 3), dtype=np.uint8)  # 2 200 bands of one row: more bands than blocks fit the GPU at once,
    big[:, :, 0] = (np.arange(7680) // 11 % 256).astype(np.uint8)  # so late blocks wait on the sizes of blocks long gone
    big[500:900, 1000:5000] = rng.integers(0, 256, (400, 4000, 3), dtype=np.uint8)
    big[1500:, :, 2] = (np.arange(700)[:, None] % 256).astype(np.uint8)
    for name, img in list(_images(rng)) + [("noise 1920x1080 (every band stored)", rng.integers(0, 256, (1080, 1920, 3), dtype=np.uint8)),
                                            ("mixed 7680x2200 (2 200 bands)", big)]:
        h, w, _ = img.shape
        pixels = torch.from_numpy(np.ascontiguousarray(img)).cuda()
        scratch_bytes, capacity = int(lib.rt_png_scratch_bytes(w, h)), int(lib.rt_png_file_bound(w, h))
        scratch = torch.full((scratch_bytes,), 0xA5, dtype=torch.uint8, device="cuda")  # (nothing relies on what the scratch held)
        want_file = torch.zeros(capacity, dtype=torch.uint8, device="cuda")
        want_size = torch.zeros(1, dtype=torch.int64, device="cuda")
        abi.check(lib.rt_png_encode_bands_dev(0, pixels.data_ptr(), 3, 3 * w, h, w, h, 0, h, 1, scratch.data_ptr(), scratch_bytes, stream), "bands")
        abi.check(lib.rt_png_pack_dev(0, w, h, scratch.data_ptr(), want_file.data_ptr(), capacity, want_size.data_ptr(), stream), "pack")
        torch.cuda.synchronize()
        n = int(want_size.item())
        want = want_file[:n].cpu().numpy()
        assert n > 0 and np.array_equal(_decode(want.tobytes(), tmp_path), img), name
        guard = 64
        for round_ in range(3):
            got_file = torch.full((capacity + guard,), 0xEE, dtype=torch.uint8, device="cuda")
            got_size = torch.full((1,), -1, dtype=torch.int64, device="cuda")
            abi.check(lib.rt_png_encode_file_dev(0, pixels.data_ptr(), 3, 3 * w, w, h, scratch.data_ptr(), scratch_bytes,
                                                 got_file.data_ptr(), capacity, got_size.data_ptr(), stream), "file")
            torch.cuda.synchronize()
            assert int(got_size.item()) == n, (name, round_)
            assert np.array_equal(got_file[:n].cpu().numpy(), want), (name, round_)
            assert bool((got_file[capacity:] == 0xEE).all().item()), name
        # a buffer that cannot hold the file: 0 bytes reported, nothing written past its end
        if n < 140:
            continue
        small = n - 5
        got_file = torch.full((small + guard,), 0xEE, dtype=torch.uint8, device="cuda")
        got_size = torch.full((1,), -1, dtype=torch.int64, device="cuda")
        abi.check(lib.rt_png_encode_file_dev(0, pixels.data_ptr(), 3, 3 * w, w, h, scratch.data_ptr(), scratch_bytes,
                                             got_file.data_ptr(), small, got_size.data_ptr(), stream), "file")
        torch.cuda.synchronize()
        assert int(got_size.item()) == 0 and bool((got_file[small:] == 0xEE).all().item()), name
    # RGBA input and a pinned host destination
    img = rng.integers(0, 256, (97, 333, 4), dtype=np.uint8)
    img[:, 100:250] = 9
    h, w, _ = img.shape
    pixels = torch.from_numpy(img).cuda()
    scratch_bytes, capacity = int(lib.rt_png_scratch_bytes(w, h)), int(lib.rt_png_file_bound(w, h))
    scratch = torch.empty(scratch_bytes, dtype=torch.uint8, device="cuda")
    host_file = torch.zeros(capacity, dtype=torch.uint8).pin_memory()
    host_size = torch.zeros(1, dtype=torch.int64).pin_memory()
    abi.check(lib.rt_png_encode_file_dev(0, pixels.data_ptr(), 4, 4 * w, w, h, scratch.data_ptr(), scratch_bytes, host_file.data_ptr(),
                                         capacity, host_size.data_ptr(), stream), "file")
    torch.cuda.synchronize()
    n = int(host_size[0])
    assert n > 0 and np.array_equal(_decode(host_file.numpy()[:n].tobytes(), tmp_path), img[:, :, :3])


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
