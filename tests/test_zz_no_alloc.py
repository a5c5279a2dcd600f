"""rt_png_encode_static_dev: the device PNG encoder with its band scratch in a file-scope device array of the
library (RT_PNG_STATIC_SCRATCH_BYTES) instead of memory the caller allocated -- the entry point for callers that
cannot call cudaMalloc.  It has no counterpart in the reference.

Collected last on purpose.  The GPU test calls the new entry point through
the C ABI twice in a row, on different images, and wants byte for byte the file of the existing entry points
(rt_png_encode_bands_dev + rt_png_pack_dev with a scratch buffer of the caller's); the encoder is integer work,
so the tolerance is the one tests/test_png_device.py uses: none."""

import ctypes as C

import numpy as np
import pytest

from python_ray_tracer_b200.infrastructure.cuda import abi


def test_static_encoder_rejects_what_does_not_fit_without_a_gpu():
    lib = abi.load()
    limit = abi.RT_PNG_STATIC_SCRATCH_BYTES
    assert lib.rt_png_scratch_bytes(1920, 1080) <= limit < lib.rt_png_scratch_bytes(2560, 1440)
    assert lib.rt_png_scratch_bytes(2048, 1024) <= limit
    # (sizes are checked before any CUDA call: pointers only have to be non-NULL here)
    rc = lib.rt_png_encode_static_dev(0, 8, 3, 3 * 2560, 2560, 1440, 8, 1 << 30, 8, None)
    assert rc == -4
    message = lib.rt_last_error().decode()
    assert "2560 x 1440" in message and str(limit) in message and "rt_png_encode_bands_dev" in message
    rc = lib.rt_png_encode_static_dev(0, 8, 3, 3 * 30000, 30000, 10, 8, 1 << 30, 8, None)  # a row no band can hold
    assert rc == -4
    assert lib.rt_png_encode_static_dev(0, None, 3, 48, 16, 9, 8, 1 << 20, 8, None) == -1
    assert lib.rt_png_encode_static_dev(0, 8, 3, 48, 16, 9, None, 1 << 20, 8, None) == -1
    assert lib.rt_png_encode_static_dev(0, 8, 3, 48, 16, 9, 8, 1 << 20, None, None) == -1
    assert lib.rt_png_encode_static_dev(0, 8, 2, 48, 16, 9, 8, 1 << 20, 8, None) == -1   # pixel_bytes
    assert lib.rt_png_encode_static_dev(0, 8, 3, 47, 16, 9, 8, 1 << 20, 8, None) == -1   # stride shorter than a row
    assert lib.rt_png_encode_static_dev(-1, 8, 3, 48, 16, 9, 8, 1 << 20, 8, None) == -3


def _frames():
    """Two different images of different sizes: the second is smaller, so the shared array still holds bands of
    the first while it is encoded."""
    from python_ray_tracer_b200 import scenes
    from python_ray_tracer_b200.infrastructure import CudaRenderer

    rng = np.random.default_rng(2024)
    renderer = CudaRenderer(**scenes.README_GAINS)
    scene = scenes.readme_scene(1920, 1080)
    color = renderer.raytrace_scene(scene.camera.position, renderer.get_ray_directions(scene.camera), scene)
    first = color.rgba8.cpu().numpy().reshape(1080, 1920, -1)[:, :, :3].copy()  # a rendered frame of the largest size the header names
    mix = np.zeros((333, 701, 3), dtype=np.uint8)
    mix[50:200, 100:600] = rng.integers(0, 256, (150, 500, 3))   # noise: stored bands
    mix[200:, :, 1] = (np.arange(701) // 5 % 256).astype(np.uint8)  # runs and gradients; the rest one long run
    return [("README 1920x1080", first), ("mixed 701x333", mix), ("README 1920x1080 again", first)]


@pytest.mark.gpu
def test_static_encoder_equals_the_callers_scratch_encoder_two_calls_in_a_row(tmp_path):
    import torch

    from python_ray_tracer_b200.infrastructure.png import read_png

    lib = abi.load()
    device = 0
    stream = torch.cuda.current_stream(device)
    handle = int(stream.cuda_stream)
    for name, img in _frames():
        img = np.ascontiguousarray(img)
        h, w, _ = img.shape
        pixels = torch.from_numpy(img).cuda(device)
        capacity = int(lib.rt_png_file_bound(w, h))
        # ---- the existing entry points, scratch allocated by the caller
        scratch_bytes = int(lib.rt_png_scratch_bytes(w, h))
        scratch = torch.empty(scratch_bytes, dtype=torch.uint8, device=pixels.device)
        want_file = torch.zeros(capacity, dtype=torch.uint8, device=pixels.device)
        want_size = torch.zeros(1, dtype=torch.int64, device=pixels.device)
        abi.check(lib.rt_png_encode_bands_dev(device, pixels.data_ptr(), 3, 3 * w, h, w, h, 0, h, 1, scratch.data_ptr(),
                                              scratch_bytes, handle), "rt_png_encode_bands_dev")
        abi.check(lib.rt_png_pack_dev(device, w, h, scratch.data_ptr(), want_file.data_ptr(), capacity, want_size.data_ptr(),
                                      handle), "rt_png_pack_dev")
        # ---- the new one: inputs and outputs only (allocated before the call; the call itself allocates nothing)
        got_file = torch.zeros(capacity, dtype=torch.uint8, device=pixels.device)
        got_size = torch.zeros(1, dtype=torch.int64, device=pixels.device)
        torch.cuda.synchronize(device)
        allocated = torch.cuda.memory_allocated(device)
        free_before, _ = torch.cuda.mem_get_info(device)
        abi.check(lib.rt_png_encode_static_dev(device, pixels.data_ptr(), 3, 3 * w, w, h, got_file.data_ptr(), capacity,
                                               got_size.data_ptr(), handle), "rt_png_encode_static_dev")
        stream.synchronize()
        assert torch.cuda.memory_allocated(device) == allocated, name
        if name.endswith("again"):  # (the first call on a device may load the module; later ones must not move the free-memory figure)
            assert torch.cuda.mem_get_info(device)[0] >= free_before, name  # (nothing was taken from the device)
        n_want, n_got = int(want_size.item()), int(got_size.item())
        assert n_want > 0 and n_got == n_want, (name, n_want, n_got)
        want = want_file[:n_want].cpu().numpy()
        got = got_file[:n_got].cpu().numpy()
        assert np.array_equal(got, want), name
        path = tmp_path / "static.png"
        path.write_bytes(got.tobytes())
        assert np.array_equal(read_png(path), img), name
        abi.check(lib.rt_device_status(device), "rt_device_status")


@pytest.mark.gpu
def test_static_encoder_refuses_a_frame_beyond_its_array_on_a_gpu():
    import torch

    lib = abi.load()
    w, h = 2560, 1440
    pixels = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda:0")
    out = torch.zeros(int(lib.rt_png_file_bound(w, h)), dtype=torch.uint8, device="cuda:0")
    size = torch.full((1,), -7, dtype=torch.int64, device="cuda:0")
    rc = lib.rt_png_encode_static_dev(0, pixels.data_ptr(), 3, 3 * w, w, h, out.data_ptr(), out.numel(), size.data_ptr(), None)
    assert rc == -4 and b"RT_PNG_STATIC_SCRATCH_BYTES" in lib.rt_last_error()
    torch.cuda.synchronize()
    assert int(size.item()) == -7 and int(out.max().item()) == 0  # nothing was launched
