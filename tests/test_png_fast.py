"""rt_png_deflate_rows / encode_png_fast: the library's own PNG deflate (save_image's fast path,
base.py:143-151).  What matters is the decoded image: every PNG written here is decoded by zlib
(which also checks the stitched Adler-32), by this repo's reader and by Pillow."""
import ctypes as C
import io
import struct
import zlib

import numpy as np
import pytest

from python_ray_tracer_b200.infrastructure import png
from python_ray_tracer_b200.infrastructure.cuda import abi


def _decode_all(data: bytes, tmp_path, want: np.ndarray) -> None:
    path = tmp_path / "x.png"
    path.write_bytes(data)
    assert np.array_equal(png.read_png(path), want)
    try:
        from PIL import Image
    except ImportError:  # Pillow is optional here; zlib + our reader already decoded it
        return
    assert np.array_equal(np.array(Image.open(io.BytesIO(data)).convert("RGB")), want)


def _images():
    rng = np.random.default_rng(7)
    yield "noise (stored blocks win)", rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    yield "black", np.zeros((64, 200, 3), dtype=np.uint8)
    yield "white 1x1", np.full((1, 1, 3), 255, dtype=np.uint8)
    yield "column", rng.integers(0, 256, (300, 1, 3), dtype=np.uint8)
    yield "row", rng.integers(0, 256, (1, 700, 3), dtype=np.uint8)
    ramp = np.zeros((90, 640, 3), dtype=np.uint8)
    ramp[..., 0] = np.arange(640) // 3
    ramp[..., 1] = (np.arange(90) * 2)[:, None]
    ramp[..., 2] = 17
    yield "ramps (period-3 runs)", ramp
    checker = (((np.arange(480)[None, :] // 16) + (np.arange(270)[:, None] // 16)) % 2 * 255).astype(np.uint8)
    yield "checker", np.repeat(checker[:, :, None], 3, axis=2)
    long_runs = np.zeros((8, 5000, 3), dtype=np.uint8)
    long_runs[:, 1234:] = (9, 200, 31)  # runs far longer than 258, across row ends too
    yield "long runs", long_runs
    smooth = (np.add.outer(np.arange(200), np.arange(300)) % 251).astype(np.uint8)
    yield "smooth + speckle", np.stack([smooth, smooth[::-1], rng.integers(0, 4, smooth.shape, dtype=np.uint8)], axis=2)


@pytest.mark.parametrize("name, image", list(_images()))
def test_fast_png_decodes_to_the_input(name, image, tmp_path):
    for threads in (1, 2, 5):
        _decode_all(png.encode_png_fast(image, threads), tmp_path, image)


def test_fast_png_of_a_rendered_frame_is_smaller_than_zlib_level_1(tmp_path):
    from _cases import load_case

    case = load_case("readme_160x90")
    rgb8 = np.clip(case.rgb.T.reshape(case.height, case.width, 3), 0.0, 1.0)
    rgb8 = (255.0 * rgb8).astype(np.uint8)
    fast = png.encode_png_fast(rgb8, 2)
    _decode_all(fast, tmp_path, rgb8)
    assert len(fast) < len(png.encode_png_parallel(rgb8, 1, 2))


def test_fast_png_drops_alpha_and_takes_strided_rows(tmp_path):
    rng = np.random.default_rng(11)
    rgba = rng.integers(0, 256, (41, 67, 4), dtype=np.uint8)
    _decode_all(png.encode_png_fast(rgba, 3), tmp_path, rgba[:, :, :3])
    wide = rng.integers(0, 256, (41, 100, 3), dtype=np.uint8)
    view = wide[:, 10:77]  # rows 300 bytes apart, 201 used
    assert not view.flags["C_CONTIGUOUS"]
    _decode_all(png.encode_png_fast(view, 2), tmp_path, view)
    flipped = wide[:, ::-1]  # negative pixel stride: copied first
    _decode_all(png.encode_png_fast(flipped, 2), tmp_path, flipped)


def test_skewed_histogram_is_length_limited(tmp_path):
    """Residual frequencies growing like Fibonacci numbers make a Huffman tree deeper than the 15
    bits deflate allows; the encoder flattens the histogram until the tree fits."""
    fib = [1, 1]
    while len(fib) < 27:
        fib.append(fib[-1] + fib[-2])
    values = np.concatenate([np.full(f, 10 + 7 * k, dtype=np.uint8) for k, f in enumerate(fib)])
    rng = np.random.default_rng(3)
    rng.shuffle(values)
    n = values.size // 3
    residual = values[: 3 * n].reshape(1, n, 3)  # what the Sub filter must produce
    image = np.cumsum(residual, axis=1, dtype=np.uint64).astype(np.uint8)
    data = png.encode_png_fast(image, 1)
    _decode_all(data, tmp_path, image)
    # it really was a Huffman block with long codes, not the stored fallback
    idat = data[data.index(b"IDAT") + 4:]
    assert (idat[2] >> 1) & 3 == 2 and len(data) < image.size


def test_deflate_rows_rejects_bad_arguments():
    lib = abi.load()
    img = np.zeros((4, 4, 3), dtype=np.uint8)
    cap = int(lib.rt_png_deflate_bound(4, 4))
    out = np.zeros(cap, dtype=np.uint8)
    used, adler, raw = C.c_int64(0), C.c_uint32(0), C.c_int64(0)
    args = lambda **kw: [kw.get("pixels", img.ctypes.data), kw.get("width", 4), kw.get("rows", 4), kw.get("pixel_bytes", 3),  # noqa: E731
                         kw.get("stride", 12), 1, out.ctypes.data, kw.get("capacity", cap), C.byref(used), C.byref(adler), C.byref(raw)]
    assert lib.rt_png_deflate_rows(*args()) == 0
    for bad in ({"pixels": None}, {"width": 0}, {"rows": -1}, {"pixel_bytes": 2}, {"stride": 11}, {"capacity": cap - 1}):
        assert lib.rt_png_deflate_rows(*args(**bad)) == -1, bad
    assert lib.rt_png_deflate_bound(0, 4) == -1
    with pytest.raises(ValueError):
        png.encode_png_fast(np.zeros((4, 4), dtype=np.uint8))


def test_adler32_combine_matches_zlib():
    rng = np.random.default_rng(5)
    for n1, n2 in ((0, 1), (1, 0), (5, 7), (70000, 131071), (65521, 65521)):
        a, b = rng.integers(0, 256, n1, dtype=np.uint8).tobytes(), rng.integers(0, 256, n2, dtype=np.uint8).tobytes()
        assert png.adler32_combine(zlib.adler32(a), zlib.adler32(b), n2) == zlib.adler32(a + b)


def test_bands_concatenate_into_one_raw_deflate_stream():
    """The C entry point alone: two bands, stitched by hand, inflate to the Sub-filtered rows."""
    lib = abi.load()
    rng = np.random.default_rng(9)
    img = (rng.integers(0, 6, (10, 33, 3)) * 40).astype(np.uint8)
    parts, adlers = [], []
    for r0, r1, last in ((0, 4, 0), (4, 10, 1)):
        band = np.ascontiguousarray(img[r0:r1])
        cap = int(lib.rt_png_deflate_bound(33, r1 - r0))
        out = np.zeros(cap, dtype=np.uint8)
        used, adler, raw = C.c_int64(0), C.c_uint32(0), C.c_int64(0)
        abi.check(lib.rt_png_deflate_rows(band.ctypes.data, 33, r1 - r0, 3, 99, last, out.ctypes.data, cap, C.byref(used),
                                          C.byref(adler), C.byref(raw)), "rt_png_deflate_rows")
        assert raw.value == (r1 - r0) * 100
        parts.append(out[: used.value].tobytes())
        adlers.append((adler.value, raw.value))
    rows = zlib.decompress(b"".join(parts), -15)
    flat = img.reshape(10, 99).astype(np.int16)
    want = np.concatenate([np.ones((10, 1), dtype=np.uint8), np.concatenate([flat[:, :3], flat[:, 3:] - flat[:, :-3]], axis=1).astype(np.uint8)], axis=1)
    assert rows == want.tobytes()
    assert png.adler32_combine(adlers[0][0], adlers[1][0], adlers[1][1]) == zlib.adler32(rows)
    assert struct.pack(">I", zlib.adler32(rows)) == struct.pack(">I", png.adler32_combine(adlers[0][0], adlers[1][0], adlers[1][1]))


def test_turntable_encode_stage_both_encoders(tmp_path):
    """The encode stage of the batch driver on a frame as the GPU delivers it ([n, 4] RGBA8):
    the library's deflate (default) and zlib decode to the same RGB."""
    from concurrent.futures import ThreadPoolExecutor

    from python_ray_tracer_b200 import turntable

    rng = np.random.default_rng(13)
    w, h = 96, 54
    rgba = np.zeros((h, w, 4), dtype=np.uint8)
    rgba[..., :3] = (np.add.outer(np.arange(h), np.arange(w))[:, :, None] // np.array([1, 2, 3])) % 256
    rgba[10:20, 30:60, :3] = rng.integers(0, 256, (10, 30, 3))
    rgba[..., 3] = 255
    flat = rgba.reshape(-1, 4)
    with ThreadPoolExecutor(3) as pool:
        fast = turntable.encode_frame(flat, w, h, None, 3, pool)
        slow = turntable.encode_frame(flat, w, h, 1, 3, pool)
    _decode_all(fast, tmp_path, rgba[:, :, :3])
    _decode_all(slow, tmp_path, rgba[:, :, :3])


def test_write_png_with_the_library_deflate(tmp_path):
    rng = np.random.default_rng(17)
    img = (rng.integers(0, 3, (33, 47, 3)) * 100).astype(np.uint8)
    png.write_png(tmp_path / "fast.png", img, None)
    png.write_png(tmp_path / "zlib.png", img, 6)
    assert np.array_equal(png.read_png(tmp_path / "fast.png"), img)
    assert np.array_equal(png.read_png(tmp_path / "zlib.png"), img)
    from python_ray_tracer_b200.infrastructure import CudaRenderer

    assert CudaRenderer().png_compress_level is None  # the library's deflate is save_image's default
    assert CudaRenderer(png_compress_level=6).png_compress_level == 6


def test_turntable_resume_ignores_partial_frames(tmp_path):
    """Resume counts a frame as done only under its final name: a leftover ``.png.tmp`` (a run killed
    mid-write) is deleted and its frame rendered again; finished frames are skipped."""
    from python_ray_tracer_b200 import turntable

    turntable.frame_path(tmp_path, 0).write_bytes(b"done")
    turntable.frame_path(tmp_path, 2).with_suffix(".png.tmp").write_bytes(b"\x89PNG truncated")
    turntable.frame_path(tmp_path, 3).write_bytes(b"done")
    turntable.frame_path(tmp_path, 3).with_suffix(".png.tmp").write_bytes(b"stale")
    assert turntable.frames_to_render(tmp_path, 5) == [1, 2, 4]
    assert not list(tmp_path.glob("*.tmp"))
    assert turntable.frames_to_render(tmp_path, 5, resume=False) == [0, 1, 2, 3, 4]
    assert turntable.frames_to_render(tmp_path, 5, rank=1, world=2) == [1]
