"""Minimal PNG writer/reader (8-bit RGB, zlib) for ``save_image``.

The reference writes its PNG through Pillow with default settings (``base.py:147-151``);
what a consumer compares is the decoded pixels, which do not depend on the encoder, so
the container is written here with the standard library only (no Pillow dependency on
the render path).  ``compress_level`` 6 is zlib's (and Pillow's) default.
"""

from __future__ import annotations

import struct
import zlib
from pathlib import Path
from typing import Optional

import numpy as np

_SIGNATURE = b"\x89PNG\r\n\x1a\n"


def _chunk(tag: bytes, payload: bytes) -> bytes:
    return struct.pack(">I", len(payload)) + tag + payload + struct.pack(">I", zlib.crc32(tag + payload) & 0xFFFFFFFF)


def encode_png(rgb8: np.ndarray, compress_level: int = 6) -> bytes:
    """``rgb8``: uint8 ``[H, W, 3]`` -> PNG bytes (colour type 2, filter 0 on every row)."""
    if rgb8.dtype != np.uint8 or rgb8.ndim != 3 or rgb8.shape[2] != 3:
        raise ValueError("expected uint8 [H, W, 3]")
    height, width, _ = rgb8.shape
    rows = np.zeros((height, 1 + 3 * width), dtype=np.uint8)
    rows[:, 1:] = rgb8.reshape(height, 3 * width)
    header = struct.pack(">IIBBBBB", width, height, 8, 2, 0, 0, 0)
    return b"".join(
        (_SIGNATURE, _chunk(b"IHDR", header), _chunk(b"IDAT", zlib.compress(rows.tobytes(), compress_level)),
         _chunk(b"IEND", b""))
    )


def _deflate_part(args) -> bytes:
    """Raw-deflate one band of filtered rows; every band but the last ends on a sync flush
    (byte-aligned, no BFINAL), so the parts concatenate into one valid deflate stream."""
    data, level, last = args
    comp = zlib.compressobj(level, zlib.DEFLATED, -15)
    return comp.compress(data) + comp.flush(zlib.Z_FINISH if last else zlib.Z_SYNC_FLUSH)


def encode_png_parallel(rgb8: np.ndarray, compress_level: int = 1, threads: int = 4, pool=None) -> bytes:
    """Same container as ``encode_png``, with the zlib stream produced in ``threads`` row
    bands compressed concurrently (zlib releases the GIL) and stitched together the way
    pigz does: independent raw-deflate parts, one zlib header, one Adler-32 of the whole.
    Decoded pixels are identical to ``encode_png``'s whatever the level or band count --
    this is the fast path of ``save_image`` for animation batches, where the encoder, not
    the GPU, bounds the frame rate."""
    if rgb8.dtype != np.uint8 or rgb8.ndim != 3 or rgb8.shape[2] != 3:
        raise ValueError("expected uint8 [H, W, 3]")
    height, width, _ = rgb8.shape
    rows = np.zeros((height, 1 + 3 * width), dtype=np.uint8)
    rows[:, 1:] = rgb8.reshape(height, 3 * width)
    parts = max(1, min(int(threads), height))
    bounds = [height * i // parts for i in range(parts + 1)]
    raw = rows.reshape(-1)
    jobs = [(raw[bounds[i] * rows.shape[1]:bounds[i + 1] * rows.shape[1]].tobytes(), compress_level, i + 1 == parts)
            for i in range(parts)]
    if parts == 1:
        bodies = [_deflate_part(jobs[0])]
    elif pool is not None:
        bodies = list(pool.map(_deflate_part, jobs))
    else:
        from concurrent.futures import ThreadPoolExecutor

        with ThreadPoolExecutor(parts) as own:
            bodies = list(own.map(_deflate_part, jobs))
    stream = b"\x78\x01" + b"".join(bodies) + struct.pack(">I", zlib.adler32(raw) & 0xFFFFFFFF)
    header = struct.pack(">IIBBBBB", width, height, 8, 2, 0, 0, 0)
    return b"".join((_SIGNATURE, _chunk(b"IHDR", header), _chunk(b"IDAT", stream), _chunk(b"IEND", b"")))


def adler32_combine(adler1: int, adler2: int, len2: int) -> int:
    """Adler-32 of two concatenated byte strings from their own checksums and the length of
    the second (zlib's ``adler32_combine``, which Python's ``zlib`` does not expose)."""
    base = 65521
    rem = len2 % base
    sum1 = adler1 & 0xFFFF
    sum2 = (rem * sum1) % base
    sum1 += (adler2 & 0xFFFF) + base - 1
    sum2 += ((adler1 >> 16) & 0xFFFF) + ((adler2 >> 16) & 0xFFFF) + base - rem
    if sum1 >= base:
        sum1 -= base
    if sum1 >= base:
        sum1 -= base
    if sum2 >= (base << 1):
        sum2 -= base << 1
    if sum2 >= base:
        sum2 -= base
    return (sum2 << 16) | sum1


def _fast_band(args):
    """One band through ``rt_png_deflate_rows`` (ctypes releases the GIL for the call)."""
    import ctypes as C

    from .cuda import abi

    lib, band, last = args
    rows, width, channels = band.shape
    capacity = int(lib.rt_png_deflate_bound(width, rows))
    out = np.empty(capacity, dtype=np.uint8)
    used, adler, raw = C.c_int64(0), C.c_uint32(0), C.c_int64(0)
    abi.check(lib.rt_png_deflate_rows(band.ctypes.data, width, rows, channels, band.strides[0], 1 if last else 0,
                                      out.ctypes.data, capacity, C.byref(used), C.byref(adler), C.byref(raw)),
              "rt_png_deflate_rows")
    return out[: used.value].tobytes(), int(adler.value), int(raw.value)


def encode_png_fast(pixels: np.ndarray, threads: int = 1, pool=None) -> bytes:
    """PNG of ``pixels`` (uint8 ``[H, W, 3]`` or ``[H, W, 4]``; alpha is dropped) through the
    library's own deflate (``rt_png_deflate_rows``: Sub filter, runs at distance 1 and 3,
    one dynamic-Huffman block per band): about 8x faster than zlib level 1 on rendered
    frames at the same size.  Bands are compressed concurrently and concatenated like
    ``encode_png_parallel``'s.  Decoded pixels are identical to every other encoder's."""
    from .cuda import abi

    if pixels.dtype != np.uint8 or pixels.ndim != 3 or pixels.shape[2] not in (3, 4):
        raise ValueError("expected uint8 [H, W, 3] or [H, W, 4]")
    if pixels.strides[2] != 1 or pixels.strides[1] != pixels.shape[2]:
        pixels = np.ascontiguousarray(pixels)
    height, width, _ = pixels.shape
    if height == 0 or width == 0:
        raise ValueError("empty image")
    lib = abi.load()
    parts = max(1, min(int(threads), height))
    bounds = [height * i // parts for i in range(parts + 1)]
    jobs = [(lib, pixels[bounds[i]:bounds[i + 1]], i + 1 == parts) for i in range(parts)]
    if parts == 1:
        done = [_fast_band(jobs[0])]
    elif pool is not None:
        done = list(pool.map(_fast_band, jobs))
    else:
        from concurrent.futures import ThreadPoolExecutor

        with ThreadPoolExecutor(parts) as own:
            done = list(own.map(_fast_band, jobs))
    adler = done[0][1]
    for _, a, n in done[1:]:
        adler = adler32_combine(adler, a, n)
    stream = b"\x78\x01" + b"".join(d[0] for d in done) + struct.pack(">I", adler & 0xFFFFFFFF)
    header = struct.pack(">IIBBBBB", width, height, 8, 2, 0, 0, 0)
    return b"".join((_SIGNATURE, _chunk(b"IHDR", header), _chunk(b"IDAT", stream), _chunk(b"IEND", b"")))


class DevicePngEncoder:
    """PNG files made on the GPU (``rt_png_encode_bands_dev`` + ``rt_png_pack_dev``, or ``rt_png_encode_file_dev`` in one launch for frames of many bands; csrc/rt_png_dev.cu): the
    8-bit image stays on the device, only the ~10x smaller file crosses PCIe -- written by the pack kernel
    straight into a pinned host buffer.  One encoder holds the scratch and landing buffers of one frame size
    on one device; ``encode`` returns the file's bytes.  Decoded pixels equal the input."""

    def __init__(self, width: int, height: int, device: int = 0) -> None:
        import torch

        from .cuda import abi

        self.lib = abi.load()
        self.width, self.height, self.device = int(width), int(height), int(device)
        self.rows_per_band = int(self.lib.rt_png_rows_per_band(self.width, self.height))
        if self.rows_per_band < 0:
            raise ValueError(f"the device encoder cannot take {width}x{height} frames (rows wider than 21 666 pixels)")
        self.scratch_bytes = int(self.lib.rt_png_scratch_bytes(self.width, self.height))
        self.capacity = int(self.lib.rt_png_file_bound(self.width, self.height))
        self.scratch = torch.empty(self.scratch_bytes, dtype=torch.uint8, device=torch.device("cuda", self.device))
        self.file = torch.empty(self.capacity, dtype=torch.uint8).pin_memory()
        self.size = torch.zeros(1, dtype=torch.int64).pin_memory()

    def enqueue(self, pixels, stream: int) -> None:
        """``pixels``: uint8 device tensor ``[H * W, 3 or 4]`` (or ``[H, W, C]``), contiguous.  Enqueues the
        encoder on ``stream``; ``finish()`` after synchronising that stream."""
        from .cuda import abi

        channels = int(pixels.shape[-1])
        if pixels.numel() != self.width * self.height * channels or channels not in (3, 4) or not pixels.is_contiguous():
            raise ValueError("expected a contiguous uint8 image of the encoder's size with 3 or 4 channels")
        self.size[0] = -1
        if -(-self.height // self.rows_per_band) >= abi.RT_PNG_ONE_LAUNCH_BANDS:  # (measured: include/rt_abi.h)
            abi.check(self.lib.rt_png_encode_file_dev(self.device, pixels.data_ptr(), channels, self.width * channels, self.width,
                                                      self.height, self.scratch.data_ptr(), self.scratch_bytes, self.file.data_ptr(),
                                                      self.capacity, self.size.data_ptr(), stream), "rt_png_encode_file_dev")
            return
        abi.check(self.lib.rt_png_encode_bands_dev(self.device, pixels.data_ptr(), channels, self.width * channels, self.height,
                                                   self.width, self.height, 0, self.height, 1, self.scratch.data_ptr(),
                                                   self.scratch_bytes, stream), "rt_png_encode_bands_dev")
        abi.check(self.lib.rt_png_pack_dev(self.device, self.width, self.height, self.scratch.data_ptr(), self.file.data_ptr(),
                                           self.capacity, self.size.data_ptr(), stream), "rt_png_pack_dev")

    def finish(self) -> memoryview:
        """The file the last ``enqueue`` made (a view of the pinned buffer: copy or write it before the next one)."""
        n = int(self.size[0])
        if n <= 0:
            raise RuntimeError("the device PNG encoder reported no file (stream not synchronised, or capacity too small)")
        return memoryview(self.file.numpy())[:n]

    def encode(self, pixels) -> bytes:
        import torch

        stream = torch.cuda.current_stream(self.device)
        self.enqueue(pixels, int(stream.cuda_stream))
        stream.synchronize()
        return bytes(self.finish())


def write_png(path: Path, rgb8: np.ndarray, compress_level: Optional[int] = 6, threads: int = 1, pool=None) -> None:
    """``compress_level=None``: the library's own deflate (``encode_png_fast``); an integer: zlib."""
    rgb8 = np.ascontiguousarray(rgb8)
    if compress_level is None:
        Path(path).write_bytes(encode_png_fast(rgb8, threads, pool))
    elif threads > 1 or pool is not None:
        Path(path).write_bytes(encode_png_parallel(rgb8, compress_level, threads, pool))
    else:
        Path(path).write_bytes(encode_png(rgb8, compress_level))


def read_png(path: Path) -> np.ndarray:
    """Decode an 8-bit RGB/RGBA non-interlaced PNG to uint8 ``[H, W, C]`` (all five
    filter types), enough to read back what ``write_png`` or Pillow wrote."""
    data = Path(path).read_bytes()
    if data[:8] != _SIGNATURE:
        raise ValueError(f"{path}: not a PNG")
    pos, idat, width, height, channels = 8, [], 0, 0, 0
    while pos < len(data):
        (length,) = struct.unpack(">I", data[pos:pos + 4])
        tag = data[pos + 4:pos + 8]
        payload = data[pos + 8:pos + 8 + length]
        pos += 12 + length
        if tag == b"IHDR":
            width, height, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", payload)
            if depth != 8 or ctype not in (2, 6) or interlace != 0:
                raise ValueError(f"{path}: only 8-bit RGB/RGBA non-interlaced PNGs are supported")
            channels = 3 if ctype == 2 else 4
        elif tag == b"IDAT":
            idat.append(payload)
        elif tag == b"IEND":
            break
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), dtype=np.uint8).reshape(height, 1 + channels * width)
    out = np.zeros((height, channels * width), dtype=np.uint8)
    bpp = channels
    prev = np.zeros(channels * width, dtype=np.int32)
    for r in range(height):
        ftype, line = int(raw[r, 0]), raw[r, 1:].astype(np.int32)
        if ftype == 0:
            cur = line
        elif ftype == 2:
            cur = (line + prev) & 255
        elif ftype == 1:
            cur = line.reshape(width, bpp).cumsum(axis=0).reshape(-1) & 255
        else:  # 3 (average) and 4 (Paeth) depend on the pixel to the left: sequential
            cur = np.zeros_like(line)
            for i in range(line.size):
                a = int(cur[i - bpp]) if i >= bpp else 0
                b = int(prev[i])
                c = int(prev[i - bpp]) if i >= bpp else 0
                if ftype == 3:
                    pred = (a + b) >> 1
                else:
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pred = a if pa <= pb and pa <= pc else (b if pb <= pc else c)
                cur[i] = (int(line[i]) + pred) & 255
        out[r] = cur
        prev = cur
    return out.reshape(height, width, channels)
