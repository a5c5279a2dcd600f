"""Animation batches: BASELINE config 5 (a 256-frame turntable, one PNG per frame).

``render_image_pipeline`` (``application.py:43-52``) renders one frame and blocks on its PNG.
For a batch the three stages are overlapped instead: while the GPU traces frame ``k+1``
(``rt_render_rows_dev`` with the fused RGBA8 store), frame ``k``'s 8-bit image is copied to
pinned host memory on a second stream and earlier frames are being PNG-encoded by a pool
of host threads (the library's own deflate, ``rt_png_deflate_rows``: ~10 ms per 1080p frame
per core; ctypes releases the GIL).  The encoder, not the GPU, bounds the frame rate
(~0.1-0.4 s of zlib per 1080p frame per core against ~0.1 ms
of tracing), so frames are sharded over ranks (frame ``k`` -> rank ``k mod world``, one
process per GPU, nothing exchanged) and every rank runs its own encoder pool.

Frames whose PNG already exists are skipped (``resume=True``): the cheap restart the
reference lacks.

    python -m python_ray_tracer_b200.turntable --frames 256 --out frames/ [--width 1920 --height 1080]
    torchrun --nproc-per-node 8 -m python_ray_tracer_b200.turntable --frames 256 --out frames/
"""

from __future__ import annotations

import argparse
import os
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import Callable, Dict, Optional

import numpy as np

from . import scenes
from .domain import Scene3D
from .infrastructure.cuda.sharding import frames_of_rank
from .infrastructure.png import encode_png_fast, encode_png_parallel


def frame_path(out_dir: Path, frame: int) -> Path:
    return Path(out_dir) / f"frame_{frame:04d}.png"


def encode_frame(rgba: np.ndarray, width: int, height: int, compress_level: Optional[int] = None, bands: int = 2,
                 pool=None) -> bytes:
    """PNG bytes of one frame as it comes off the GPU (``[n, 3]`` RGB8 or ``[n, 4]`` RGBA8, row-major).
    ``compress_level=None``: the library's own deflate (``rt_png_deflate_rows``, about four times
    faster than zlib level 1 on rendered frames and smaller; alpha dropped inside the encoder);
    an integer: zlib at that level.  Decoded pixels are the same either way."""
    channels = rgba.shape[-1]
    if compress_level is None:
        return encode_png_fast(rgba.reshape(height, width, channels), bands, pool)
    rgb8 = np.ascontiguousarray(rgba.reshape(-1, channels)[:, :3]).reshape(height, width, 3)
    return encode_png_parallel(rgb8, compress_level, bands, pool)


def frames_to_render(out_dir: Path, n_frames: int, rank: int = 0, world: int = 1, resume: bool = True) -> list:
    """This rank's frames that still have to be rendered.  A frame is done when its final name
    exists: frames are written to ``frame_NNNN.png.tmp`` and renamed, so a run that was killed
    mid-write (or whose encoder raised) leaves no truncated ``frame_NNNN.png`` behind for the next
    resume to count as done.  Stale ``.tmp`` files of this rank's frames are removed."""
    mine = list(frames_of_rank(n_frames, world, rank))
    for k in mine:
        stale = frame_path(out_dir, k).with_suffix(".png.tmp")
        if stale.exists():
            stale.unlink()
    return [k for k in mine if not (resume and frame_path(out_dir, k).exists())]


def _write_atomically(path: Path, data) -> None:
    partial = path.with_suffix(".png.tmp")
    with open(partial, "wb") as f:
        f.write(data)
    os.replace(partial, path)  # atomic within the directory


def _render_frames_device(renderer, scene_of_frame, todo, out_dir: Path, stats: Dict[str, float]) -> Dict[str, float]:
    """The batch with the PNGs made on the GPU (``DevicePngEncoder``): per frame one trace launch (8-bit image
    only), the band encoder and the pack kernel, which writes the ~0.6 MB file into a pinned host buffer.  The GPU
    needs ~0.12 ms per 1080p frame; describing the next scene and writing a file take the host longer than that, so
    the files are written by a small thread pool (the system calls release the GIL) while this thread only describes
    scenes and enqueues launches: four frames in flight -- a frame's file is handed to the pool two iterations after
    its launches, and its slot is taken again two iterations after that."""
    import torch

    from .infrastructure.cuda import abi
    from .infrastructure.png import DevicePngEncoder

    device = renderer.device
    first = scene_of_frame(todo[0]).camera
    width, height = int(first.width), int(first.height)
    stream = torch.cuda.current_stream(device)
    dev = torch.device("cuda", device)
    n_slots, lag = 4, 2
    slots = [{"rgb8": torch.empty((width * height, 3), dtype=torch.uint8, device=dev), "enc": DevicePngEncoder(width, height, device),
              "done": torch.cuda.Event(), "start": torch.cuda.Event(enable_timing=True), "end": torch.cuda.Event(enable_timing=True),
              "frame": None, "written": None} for _ in range(n_slots)]
    failed = []

    def write(frame: int, data) -> float:
        t0 = time.perf_counter()
        _write_atomically(frame_path(out_dir, frame), data)
        return time.perf_counter() - t0

    def hand_over(slot, pool) -> None:
        """The slot's frame is (or is about to be) finished on the GPU: its file goes to the writers."""
        if slot["frame"] is None:
            return
        slot["done"].synchronize()
        stats["gpu_s"] += slot["start"].elapsed_time(slot["end"]) * 1e-3
        slot["written"] = pool.submit(write, slot["frame"], slot["enc"].finish())  # (a view of the slot's pinned buffer)
        slot["frame"] = None

    def release(slot) -> None:
        """The slot's buffers may be used again once its file is on disk."""
        if slot["written"] is None:
            return
        try:
            stats["encode_s"] += slot["written"].result()
            stats["frames"] += 1
        except Exception as exc:  # noqa: BLE001 -- re-raised once every frame that can be written is written
            failed.append(exc)
        slot["written"] = None

    wall0 = time.perf_counter()
    with ThreadPoolExecutor(max(2, min(4, stats.get("encoder_threads", 2)))) as pool:
        for i, frame in enumerate(todo):
            if i >= lag:
                hand_over(slots[(i - lag) % n_slots], pool)
            slot = slots[i % n_slots]
            release(slot)
            scene = scene_of_frame(frame)
            flat = renderer.flatten(scene)
            slot["start"].record(stream)
            renderer.render_rows(flat, scene.camera, 0, height, rgba8=slot["rgb8"], flags=abi.RT_FLAG_RGB8, no_planes=True)
            slot["enc"].enqueue(slot["rgb8"], int(stream.cuda_stream))
            slot["end"].record(stream)
            slot["done"].record(stream)
            slot["frame"] = frame
        for k in range(len(todo), len(todo) + n_slots):  # the frames still in flight, oldest first
            hand_over(slots[(k - lag) % n_slots], pool)
        for slot in slots:
            release(slot)
    abi.check(abi.load().rt_device_status(device), "rt_device_status")
    stats["wall_s"] = time.perf_counter() - wall0
    stats["encoder"] = "device (rt_png_encode_bands_dev + rt_png_pack_dev); encode_s = file writes (on a thread pool)"
    if failed:
        raise failed[0]
    return stats


def render_frames(renderer, scene_of_frame: Callable[[int], Scene3D], n_frames: int, out_dir: Path, *,
                  rank: int = 0, world: int = 1, encoder_threads: Optional[int] = None,
                  compress_level: Optional[int] = None, resume: bool = True, bands_per_frame: int = 2,
                  encoder: str = "device") -> Dict[str, float]:
    """Render this rank's frames of an ``n_frames`` batch to ``out_dir/frame_NNNN.png``.

    ``scene_of_frame(k)`` builds frame ``k``'s ``Scene3D`` (``scenes.turntable_scene`` for the
    BASELINE turntable).  ``encoder``: ``"device"`` (default) makes the PNGs on the GPU;
    ``"host"`` copies the 8-bit images to the host and deflates them on a thread pool
    (``compress_level``: see ``encode_frame``; a zlib level implies the host).  Returns timing
    totals (seconds) and the number of frames written.
    """
    import torch

    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    device = renderer.device
    todo = frames_to_render(out_dir, n_frames, rank, world, resume)
    threads = encoder_threads or max(1, min(len(os.sched_getaffinity(0)), (os.cpu_count() or 2) // max(1, world)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 2) // max(1, world))
    stats = {"frames": 0, "skipped": len(list(frames_of_rank(n_frames, world, rank))) - len(todo),
             "wall_s": 0.0, "gpu_s": 0.0, "encode_s": 0.0, "encoder_threads": threads}
    if not todo:
        return stats
    first_camera = scene_of_frame(todo[0]).camera
    if encoder == "device" and compress_level is None and 1 + 3 * int(first_camera.width) <= 65000:
        return _render_frames_device(renderer, scene_of_frame, todo, out_dir, stats)

    render_stream = torch.cuda.current_stream(device)
    copy_stream = torch.cuda.Stream(device)
    first = scene_of_frame(todo[0]).camera
    width, height = int(first.width), int(first.height)
    n = width * height
    dev = torch.device("cuda", device)
    # two frames in flight: the device 8-bit image (3 bytes per pixel, what save_image hands to its
    # encoder, base.py:143-151; no linear planes are written) and its pinned host copy
    from .infrastructure.cuda import abi

    slots = []
    for _ in range(2):
        slots.append({
            "rgba8": torch.empty((n, 3), dtype=torch.uint8, device=dev),
            "host": torch.empty((n, 3), dtype=torch.uint8).pin_memory(),
            "rendered": torch.cuda.Event(enable_timing=True), "started": torch.cuda.Event(enable_timing=True),
            "copied": torch.cuda.Event(), "frame": None, "pending": None,
        })

    def encode_and_write(frame: int, rgba: np.ndarray) -> float:
        t0 = time.perf_counter()
        data = encode_frame(rgba, width, height, compress_level, bands_per_frame, band_pool)  # off the main thread
        _write_atomically(frame_path(out_dir, frame), data)
        return time.perf_counter() - t0

    def retire(slot) -> None:
        """Frame in ``slot`` has been copied: hand its pixels to the encoder pool."""
        if slot["frame"] is None:
            return
        slot["copied"].synchronize()
        stats["gpu_s"] += slot["started"].elapsed_time(slot["rendered"]) * 1e-3
        rgba = slot["host"].numpy().copy()  # one contiguous memcpy frees the pinned slot for the next copy
        while sum(1 for f in futures if not f.done()) > 2 * threads:  # the GPU is far ahead of zlib:
            time.sleep(0.001)                                          # bound the frames held in memory
        futures.append(frame_pool.submit(encode_and_write, slot["frame"], rgba))
        slot["frame"] = None

    wall0 = time.perf_counter()
    futures = []
    with ThreadPoolExecutor(max(1, threads // max(1, bands_per_frame))) as frame_pool, \
            ThreadPoolExecutor(max(1, threads)) as band_pool:
        for i, frame in enumerate(todo):
            slot = slots[i % 2]
            retire(slot)  # the frame rendered two iterations ago
            scene = scene_of_frame(frame)
            flat = renderer.flatten(scene)
            slot["started"].record(render_stream)
            renderer.render_rows(flat, scene.camera, 0, height, rgba8=slot["rgba8"], flags=abi.RT_FLAG_RGB8, no_planes=True)
            slot["rendered"].record(render_stream)
            copy_stream.wait_event(slot["rendered"])
            with torch.cuda.stream(copy_stream):
                slot["host"].copy_(slot["rgba8"], non_blocking=True)
                slot["copied"].record(copy_stream)
            slot["frame"] = frame
        for slot in slots:
            retire(slot)
        failed = None
        for f in futures:  # every frame that can be written is written before an encoder error surfaces
            try:
                stats["encode_s"] += f.result()
                stats["frames"] += 1
            except Exception as exc:  # noqa: BLE001 -- re-raised below
                failed = failed or exc
        if failed is not None:
            raise failed
    stats["wall_s"] = time.perf_counter() - wall0
    return stats


def render_turntable(renderer, n_frames: int, out_dir: Path, width: int = 1920, height: int = 1080, **kw) -> Dict[str, float]:
    """BASELINE config 5: frame ``k`` is the README scene with the two small spheres and the
    point light rotated by ``2 pi k / n_frames`` about the vertical axis through (0, ., 2)
    (the reference camera has no orientation, ``base.py:130-141``, so the scene turns)."""
    return render_frames(renderer, lambda k: scenes.turntable_scene(k, n_frames, width, height), n_frames, out_dir, **kw)


def main() -> None:
    parser = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    parser.add_argument("--frames", type=int, default=256)
    parser.add_argument("--out", type=Path, required=True)
    parser.add_argument("--width", type=int, default=1920)
    parser.add_argument("--height", type=int, default=1080)
    parser.add_argument("--compress-level", type=int, default=None,
                        help="zlib level for the PNGs; default: the library's own faster deflate (rt_png_deflate_rows)")
    parser.add_argument("--encoder-threads", type=int, default=None)
    parser.add_argument("--encoder", choices=["device", "host"], default="device",
                        help="device: PNGs made on the GPU (default); host: 8-bit images deflated on host threads")
    parser.add_argument("--no-resume", action="store_true")
    args = parser.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    from .infrastructure import CudaRenderer
    from .infrastructure.cuda import affinity

    if world > 1:
        affinity.bind_to_device(local)  # encoder threads and pinned slots on the GPU's own socket
    renderer = CudaRenderer(**scenes.README_GAINS, device=local, output_dtype="float32")
    stats = render_turntable(renderer, args.frames, args.out, args.width, args.height, rank=rank, world=world,
                             encoder_threads=args.encoder_threads, compress_level=args.compress_level,
                             resume=not args.no_resume, encoder=args.encoder)
    fps = stats["frames"] / stats["wall_s"] if stats["wall_s"] else 0.0
    print(f"rank {rank}/{world}: {stats['frames']} frames written ({stats['skipped']} skipped) in {stats['wall_s']:.2f} s "
          f"= {fps:.1f} frames/s; GPU {stats['gpu_s'] * 1e3:.1f} ms total, "
          + (f"PNGs made on the device, {stats['encode_s']:.2f} s of file writes" if "encoder" in stats else
             f"encoder {stats['encode_s']:.2f} s over {stats['encoder_threads']} threads"))


if __name__ == "__main__":
    main()
