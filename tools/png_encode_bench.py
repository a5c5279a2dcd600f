"""Encode stage of the turntable driver alone (no GPU): frames per second of the PNG encoders on
the host cores of this machine, on one rendered 1080p README frame (from the CPU oracle).
  python tools/png_encode_bench.py [--frames 64] [--threads N]"""
import argparse
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from oracle import c_oracle  # noqa: E402  (a tool, not the product: the oracle only supplies the test frame)
from python_ray_tracer_b200 import scenes, turntable  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402


def main() -> None:
    parser = argparse.ArgumentParser()
    parser.add_argument("--frames", type=int, default=64)
    parser.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
    parser.add_argument("--bands", type=int, default=2)
    args = parser.parse_args()
    w, h = 1920, 1080
    flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
    rgb, _ = c_oracle.render(flat, w, h)
    rgba = np.full((w * h, 4), 255, dtype=np.uint8)
    rgba[:, :3] = c_oracle.quantize(rgb)
    print(f"{args.frames} frames of {w}x{h}, {args.threads} threads, {args.bands} bands per frame")
    for label, level in (("library deflate (rt_png_deflate_rows)", None), ("zlib level 1", 1), ("zlib level 6 (Pillow's default)", 6)):
        t1 = time.perf_counter()
        one = turntable.encode_frame(rgba, w, h, level, 1, None)
        t1 = time.perf_counter() - t1
        n = args.frames if level != 6 else max(8, args.frames // 8)
        with ThreadPoolExecutor(max(1, args.threads // args.bands)) as frame_pool, ThreadPoolExecutor(args.threads) as band_pool:
            t0 = time.perf_counter()
            list(frame_pool.map(lambda _: turntable.encode_frame(rgba, w, h, level, args.bands, band_pool), range(n)))
            dt = time.perf_counter() - t0
        print(f"  {label:40s} {len(one) / 1e3:7.1f} KB  one frame on one thread {t1 * 1e3:6.1f} ms   pool: {n / dt:6.1f} frames/s")


if __name__ == "__main__":
    main()
