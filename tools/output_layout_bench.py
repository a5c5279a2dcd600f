"""Trace kernel time by output layout (GPU box): README 1080p, CUDA events, L2 flushed between launches.
    python tools/output_layout_bench.py [width height]"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
lib = abi.load()
dev = torch.device("cuda", 0)
flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
cam = abi.make_camera(flat.eye, w, h)
n = w * h
planes = torch.empty((3, n), dtype=torch.float32, device=dev)
img = torch.empty(n * 4, dtype=torch.uint8, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
stream = int(torch.cuda.current_stream().cuda_stream)


def run(name, want_planes, flags):
    env = abi.make_env(flat.eye, 1, flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST, cutoff=1e-7, flags=flags,
                       spheres_host=flat.spheres, point_lights_host=flat.point_lights)
    times = []
    for i in range(60):
        flush.fill_(i & 255)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), 0, h,
                                         planes.data_ptr() if want_planes else None, abi.RT_DTYPE_F32, n, img.data_ptr(), None, stream), name)
        e1.record()
        torch.cuda.synchronize()
        if i >= 10:
            times.append(e0.elapsed_time(e1))
    print(f"{name:42s} {np.mean(times) * 1e3:7.1f} us")


print(f"{w}x{h}")
run("f32 planes + RGBA8 (the bench's `value`)", True, 0)
run("RGBA8 only", False, 0)
run("RGB8 only, 8x4 patches (the e2e path)", False, abi.RT_FLAG_RGB8)
run("RGB8 only, 32x1 patches, packed words", False, abi.RT_FLAG_RGB8 | abi.RT_FLAG_WIDE_STORES)
run("RGBA8 only, 32x1 patches", False, abi.RT_FLAG_WIDE_STORES)
