"""rt_render_frame_host, 8-bit image only, README 1080p into pinned memory: ms per call for several row-chunk
pipelines (RT_HOST_CHUNKS = "chunks,head divisor": the first chunk is rows / divisor).  GPU box only."""
import ctypes as C
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402

lib = abi.load()
w, h = 1920, 1080
flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
env = abi.make_env(flat.eye, 1, flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST, cutoff=1e-7)
cam = abi.make_camera(flat.eye, w, h)
out = torch.empty((w * h, 3), dtype=torch.uint8).pin_memory()
for setting in ["4,16", "2,8", "3,12", "4,32", "6,24", "8,32", "8,64", "12,48", "16,64", "1,2"]:
    os.environ["RT_HOST_CHUNKS"] = setting
    times = []
    for s in range(60):
        t0 = time.perf_counter()
        abi.check(lib.rt_render_frame_host(0, flat.spheres.ctypes.data, flat.n_spheres, flat.point_lights.ctypes.data, C.byref(env),
                                           C.byref(cam), 0, h, None, abi.RT_DTYPE_F32, out.data_ptr(), None), "host")
        _ = int(out[w * h // 2, 1])
        if s >= 10:
            times.append(time.perf_counter() - t0)
    print(f"chunks,head = {setting:6s}: {np.mean(times) * 1e3:.4f} ms  (min {np.min(times) * 1e3:.4f})")
