"""Fixed cost of a camera launch (GPU box): frames from 32x32 up, L2 flushed or not, CUDA events around one launch.
    python tools/launch_floor.py"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402

lib = abi.load()
dev = torch.device("cuda", 0)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
stream = torch.cuda.current_stream()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
times = []
for _ in range(20):
    e0.record(); e1.record(); torch.cuda.synchronize(); times.append(e0.elapsed_time(e1))
print(f"two events back to back: {np.median(times) * 1e3:.1f} us")
for w, h in ((32, 32), (256, 128), (960, 540), (1920, 1080)):
    flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
    sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
    cam = abi.make_camera(flat.eye, w, h)
    buf = torch.empty(w * h * 4, dtype=torch.uint8, device=dev)
    for depth in (1, 0):
        env = abi.make_env(flat.eye, 1, flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST, cutoff=1e-7, max_depth=depth,
                           spheres_host=flat.spheres, point_lights_host=flat.point_lights)
        for flushed in (True, False):
            times = []
            for s in range(30):
                if flushed:
                    flush.fill_(s & 255)
                e0.record()
                abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), 0, h, None,
                                                 abi.RT_DTYPE_F32, 0, buf.data_ptr(), None, int(stream.cuda_stream)), "rows")
                e1.record()
                torch.cuda.synchronize()
                if s >= 5:
                    times.append(e0.elapsed_time(e1))
            print(f"{w}x{h} depth {'all' if depth == 0 else depth} {'L2 flushed' if flushed else 'warm      '}: {np.median(times) * 1e3:7.1f} us")
