"""Kernel time against the depth cap at two frame sizes: separates the fixed launch cost
(launch + staging) from the latency of the longest bounce chain.  Run on the GPU box."""
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
stream = torch.cuda.current_stream(0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts = []
for i in range(30):
    e0.record(stream); e1.record(stream); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
print(f"back-to-back events: {np.median(ts) * 1e3:.1f} us")
for w, h in ((32, 32), (1920, 1080)):
    flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
    cam = abi.make_camera(flat.eye, w, h)
    sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
    n = w * h
    planes = torch.empty((3, n), dtype=torch.float32, device=dev)
    rgba8 = torch.empty((n, 4), dtype=torch.uint8, device=dev)
    for depth in (1, 2, 3, 4, 6, 8, 12, 0):
        env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, max_depth=depth,
                           mode=abi.RT_MODE_FAST, cutoff=1e-7)
        times = []
        for i in range(40):
            e0.record(stream)
            abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), 0, h,
                                             planes.data_ptr(), abi.RT_DTYPE_F32, n, rgba8.data_ptr(), None, int(stream.cuda_stream)), "render")
            e1.record(stream)
            torch.cuda.synchronize()
            if i >= 10:
                times.append(e0.elapsed_time(e1))
        print(f"{w}x{h} max_depth={depth or 'none':>4}: median {np.median(times) * 1e3:7.1f} us")
