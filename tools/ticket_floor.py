"""Floor of the work distribution: a 1080p / 8K frame whose every ray misses (all spheres behind
the camera), so an item costs next to nothing and the kernel time is launch + staging + the
ticket draws.  Run on the GPU box."""
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
for w, h in ((1920, 1080), (7680, 4320)):
    for label, shift in (("README scene", 0.0), ("every ray misses", -1000.0)):
        flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
        flat.spheres[2, :2] += shift          # the two small spheres far behind the camera
        flat.spheres[1, 2] -= 1000.0 if shift else 0.0  # and the ground far below
        env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST, cutoff=1e-7)
        cam = abi.make_camera(flat.eye, w, h)
        sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
        n = w * h
        planes = torch.empty((3, n), dtype=torch.float32, device=dev)
        rgba8 = torch.empty((n, 4), dtype=torch.uint8, device=dev)
        stats = torch.zeros(4, dtype=torch.int64, device=dev)
        times = []
        for i in range(40):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), 0, h,
                                             planes.data_ptr(), abi.RT_DTYPE_F32, n, rgba8.data_ptr(), None, int(stream.cuda_stream)), "render")
            e1.record(stream)
            torch.cuda.synchronize()
            if i >= 10:
                times.append(e0.elapsed_time(e1))
        abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), 0, h,
                                         planes.data_ptr(), abi.RT_DTYPE_F32, n, rgba8.data_ptr(), stats.data_ptr(), int(stream.cuda_stream)), "render")
        torch.cuda.synchronize()
        print(f"{w}x{h} {label:18s}: median {np.median(times) * 1e3:8.1f} us   hits {int(stats[1])}   items {((w + 7) // 8) * ((h + 3) // 4)}")
