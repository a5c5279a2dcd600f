"""Kernel time with and without the sphere BVH against scene size (CUDA events, fast policy,
1080p, depth 8): where the tree starts to pay decides RT_ACCEL_AUTO_SPHERES.  Also prints the
host build time of the blob.  Run on the GPU box."""
import ctypes as C
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
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream(0)
W, H, DEPTH = 1920, 1080, 8
n = W * H
planes = torch.empty((3, n), dtype=torch.float32, device=dev)
rgba8 = torch.empty((n, 4), dtype=torch.uint8, device=dev)
sizes = [int(a) for a in sys.argv[1:]] or [2, 7, 15, 31, 63, 127, 255, 1024, 4096, 16384]
for count in sizes:
    if count == 2:
        flat = flatten_scene(scenes.readme_scene(W, H), DefaultMaterial(**scenes.README_GAINS))
    else:
        flat = flatten_scene(scenes.random_spheres_scene(W, H, n_spheres=count), DefaultMaterial(**scenes.README_GAINS))
    cam = abi.make_camera(flat.eye, W, H)
    sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
    plain = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, max_depth=DEPTH,
                         mode=abi.RT_MODE_FAST, cutoff=1e-7)
    tree = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, max_depth=DEPTH,
                        mode=abi.RT_MODE_FAST, cutoff=1e-7)
    capacity = int(lib.rt_accel_bytes(flat.n_spheres))
    blob = np.empty(capacity, dtype=np.uint8)
    used = C.c_int64(0)
    t0 = time.perf_counter()
    for _ in range(5):
        abi.check(lib.rt_accel_build_host(flat.spheres.ctypes.data, flat.n_spheres, C.byref(plain), blob.ctypes.data,
                                          capacity, C.byref(used)), "rt_accel_build_host")
    build_us = (time.perf_counter() - t0) / 5 * 1e6
    blob_dev = torch.from_numpy(blob[: used.value]).to(dev)
    tree.accel_dev, tree.accel_bytes = blob_dev.data_ptr(), used.value
    result = {}
    for name, env in (("off", plain), ("on", tree)):
        times = []
        reps = 6 if count > 2000 and name == "off" else 25
        try:
            for i in range(reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam),
                                                 0, H, planes.data_ptr(), abi.RT_DTYPE_F32, n, rgba8.data_ptr(), None,
                                                 int(stream.cuda_stream)), "render")
                e1.record(stream)
                torch.cuda.synchronize()
                if i >= 3:
                    times.append(e0.elapsed_time(e1))
            result[name] = float(np.median(times))
        except RuntimeError as err:
            result[name] = float("nan")
            print(f"  ({name}: {str(err)[:90]})")
    print(f"{flat.n_spheres:6d} shapes: exhaustive {result['off']:9.3f} ms   tree {result['on']:9.3f} ms   "
          f"ratio {result['off'] / result['on']:6.2f}   build {build_us:8.1f} us   blob {used.value} B", flush=True)
