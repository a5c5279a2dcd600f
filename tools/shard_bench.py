"""One rank's share of the 8K frame on ONE GPU (GPU box): rt_render_strips_dev(strip_first = 0, strip_step = N)
into local memory, timed like bench.py (L2 flush, CUDA events).  What the multi-GPU split can reach at best is
the whole frame's time / N; the difference is launch ramp, tail and the ticket schedule.
    python tools/shard_bench.py [N ...]"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi, sharding  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402

lib = abi.load()
w, h = 7680, 4320
flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
dev = torch.device("cuda", 0)
sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
cam = abi.make_camera(flat.eye, w, h)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
stream = torch.cuda.current_stream()
out = []
for n in [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8]:
    for name, flags in (("rgba8 8x4", 0), ("rgb8 8x4", abi.RT_FLAG_RGB8)):
        env = abi.make_env(flat.eye, 1, flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST, cutoff=1e-7, flags=flags,
                           spheres_host=flat.spheres, point_lights_host=flat.point_lights)
        shard = sharding.StripShard(h, n, 0)
        buf = torch.empty(shard.n_rows * w * 4, dtype=torch.uint8, device=dev)
        times = []
        for s in range(25):
            flush.fill_(s & 255)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            abi.check(lib.rt_render_strips_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), shard.strip_rows,
                                               0, n, None, abi.RT_DTYPE_F32, 0, buf.data_ptr(), None, int(stream.cuda_stream)), "strips")
            e1.record()
            torch.cuda.synchronize()
            if s >= 5:
                times.append(e0.elapsed_time(e1))
        out.append((n, name, float(np.mean(times))))
base = {name: t for n, name, t in out if n == 1}
for n, name, t in out:
    print(f"1/{n} of 8K  {name:10s} {t * 1e3:8.1f} us" + (f"   ideal {base[name] / n * 1e3:7.1f} us  efficiency {base[name] / n / t:.3f}" if name in base else ""))
