"""Per-warp start/end stamps of one 1080p launch (needs a -DRT_TIMING_DEBUG build selected
with RT_B200_LIBRARY): when do warps start working, when do they finish.  GPU box only."""
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
w, h = 1920, 1080
flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST, cutoff=1e-7,
                   spheres_host=flat.spheres, point_lights_host=flat.point_lights)
cam = abi.make_camera(flat.eye, w, h)
sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
n = w * h
planes = torch.empty((3, n), dtype=torch.float32, device=dev)
rgba8 = torch.empty((n, 4), dtype=torch.uint8, device=dev)
n_items = ((w + 7) // 8) * ((h + 3) // 4)
stats = torch.zeros(8 + 2 * 148 * 32 + 2 * n_items, dtype=torch.int64, device=dev)
for rep in range(4):
    stats.zero_()
    torch.cuda.synchronize()
    abi.check(lib.rt_render_rows_dev(0, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam), 0, h,
                                     planes.data_ptr(), abi.RT_DTYPE_F32, n, rgba8.data_ptr(), stats.data_ptr(), int(stream.cuda_stream)), "render")
    torch.cuda.synchronize()
raw = stats.cpu().numpy()
t = raw[8:8 + 2 * 148 * 32].reshape(-1, 2)
items = raw[8 + 2 * 148 * 32:].reshape(-1, 2)
t = t[t[:, 0] > 0]
t0 = t[:, 0].min()
start, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
print(f"warps {len(t)}  start: p50 {np.median(start):.1f} max {start.max():.1f} us   end: min {end.min():.1f} p10 {np.percentile(end, 10):.1f} "
      f"p50 {np.median(end):.1f} p90 {np.percentile(end, 90):.1f} p99 {np.percentile(end, 99):.1f} max {end.max():.1f} us")
busy = (end - start).sum() / (len(t) * end.max())
print(f"mean warp residency / kernel span = {busy:.3f}")
hist, edges = np.histogram(end, bins=12)
for c, e in zip(hist, edges):
    print(f"  end >= {e:6.1f} us: {c}")

dur = (items[:, 1] - items[:, 0]) / 1e3
iend = (items[:, 1] - t0) / 1e3
istart = (items[:, 0] - t0) / 1e3
print(f"items {len(dur)}: duration p50 {np.median(dur):.2f} p90 {np.percentile(dur, 90):.2f} p99 {np.percentile(dur, 99):.2f} max {dur.max():.2f} us")
good = dur > 0
print(f"time inside items / warp residency = {dur[good].sum() / (end - start).sum():.3f}   "
      f"(items {dur[good].sum() / len(t):.1f} us per warp, residency {np.mean(end - start):.1f} us, {good.sum() / len(t):.1f} items per warp)")
late = np.argsort(iend)[-12:]
tiles_x = (w + 7) // 8
for i in late:
    print(f"  item {i} (tile row {i // tiles_x}, col {i % tiles_x}): start {istart[i]:.1f} end {iend[i]:.1f} dur {dur[i]:.1f} us")
rows = dur.reshape(-1, tiles_x).mean(axis=1)
print("mean patch duration by tile-row decile:", np.round([rows[int(k * len(rows) / 10):int((k + 1) * len(rows) / 10)].mean() for k in range(10)], 2))
