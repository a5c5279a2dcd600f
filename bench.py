#!/usr/bin/env python
"""Benchmark of the hot path: ``NumpyRenderer.raytrace_scene`` replaced by the sm_100a trace kernel.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload c2|c1|c3|c4|c5] [--precision fast|exact]

A step is one frame of the workload through the hot path (camera rays -> nearest hit ->
shadow ray -> shading -> bounce chain -> framebuffer stores).  The default workload is
BASELINE.json's ``configs[1]`` (README scene, 1920x1080, all effects); with N > 1 every rank
renders its own such frame per step (frame-sharded, no data-path collective, weak scaling).
One JSON line is printed by rank 0.  Besides the contract's keys it carries

  ``c4``        BASELINE configs[3]: ONE 7680x4320 frame over the N ranks (interleaved 16-row strips, the
                kernels store the 8-bit frame into rank 0's framebuffer over NVLink peer memory), with
                the same frame on rank 0 alone timed beside it and the two frames compared;
  ``pipeline``  ``render_image_pipeline(scene, path, CudaRenderer())`` at 1080p, PNG written, on rank 0;
  ``c3``        BASELINE configs[2] (1 025 shapes, 4K, depth 8) through the sphere tree, rank 0, N = 1.

Metric: Mrays/s = primary rays (= pixels) per second / 1e6; ``ms_per_step`` is ms/frame.
``value`` is device time with the scene resident in HBM (CUDA events on the launch stream,
L2 flushed between steps, max over ranks).  ``e2e`` is the same frame through the host-buffer
C-ABI entry point ``rt_render_png_host``: kernel, PNG encoder on the device, the finished file
``save_image`` writes (base.py:143-151) in pinned host memory, every step.  Beside it: the 8-bit
image ``save_image`` consumes brought back instead (``rgb8_image``, ``rt_render_frame_host``; PCIe-bound,
and bound by the host when several ranks copy at once) and the linear float planes.
"""

from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402

METRIC = "Mrays/s (primary rays = pixels per second; ms/frame in ms_per_step), README scene @1080p"
UNIT = "Mrays/s"
README_DEFAULT = DefaultMaterial(**scenes.README_GAINS)

# Algorithmic work per unit (restated in DESIGN.md):
FLOP_PER_SPHERE_TEST = 24.0   # one ray/sphere quadratic
FLOP_PER_HIT = 280.0          # NumpyShader.create without its S shadow tests
BYTES_PER_PIXEL = 12.0 + 4.0  # planar f32 RGB + RGBA8, each written once
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from this round's
# `ncu --set full` captures: profiles/r02/ncu_traffic.json names the capture each figure comes from
TRAFFIC_FILE = ROOT / "profiles" / "r02" / "ncu_traffic.json"


def scene_of(name: str, rank: int = 0, step: int = 0, world: int = 1):
    """(Scene3D, width, height, max_depth) of one step on one rank."""
    if name == "c2":
        return scenes.readme_scene(1920, 1080), 1920, 1080, 0
    if name == "c5":
        return scenes.turntable_scene((step * world + rank) % 256, 256, 1920, 1080), 1920, 1080, 0
    if name == "c3":
        return scenes.random_spheres_scene(3840, 2160), 3840, 2160, 8
    if name == "c4":
        return scenes.readme_scene(7680, 4320), 7680, 4320, 0
    if name == "c1":
        return scenes.readme_scene(960, 540), 960, 540, 0
    raise SystemExit(f"unknown workload {name}")


LABELS = {
    "c1": "README scene 960x540 (configs[0])",
    "c2": "README scene 1920x1080, all effects (configs[1])",
    "c3": "1024 random spheres + checkered ground, 4 PointLights (reference semantics: light 0 shades), depth 8, 3840x2160 (configs[2])",
    "c4": "README scene 7680x4320, one frame row-band sharded over the ranks, assembled on rank 0 (configs[3])",
    "c5": "256-frame turntable of the README scene at 1080p, frame-sharded (configs[4])",
}


def workload(name: str, rank: int, step: int, world: int):
    """(label, flat scene, width, height, max_depth) of one step on one rank."""
    scene, width, height, max_depth = scene_of(name, rank, step, world)
    return LABELS[name], flatten_scene(scene, README_DEFAULT), width, height, max_depth


def config_of(name: str, world: int) -> dict:
    """The configuration BOTH arms measure (the reference arm prints the same object)."""
    _, width, height, max_depth = scene_of(name)
    flat = workload(name, 0, 0, 1)[1]
    return {"workload": LABELS[name], "width": width, "height": height, "spheres": flat.n_spheres,
            "max_depth": max_depth or None,
            "frames_per_step": 1 if name == "c4" else world,
            "sharding": ("one frame over %d rank(s)" % world) if name == "c4" else ("one frame per rank per step, %d rank(s)" % world)}


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int, period: float = 0.001) -> None:
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_event = threading.Event()
        self.nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:  # NVML missing: clocks are reported as unknown
            self.nvml = None

    def sample(self) -> None:
        if self.nvml is None:
            return
        try:
            self.samples.append(int(self.nvml.nvmlDeviceGetClockInfo(self.handle, self.nvml.NVML_CLOCK_SM)))
            mask = int(self.nvml.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
            for bit, name in self.REASONS.items():
                if mask & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self) -> None:
        while not self._stop_event.is_set():
            self.sample()
            self._stop_event.wait(self.period)

    def mark(self) -> int:
        return len(self.samples)

    def summary(self, first: int = 0, last=None) -> dict:
        window = self.samples[first:last]
        return {"sm_mhz": int(np.median(window)) if window else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(window)}

    def finish(self) -> None:
        self._stop_event.set()
        self.join(timeout=2.0)


def measured_peaks() -> dict:
    path = ROOT / "MEASURED_PEAKS.json"
    if path.exists():
        data = json.loads(path.read_text())
        return {"hbm_gbs": float(data["hbm_gbs"]), "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def ncu_traffic(name: str):
    if TRAFFIC_FILE.exists():
        entry = json.loads(TRAFFIC_FILE.read_text()).get(name)
        if entry:
            return entry.get("dram_bytes_per_launch"), entry.get("capture")
    return None, None


# --------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/ code: as the timed baseline)
# --------------------------------------------------------------------------------------
def _rows_job(args):
    flat, width, height, rows, max_depth, repeats = args
    from oracle import numpy_port

    pixels = (rows[:, None] * width + np.arange(width)[None, :]).reshape(-1)
    total = 0.0
    for _ in range(repeats):
        rgb, counters = numpy_port.render(flat, width, height, pixels=pixels, max_depth=max_depth)
        total += float(rgb.sum())
    return total, counters[0]


def sample_rows(height: int, rows_wanted: int) -> np.ndarray:
    """Every k-th row of the frame (sky, spheres and ground in the frame's own proportions)."""
    stride = max(1, -(-height // max(1, rows_wanted)))
    return np.arange(stride // 2, height, stride)


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_single_core(flat, width, height, max_depth, budget_s: float = 20.0) -> dict:
    """The NumPy port on one core (the reference is single-threaded: elementwise ufuncs, no
    BLAS) on a bounded sample: whole frames of the workload until ``budget_s`` is used, or
    a centred row band if one frame would not fit."""
    from oracle import numpy_port

    # the README scene runs at ~0.5 Mpx/s; large scenes are slower per pixel -> fewer rows
    est = width * height / 0.5e6 * max(1.0, flat.n_spheres / 3.0) ** 2
    rows = sample_rows(height, height if est <= budget_s else max(1, int(height * budget_s / est)))
    pixels = None if len(rows) == height else (rows[:, None] * width + np.arange(width)[None, :]).reshape(-1)
    n = width * len(rows)
    t0 = time.perf_counter()
    frames = 0
    while True:
        numpy_port.render(flat, width, height, pixels=pixels, max_depth=max_depth)
        frames += 1
        elapsed = time.perf_counter() - t0
        if elapsed + elapsed / frames > budget_s or frames >= 5:
            break
    what = f"{frames} x {'full frame' if pixels is None else f'every {rows[1] - rows[0] if len(rows) > 1 else height}th row of the frame'} ({n} pixels each) in {elapsed:.1f} s"
    return {"value": n * frames / elapsed / 1e6, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": what + "; oracle/numpy_port.py = the reference's whole-array NumPy algorithm on the flattened scene"}


def run_reference_arm(args) -> None:
    """``--impl reference``: the reference's CPU algorithm (NumPy port; the reference tree
    itself does not exist on the GPU box) on every host core: the frame is cut into row
    bands (bit-identical to the full frame) and rendered by a process pool.  A step is the
    whole frame rendered ``repeats`` times, sized so that a step takes about 2.5 s (the
    single-frame steps of round 1 took 0.12 s and their rate moved by +-15 % from run to run)
    and the whole run a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp

    world = max(1, args.gpus)
    label, flat, width, height, max_depth = workload(args.workload, 0, 0, world)
    cores = host_cores()
    # rows: the whole frame when the scene is small enough, evenly spaced rows otherwise
    target_px = int(min(width * height, 0.4e6 * 2.0 * cores / max(1.0, flat.n_spheres / 3.0) ** 2))
    rows = sample_rows(height, max(cores, min(height, target_px // width)))
    n = width * len(rows)
    total_steps = args.warmup + args.steps
    step_budget = min(3.0, max(0.5, 150.0 / max(1, total_steps)))
    with mp.get_context("fork").Pool(cores) as pool:
        jobs = [(flat, width, height, rows[j::cores], max_depth, 1) for j in range(cores) if len(rows[j::cores])]
        t0 = time.perf_counter()
        pool.map(_rows_job, jobs, chunksize=1)  # calibration (untimed): one pass over the sample
        once = time.perf_counter() - t0
        repeats = max(1, int(round(step_budget / max(once, 1e-3))))
        jobs = [j[:5] + (repeats,) for j in jobs]
        times = []
        for step in range(total_steps):
            t0 = time.perf_counter()
            pool.map(_rows_job, jobs, chunksize=1)
            dt = time.perf_counter() - t0
            if step >= args.warmup:
                times.append(dt)
    sec = float(np.mean(times))
    value = n * repeats / sec / 1e6
    sample = (f"{repeats} x {len(rows)} evenly spaced rows of {width}x{height} ({n} pixels) per step, interleaved over {len(jobs)} processes; "
              "oracle/numpy_port.py = the reference's whole-array NumPy algorithm on the flattened scene")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "ms_per_frame_extrapolated": width * height / (value * 1e6) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(args.workload, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
class Gpu:
    """What every part of the GPU arm needs: the library, this rank's device and stream."""

    def __init__(self, args) -> None:
        import torch
        import torch.distributed as dist

        from python_ray_tracer_b200.infrastructure.cuda import abi, affinity, sharding

        self.torch, self.dist, self.abi, self.sharding = torch, dist, abi, sharding
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.bound = affinity.bind_to_device(self.local) if self.world > 1 else None  # pinned buffers next to this rank's GPU
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.lib = abi.load()
        self.stream = torch.cuda.current_stream(self.local)
        self.sptr = int(self.stream.cuda_stream)
        self.mode = abi.RT_MODE_FAST if args.precision == "fast" else abi.RT_MODE_EXACT
        self.flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.launches = 0  # kernels of librt_b200.so launched inside timed regions

    def barrier(self) -> None:
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.local)

    def max_over_ranks(self, value: float) -> float:
        t = self.torch.tensor([float(value)], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def gather_floats(self, value: float) -> list:
        t = self.torch.tensor([float(value)], dtype=self.torch.float64, device=self.dev)
        if self.world == 1:
            return [float(value)]
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(v.item()) for v in out]

    def env_of(self, flat, max_depth, flags=0):
        abi = self.abi
        return abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, max_depth=max_depth,
                            mode=self.mode, cutoff=self.args.cutoff if self.mode == abi.RT_MODE_FAST else 0.0, flags=flags,
                            spheres_host=flat.spheres, point_lights_host=flat.point_lights)

    def tree_of(self, flat, env):
        """rt_accel_build_host + upload; names the blob in env.  Returns the device tensor (keep it alive)."""
        abi, lib = self.abi, self.lib
        capacity = int(lib.rt_accel_bytes(flat.n_spheres))
        blob = np.empty(capacity, dtype=np.uint8)
        used = C.c_int64(0)
        abi.check(lib.rt_accel_build_host(flat.spheres.ctypes.data, flat.n_spheres, C.byref(env), blob.ctypes.data,
                                          capacity, C.byref(used)), "rt_accel_build_host")
        dev_blob = self.torch.from_numpy(blob[: used.value]).to(self.dev)
        env.accel_dev, env.accel_bytes = dev_blob.data_ptr(), dev_blob.numel()
        return dev_blob

    def timed_steps(self, launch, steps: int, warmup: int, together: bool = False) -> np.ndarray:
        """``warmup`` untimed launches, then ``steps`` launches each between its own CUDA event pair
        with the L2 flushed before it; per-step ms of THIS rank.  ``together``: all ranks start each
        step after a barrier (one frame over all ranks)."""
        torch = self.torch
        for s in range(warmup):
            launch(s)
        self.barrier()
        events = []
        for s in range(warmup, warmup + steps):
            self.flush.fill_(s & 0xFF)  # evict the previous frame's lines from L2 (not timed)
            if together and self.world > 1:
                # a rank still flushing would compete with the others' peer stores into rank 0
                self.dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self.stream)
            launch(s)
            self.launches += 1
            e1.record(self.stream)
            events.append((e0, e1))
        self.barrier()
        return np.array([a.elapsed_time(b) for a, b in events])


def device_frames(g: Gpu, name: str, use_tree: bool):
    """Scene tables of every step resident in HBM, and launch(step, stats) for workload ``name``
    (one frame per rank per step; rows_dev entry point, planar f32 + RGBA8 outputs)."""
    torch, abi, lib, args = g.torch, g.abi, g.lib, g.args
    total = args.warmup + args.steps
    distinct = total if name == "c5" else 1
    tables, keep = [], []
    for s in range(distinct):
        _, flat, width, height, max_depth = workload(name, g.rank, s, g.world)
        env = g.env_of(flat, max_depth)
        if use_tree:
            keep.append(g.tree_of(flat, env))
        tables.append((torch.from_numpy(flat.spheres).to(g.dev), torch.from_numpy(flat.point_lights).to(g.dev), env, flat))
    n = width * height
    cam = abi.make_camera(tables[0][3].eye, width, height)
    planes = torch.empty((3, n), dtype=torch.float32, device=g.dev)
    rgba8 = torch.empty((n, 4), dtype=torch.uint8, device=g.dev)
    stats = torch.zeros(4, dtype=torch.int64, device=g.dev)

    def launch(step: int, with_stats: bool = False) -> None:
        spheres, lights, env, flat = tables[step % distinct]
        abi.check(lib.rt_render_rows_dev(g.local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                         C.byref(cam), 0, height, planes.data_ptr(), abi.RT_DTYPE_F32, n,
                                         rgba8.data_ptr(), stats.data_ptr() if with_stats else None, g.sptr),
                  "rt_render_rows_dev")

    def counters() -> dict:
        stats.zero_()
        launch(total - 1, with_stats=True)
        torch.cuda.synchronize(g.local)
        traced, hits, deepest, tests = (int(v) for v in stats.cpu().tolist())
        return {"rays_traced": traced, "hits_shaded": hits, "generations": deepest, "sphere_tests": tests}

    return launch, counters, tables, cam, width, height, (planes, rgba8, keep)


def host_path(g: Gpu, tables, cam, width, height, use_tree: bool) -> dict:
    """End to end through rt_render_frame_host with pinned host outputs, every step: scene upload (and the
    tree build for large scenes), kernel, copy to the host, one value read on the host."""
    torch, abi, lib, args = g.torch, g.abi, g.lib, g.args
    n = width * height
    rgb_host = torch.empty((3, n), dtype=torch.float32).pin_memory()
    rgb8_host = torch.empty((n, 3), dtype=torch.uint8).pin_memory()
    total = args.warmup + args.steps
    envs = []
    for _, _, env, flat in tables:  # the host entry point builds its own tree (or none: --accel off)
        e = g.env_of(flat, env.max_depth, 0 if use_tree else abi.RT_FLAG_NO_ACCEL)
        envs.append(e)

    def one_pass(with_rgb8: bool, with_planes: bool) -> float:
        times = []
        for s in range(total):
            env, flat = envs[s % len(envs)], tables[s % len(tables)][3]
            t0 = time.perf_counter()
            abi.check(lib.rt_render_frame_host(g.local, flat.spheres.ctypes.data, flat.n_spheres, flat.point_lights.ctypes.data,
                                               C.byref(env), C.byref(cam), 0, height,
                                               rgb_host.data_ptr() if with_planes else None, abi.RT_DTYPE_F32,
                                               rgb8_host.data_ptr() if with_rgb8 else None, None), "rt_render_frame_host")
            # the result is read on the host inside the timed region
            _ = (float(rgb_host[0, n // 2]) + float(rgb_host[2, n - 1])) if with_planes else int(rgb8_host[n // 2, 1])
            if s >= args.warmup:
                times.append(time.perf_counter() - t0)
        return g.max_over_ranks(float(np.mean(times)) * 1e3)

    rgb8_ms = one_pass(True, False)
    planes_ms = one_pass(False, True)
    both_ms = one_pass(True, True)
    # save_image's FILE made on the device (rt_render_png_host): only the PNG crosses PCIe
    png_cap = int(lib.rt_png_file_bound(width, height))
    png_ms, png_bytes = None, None
    if png_cap > 0:
        png_host = torch.empty(png_cap, dtype=torch.uint8).pin_memory()
        used = C.c_int64(0)
        times = []
        for s in range(total):
            env, flat = envs[s % len(envs)], tables[s % len(tables)][3]
            t0 = time.perf_counter()
            abi.check(lib.rt_render_png_host(g.local, flat.spheres.ctypes.data, flat.n_spheres, flat.point_lights.ctypes.data,
                                             C.byref(env), C.byref(cam), png_host.data_ptr(), png_cap, C.byref(used)), "rt_render_png_host")
            _ = int(png_host[used.value - 1])  # the file's last byte, read on the host
            if s >= args.warmup:
                times.append(time.perf_counter() - t0)
        png_ms, png_bytes = g.max_over_ranks(float(np.mean(times)) * 1e3), int(used.value)
    flat = tables[-1][3]
    px = width * height * g.world
    rate = lambda ms: px / (ms * 1e-3) / 1e6  # noqa: E731
    table_bytes = int(flat.spheres.nbytes + flat.point_lights.nbytes)
    rgb8 = {
        "value": rate(rgb8_ms), "unit": UNIT, "ms_per_step": rgb8_ms, "h2d_bytes_per_step": table_bytes,
        "d2h_bytes_per_step": int(rgb8_host.numel()), "d2h_gb_per_s_per_gpu": rgb8_host.numel() / (rgb8_ms * 1e-3) / 1e9,
        "api": "rt_render_frame_host(rgb8_host): kernel (8-bit image quantised by the trace kernel, no linear planes written) + D2H "
               "of the RGB8 image save_image consumes (base.py:143-151) into pinned host memory, 8 row chunks pipelined",
    }
    extras = {
        "host_cpus_bound": len(g.bound) if g.bound else None,
        # what raytrace_scene returns to a caller that reads .x/.y/.z: the linear planes
        "linear_planes_f32": {"ms_per_step": planes_ms, "value": rate(planes_ms), "d2h_bytes_per_step": int(rgb_host.numel() * 4),
                              "d2h_gb_per_s_per_gpu": rgb_host.numel() * 4 / (planes_ms * 1e-3) / 1e9},
        "planes_and_rgb8": {"ms_per_step": both_ms, "value": rate(both_ms),
                            "d2h_bytes_per_step": int(rgb_host.numel() * 4 + rgb8_host.numel())},
        "d2h_note": "with N ranks copying at once this box moves 130-170 GB/s to the host in all (plain cudaMemcpyAsync, no kernel: "
                    "profiles/r02/d2h_aggregate_n8.txt), so paths that bring pixels back stop scaling; the PNG path brings back a tenth",
    }
    if png_ms is None:  # rows too wide for the device encoder: the 8-bit image is the headline
        return {**rgb8, **extras, "png_file": None}
    # Headline: save_image's FILE in host memory -- the end of the reference's pipeline (application.py:43-52), one step
    # further than the 8-bit image it is made from: traced, quantised, filtered, deflated and packed on the device.
    return {
        "value": rate(png_ms), "unit": UNIT, "ms_per_step": png_ms, "h2d_bytes_per_step": table_bytes,
        "d2h_bytes_per_step": png_bytes,
        "api": "rt_render_png_host: trace kernel (8-bit image, stays on the device) + rt_png_encode_bands_dev + rt_png_pack_dev "
               "writing the finished PNG into the caller's pinned buffer; the file's last byte read on the host",
        "rgb8_image": rgb8,
        **extras,
    }


def pipeline_object(g: Gpu, frames: int = 8) -> dict:
    """``render_image_pipeline(scene, path, CudaRenderer())`` at 1080p with the PNG written: the
    reference's only timed entry point (main.py:55-57, application.py:43-52)."""
    from python_ray_tracer_b200 import render_image_pipeline
    from python_ray_tracer_b200.infrastructure import CudaRenderer

    scene = scenes.readme_scene(1920, 1080)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for label, kw in (("library_deflate", {}), ("zlib6_like_pillow", {"png_compress_level": 6, "png_threads": 1})):
            renderer = CudaRenderer(**scenes.README_GAINS, device=g.local, **kw)
            n = frames if label == "library_deflate" else 2
            walls, trace_d2h, png = [], [], []
            for i in range(n + 1):
                path = Path(tmp) / f"{label}_{i}.png"
                t0 = time.perf_counter()
                render_image_pipeline(scene, path, renderer)
                dt = time.perf_counter() - t0
                if i:  # the first call creates buffers, the pinned landing zone and the encoder pool
                    walls.append(dt * 1e3)
                    trace_d2h.append(renderer.last_timing["kernel_and_d2h_ms"])
                    png.append(renderer.last_timing["png_ms"])
            size = path.stat().st_size
            out[label] = {"ms_per_frame": float(np.mean(walls)), "flatten_launch_kernel_d2h_ms": float(np.mean(walls) - np.mean(png)),
                          "kernel_and_d2h_ms": float(np.mean(trace_d2h)), "png_ms": float(np.mean(png)),
                          "encoder": renderer.last_timing["encoder"], "png_bytes": int(size), "frames": n}
    lib_d = out["library_deflate"]
    return {"api": "render_image_pipeline(readme_scene(1920, 1080), path, CudaRenderer(**README gains)): get_ray_directions -> "
                   "raytrace_scene (8-bit image only, linear planes lazy) -> save_image (pinned D2H + PNG to a tmpfs file)",
            "ms_per_frame": lib_d["ms_per_frame"], "png_ms": lib_d["png_ms"], "d2h_ms": None,
            "kernel_ms": None, "without_png_ms": lib_d["flatten_launch_kernel_d2h_ms"], "encoder": lib_d["encoder"],
            "variants": out}


def c4_object(g: Gpu, steps: int, warmup: int) -> dict:
    """BASELINE configs[3]: one 7680x4320 README frame over all ranks, assembled on rank 0."""
    torch, abi, lib, sharding, args = g.torch, g.abi, g.lib, g.sharding, g.args
    _, flat, width, height, max_depth = workload("c4", 0, 0, g.world)
    world, rank = g.world, g.rank
    layout = args.peer_layout if args.peer_layout != "auto" else ("rgb8-wide" if world >= 8 else "rgba8")
    channels = 3 if layout == "rgb8-wide" else 4
    flags = abi.RT_FLAG_IMAGE_ORDER | ((abi.RT_FLAG_WIDE_STORES | abi.RT_FLAG_RGB8 | abi.RT_FLAG_EVEN_STORES) if layout == "rgb8-wide" else 0)
    env = g.env_of(flat, max_depth, flags)
    cam = abi.make_camera(flat.eye, width, height)
    spheres, lights = torch.from_numpy(flat.spheres).to(g.dev), torch.from_numpy(flat.point_lights).to(g.dev)
    shard = sharding.StripShard(height, world, rank)
    peer = sharding.PeerFrame(height, width, rank, world, g.local, dst=0, channels=channels)

    def launch_shard(_step=0) -> None:
        # the exchange step is inside the kernel: stores go to rank 0's framebuffer
        abi.check(lib.rt_render_strips_dev(g.local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                           C.byref(cam), shard.strip_rows, rank, world, None, abi.RT_DTYPE_F32, 0,
                                           peer.ptr, None, g.sptr), "rt_render_strips_dev")

    # ---- the single-GPU frame (rank 0 alone, its fastest 8-bit layout: RGBA8, 8x4 patches) and the comparison
    solo_ms, matches = None, None
    solo_env = g.env_of(flat, max_depth, abi.RT_FLAG_IMAGE_ORDER)
    solo = torch.empty((height, width, 4), dtype=torch.uint8, device=g.dev) if rank == 0 else None

    def launch_solo(_step=0) -> None:
        abi.check(lib.rt_render_strips_dev(g.local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(solo_env),
                                           C.byref(cam), shard.strip_rows, 0, 1, None, abi.RT_DTYPE_F32, 0,
                                           solo.data_ptr(), None, g.sptr), "rt_render_strips_dev")

    if rank == 0:
        launch_solo()
    launch_shard()
    frame = peer.complete()
    if rank == 0:
        matches = bool(torch.equal(frame[:, :, :3], solo[:, :, :3]))
    # every rank times its shard; then rank 0 times the whole frame alone while the others wait
    shard_ms = g.timed_steps(launch_shard, steps, warmup, together=True)
    mine = float(shard_ms.mean())
    per_rank = g.gather_floats(mine)
    ms_per_step = max(per_rank)
    if rank == 0:
        was = g.world
        g.world = 1  # no barriers: the other ranks are parked at the barrier below
        solo_ms = float(g.timed_steps(launch_solo, steps, warmup).mean())
        g.world = was
    g.barrier()

    # ---- end to end: all ranks render, rank 0 copies the assembled frame to pinned host memory
    host = torch.empty((height, width, channels), dtype=torch.uint8).pin_memory() if rank == 0 else None
    e2e = []
    for s in range(warmup + steps):
        g.barrier()
        t0 = time.perf_counter()
        launch_shard()
        frame = peer.complete()  # stream sync on every rank + barrier
        if rank == 0:
            host.copy_(frame, non_blocking=True)
            torch.cuda.current_stream(g.local).synchronize()
            _ = int(host[height // 2, width // 2, 1])
        if s >= warmup:
            e2e.append((time.perf_counter() - t0) * 1e3)
    e2e_ms = float(np.mean(e2e))
    peer.close()
    png = c4_png_exchange(g, flat, cam, spheres, lights, width, height, max_depth, shard, solo, steps, warmup)
    if rank != 0:
        return {}
    px = width * height
    return {
        "workload": LABELS["c4"], "n_gpus": world, "ms_per_step": ms_per_step, "value": px / (ms_per_step * 1e-3) / 1e6, "unit": UNIT,
        "scaling": "strong", "exchange": {"kind": "peer stores over NVLink into rank 0's framebuffer (CUDA IPC), no gather",
                                          "layout": "RGB8, 32x1 patches, patches from both ends" if layout == "rgb8-wide" else "RGBA8, 8x4 patches",
                                          "strip_rows": shard.strip_rows},
        "per_rank_ms": per_rank, "own_n1_ms": solo_ms,
        "own_n1_note": "the same frame on rank 0 alone in the same run (RGBA8, 8x4 patches: its fastest 8-bit layout)",
        "efficiency_vs_own_n1": solo_ms / (world * ms_per_step),
        "frame_matches_single_gpu": matches,
        "e2e": {"ms_per_step": e2e_ms, "value": px / (e2e_ms * 1e-3) / 1e6,
                "what": "launch on every rank -> stream sync + barrier -> D2H of the assembled frame on rank 0 into pinned host memory -> one pixel read",
                "d2h_bytes_per_step": px * channels},
        "png_exchange": png,
        "steps": steps, "warmup": warmup,
    }


def c4_png_exchange(g: Gpu, flat, cam, spheres, lights, width, height, max_depth, shard, solo, steps: int, warmup: int):
    """The same frame with the exchange step compressed: every rank renders its strips into LOCAL memory, encodes
    their bands (rt_png_encode_bands_dev: each band of the image is one IDAT chunk) straight into rank 0's scratch
    area over NVLink -- a tenth of the bytes -- and rank 0 packs the file into pinned host memory."""
    torch, abi, lib, sharding = g.torch, g.abi, g.lib, g.sharding
    world, rank = g.world, g.rank
    r = int(lib.rt_png_rows_per_band(width, height))
    if r < 0 or shard.strip_rows % r != 0:
        return None
    scratch_bytes = int(lib.rt_png_scratch_bytes(width, height))
    cap = int(lib.rt_png_file_bound(width, height))
    env = g.env_of(flat, max_depth, abi.RT_FLAG_RGB8)
    local = torch.empty((shard.n_rows * width, 3), dtype=torch.uint8, device=g.dev)
    peer = sharding.PeerBuffer(scratch_bytes, rank, world, g.local, dst=0)
    file_host = torch.empty(cap, dtype=torch.uint8).pin_memory() if rank == 0 else None
    size_host = torch.zeros(1, dtype=torch.int64).pin_memory() if rank == 0 else None

    def launch(_step=0) -> None:
        abi.check(lib.rt_render_strips_dev(g.local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                           C.byref(cam), shard.strip_rows, rank, world, None, abi.RT_DTYPE_F32, 0,
                                           local.data_ptr(), None, g.sptr), "rt_render_strips_dev")
        abi.check(lib.rt_png_encode_bands_dev(g.local, local.data_ptr(), 3, width * 3, shard.n_rows, width, height,
                                              rank * shard.strip_rows, shard.strip_rows, world, peer.ptr, scratch_bytes, g.sptr),
                  "rt_png_encode_bands_dev")

    def pack() -> int:
        size_host[0] = -1
        abi.check(lib.rt_png_pack_dev(g.local, width, height, peer.ptr, file_host.data_ptr(), cap, size_host.data_ptr(), g.sptr), "rt_png_pack_dev")
        torch.cuda.current_stream(g.local).synchronize()
        return int(size_host[0])

    # the assembled file decodes to the single-GPU frame
    launch()
    peer.complete()
    matches, png_bytes = None, None
    if rank == 0:
        from python_ray_tracer_b200.infrastructure.png import read_png

        png_bytes = pack()
        with tempfile.TemporaryDirectory() as tmp:
            path = Path(tmp) / "c4.png"
            path.write_bytes(bytes(file_host.numpy()[:png_bytes]))
            decoded = read_png(path)
        matches = bool(np.array_equal(decoded, solo[:, :, :3].cpu().numpy()))
    ms = g.timed_steps(launch, steps, warmup, together=True)
    per_rank = g.gather_floats(float(ms.mean()))
    # the same two kernels for the WHOLE frame on rank 0 alone (the others wait): this path's own N = 1 time
    solo_ms = None
    if rank == 0:
        whole = local if world == 1 else torch.empty((height * width, 3), dtype=torch.uint8, device=g.dev)

        def launch_solo(_step=0) -> None:
            abi.check(lib.rt_render_strips_dev(g.local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                               C.byref(cam), shard.strip_rows, 0, 1, None, abi.RT_DTYPE_F32, 0,
                                               whole.data_ptr(), None, g.sptr), "rt_render_strips_dev")
            abi.check(lib.rt_png_encode_bands_dev(g.local, whole.data_ptr(), 3, width * 3, height, width, height, 0, height, 1,
                                                  peer.ptr, scratch_bytes, g.sptr), "rt_png_encode_bands_dev")

        was = g.world
        g.world = 1
        solo_ms = float(g.timed_steps(launch_solo, steps, warmup).mean())
        g.world = was
    g.barrier()
    e2e = []
    for s in range(warmup + steps):
        g.barrier()
        t0 = time.perf_counter()
        launch()
        peer.complete()
        if rank == 0:
            n = pack()
            _ = int(file_host[n - 1])
        if s >= warmup:
            e2e.append((time.perf_counter() - t0) * 1e3)
    peer.close()
    if rank != 0:
        return None
    px = width * height
    return {"ms_per_step": max(per_rank), "value": px / (max(per_rank) * 1e-3) / 1e6, "per_rank_ms": per_rank,
            "own_n1_ms": solo_ms, "efficiency_vs_own_n1": solo_ms / (world * max(per_rank)),
            "what": "per rank: trace kernel into local RGB8 + band encoder storing IDAT chunks into rank 0's scratch over NVLink",
            "rows_per_band": r, "png_bytes": png_bytes, "file_decodes_to_single_gpu_frame": matches,
            "e2e": {"ms_per_step": float(np.mean(e2e)), "value": px / (float(np.mean(e2e)) * 1e-3) / 1e6,
                    "what": "launches on every rank -> stream sync + barrier -> pack kernel on rank 0 writes the PNG into pinned host memory -> its last byte read",
                    "d2h_bytes_per_step": png_bytes}}


def c3_object(g: Gpu, steps: int, warmup: int) -> dict:
    """BASELINE configs[2] through the sphere tree on one GPU."""
    saved = (g.args.steps, g.args.warmup)
    g.args.steps, g.args.warmup = steps, warmup
    try:
        launch, counters, tables, cam, width, height, keep = device_frames(g, "c3", use_tree=True)
        ms = float(g.timed_steps(launch, steps, warmup).mean())
        k = counters()
    finally:
        g.args.steps, g.args.warmup = saved
    return {"workload": LABELS["c3"], "ms_per_step": ms, "value": width * height / (ms * 1e-3) / 1e6, "unit": UNIT,
            "accel": "sphere BVH (rt_accel_build_host), built before the timed region", "counters": k, "steps": steps, "warmup": warmup}


def run_b200_arm(args) -> None:
    g = Gpu(args)
    torch, abi, lib = g.torch, g.abi, g.lib
    name = args.workload
    if name == "c4":
        obj = c4_object(g, args.steps, args.warmup)
        if g.rank == 0:
            obj["config"] = config_of("c4", g.world)
            obj["metric"], obj["warmup"], obj["higher_is_better"], obj["dtype"], obj["data"] = METRIC, args.warmup, True, "f64", "synthetic"
            obj["gpu_launches"] = g.launches
            print(json.dumps(obj))
        if g.world > 1:
            g.dist.destroy_process_group()
        return

    label, flat0, width, height, max_depth = workload(name, g.rank, 0, g.world)
    use_tree = args.accel == "on" or (args.accel == "auto" and flat0.n_spheres >= abi.RT_ACCEL_AUTO_SPHERES)
    launch, counters, tables, cam, width, height, keep = device_frames(g, name, use_tree)
    n = width * height

    sampler = ClockSampler(g.local)
    sampler.start()
    wall0 = time.perf_counter()
    step_ms = g.timed_steps(launch, args.steps, args.warmup)
    wall = time.perf_counter() - wall0
    timed_mark = sampler.mark()
    # the ALU peaks of this GPU, measured in the same clock-sampled window (rank 0's denominators)
    fp64, fp32 = C.c_double(0.0), C.c_double(0.0)
    abi.check(lib.rt_measure_fma_peak(g.local, abi.RT_DTYPE_F64, C.byref(fp64)), "rt_measure_fma_peak f64")
    abi.check(lib.rt_measure_fma_peak(g.local, abi.RT_DTYPE_F32, C.byref(fp32)), "rt_measure_fma_peak f32")
    sampler.finish()
    clocks = sampler.summary(0, timed_mark)
    peak_clocks = sampler.summary(timed_mark)
    main_launches = g.launches

    ms_per_step = g.max_over_ranks(float(step_ms.mean()))
    pixels_per_step = n * g.world
    value = pixels_per_step / (ms_per_step * 1e-3) / 1e6
    k = counters()
    flop = k["sphere_tests"] * FLOP_PER_SPHERE_TEST + k["hits_shaded"] * FLOP_PER_HIT

    e2e = host_path(g, tables, cam, width, height, use_tree)
    c4 = c4_object(g, args.c4_steps, 3) if (args.c4 and name == "c2") else None
    extra = {}
    if g.rank == 0 and name == "c2":
        extra["pipeline"] = pipeline_object(g)
        if g.world == 1 and args.c3:
            extra["c3"] = c3_object(g, 20, 3)
    if g.world > 1:
        g.barrier()
    if g.rank != 0:
        if g.world > 1:
            g.dist.destroy_process_group()
        return

    # ---- rooflines of the trace kernel (rank 0's launch)
    kernel_ms = float(step_ms.mean())
    peaks = measured_peaks()
    achieved_gbs = BYTES_PER_PIXEL * n / (kernel_ms * 1e-3) / 1e9
    achieved_tflops = flop / (kernel_ms * 1e-3) / 1e12
    traffic, capture = ncu_traffic(name)
    cpu = None
    if g.world == 1 and not args.no_cpu_baseline:
        cpu = cpu_single_core(flat0, width, height, max_depth)
    small = (args.precision == "fast" and flat0.n_spheres <= abi.RT_SMALL_SCENE_SPHERES and not use_tree)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": g.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(name, g.world),
        "arm": {"precision": args.precision, "cutoff": args.cutoff,
                "arithmetic": "FP64 geometry (rays, roots, hit points, shadow decisions, GGX inputs), FP32 shading tail",
                "accel": ("sphere BVH (rt_accel_build_host): built before the timed region for `value`, inside it for `e2e`"
                          if use_tree else "none: every ray tests every sphere"),
                "kernel": "render_rows_small_kernel (sphere table as a kernel parameter)" if small else "render_rows_kernel",
                "output": "planar f32 linear RGB + RGBA8 (fused quantisation), 16 B per pixel",
                "parallelism": "frame-sharded x%d, no collective" % g.world,
                "l2": "512 MiB buffer written between timed steps (L2 flush); each step timed by its own CUDA event pair"},
        "e2e": e2e,
        "gpu_launches": main_launches,
        # The binding roofline of this path is the FP64 pipe (no contraction, 16 B written per
        # pixel): algorithmic flops / kernel time against the FP64 FMA rate measured on this GPU
        # in this run.  The HBM reading (MEASURED_PEAKS.json denominator) is kept beside it.
        "roofline": {"bound": "fp64-alu", "achieved": achieved_tflops, "peak": fp64.value / 1e12, "unit": "TFLOP/s",
                     "frac": achieved_tflops / (fp64.value / 1e12) if fp64.value else None,
                     "traffic": traffic, "traffic_capture": capture,
                     "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel under ncu --set full: "
                                     "the frame stays in the 126 MB L2 while the kernel runs, so DRAM sees almost nothing of the "
                                     "33 MB the launch writes",
                     "peak_source": "rt_measure_fma_peak (register-resident FP64 FMA chains), same GPU, same run, inside the clock sampler's window; "
                                    "MEASURED_PEAKS.json holds no ALU peak",
                     "peak_clocks": peak_clocks, "fp32_peak": fp32.value / 1e12, "kernel_ms": kernel_ms,
                     "algorithmic_flop_per_step": flop, "counters": k},
        "roofline_hbm": {"bound": "hbm", "achieved": achieved_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved_gbs / peaks["hbm_gbs"], "peak_source": peaks["source"],
                         "algorithmic_bytes_per_step": BYTES_PER_PIXEL * n},
        "clocks": clocks,
        "wall_s_timed_region": wall,
    }
    if c4:
        line["c4"] = c4
    line.update(extra)
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    if g.world > 1:
        g.dist.destroy_process_group()


def main() -> None:
    parser = argparse.ArgumentParser()
    parser.add_argument("--gpus", type=int, default=1)
    parser.add_argument("--steps", type=int, default=200)
    parser.add_argument("--warmup", type=int, default=5)
    parser.add_argument("--impl", choices=["b200", "reference"], default="b200")
    parser.add_argument("--workload", default="c2")
    parser.add_argument("--precision", choices=["fast", "exact"], default="fast")
    parser.add_argument("--cutoff", type=float, default=1e-7)
    parser.add_argument("--no-cpu-baseline", action="store_true")
    parser.add_argument("--accel", choices=["auto", "on", "off"], default="auto",
                        help="sphere BVH: auto = scenes of RT_ACCEL_AUTO_SPHERES shapes or more")
    parser.add_argument("--peer-layout", choices=["auto", "rgba8", "rgb8-wide"], default="auto",
                        help="c4: how the frame is stored into rank 0's buffer (auto: rgb8-wide from 8 ranks on)")
    parser.add_argument("--no-c4", dest="c4", action="store_false", help="skip the c4 object of the default run")
    parser.add_argument("--no-c3", dest="c3", action="store_false", help="skip the c3 object of the default N = 1 run")
    parser.add_argument("--c4-steps", type=int, default=30)
    args = parser.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
