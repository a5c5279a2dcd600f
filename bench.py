#!/usr/bin/env python
"""Benchmark of the hot path: ``NumpyRenderer.raytrace_scene`` replaced by the sm_100a trace kernel.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload c2|c3|c4|c5] [--precision fast|exact]

A step is one frame of the workload through the hot path (camera rays -> nearest hit ->
shadow ray -> shading -> bounce chain -> framebuffer stores).  The default workload is
BASELINE.json's ``configs[1]`` (README scene, 1920x1080, all effects); with N > 1 every rank
renders its own frame of the turntable (``configs[4]``: frame-sharded, no data-path
collective, weak scaling).  ``--workload c4`` is the 8K row-band configuration with an NCCL
gather of the bands to rank 0.  One JSON line is printed by rank 0.

Metric: Mrays/s = primary rays (= pixels) per second / 1e6; ``ms_per_step`` is ms/frame.
``value`` is device time with the scene resident in HBM (CUDA events on the launch stream,
L2 flushed between steps, max over ranks); ``e2e`` is the same frame through the host-buffer
C-ABI entry point (``rt_render_frame_host``): scene upload, kernel, linear RGB + 8-bit image
copied back into pinned host memory, every step.
"""

from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
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
# dram__bytes_read.sum + dram__bytes_write.sum per launch, one `ncu --set full` capture per workload
# (profiles/r01/*_details.txt); None where no capture was taken
NCU_DRAM_BYTES_PER_LAUNCH = {"c2": 639488}


def workload(name: str, rank: int, step: int, world: int):
    """(label, flat scene, width, height, max_depth) of one step on one rank."""
    if name == "c2":
        # N > 1: a batch of N frames of configs[1] per step, one per rank (frame-sharded, nothing
        # exchanged): per-GPU work is exactly the N = 1 work, which is what weak scaling compares.
        # (Turntable frames differ in cost by +-20 %; that batch is workload c5.)
        label = "README scene 1920x1080, all effects (configs[1])"
        if world > 1:
            label += "; batch of %d such frames per step, one per rank" % world
        return label, flatten_scene(scenes.readme_scene(1920, 1080), README_DEFAULT), 1920, 1080, 0
    if name == "c5":
        frame = (step * world + rank) % 256
        return ("256-frame turntable of the README scene at 1080p, frame-sharded (configs[4])",
                flatten_scene(scenes.turntable_scene(frame, 256, 1920, 1080), README_DEFAULT), 1920, 1080, 0)
    if name == "c3":
        return ("1024 random spheres + checkered ground, 4 PointLights (reference semantics: light 0 shades), depth 8, 3840x2160 (configs[2])",
                flatten_scene(scenes.random_spheres_scene(3840, 2160), README_DEFAULT), 3840, 2160, 8)
    if name == "c4":
        return ("README scene 7680x4320, row-band sharded with gather to rank 0 (configs[3])",
                flatten_scene(scenes.readme_scene(7680, 4320), README_DEFAULT), 7680, 4320, 0)
    if name == "c1":
        return "README scene 960x540 (configs[0])", flatten_scene(scenes.readme_scene(960, 540), README_DEFAULT), 960, 540, 0
    raise SystemExit(f"unknown workload {name}")


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int, period: float = 0.002) -> None:
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

    def finish(self) -> dict:
        self._stop_event.set()
        self.join(timeout=2.0)
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peaks() -> dict:
    path = ROOT / "MEASURED_PEAKS.json"
    if path.exists():
        data = json.loads(path.read_text())
        return {"hbm_gbs": float(data["hbm_gbs"]), "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


# --------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/ code: as the timed baseline)
# --------------------------------------------------------------------------------------
def _rows_job(args):
    flat, width, height, rows, max_depth = args
    from oracle import numpy_port

    pixels = (rows[:, None] * width + np.arange(width)[None, :]).reshape(-1)
    rgb, counters = numpy_port.render(flat, width, height, pixels=pixels, max_depth=max_depth)
    return float(rgb.sum()), counters[0]


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
    bands (bit-identical to the full frame) and rendered by a process pool."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp

    label, flat, width, height, max_depth = workload(args.workload, 0, 0, max(1, args.gpus))
    cores = host_cores()
    # bounded sample: evenly spaced rows sized for ~2 s per step per core at ~0.4 Mpx/s/core
    target_px = int(min(width * height, 0.4e6 * 2.0 * cores / max(1.0, flat.n_spheres / 3.0) ** 2))
    rows = sample_rows(height, max(cores, min(height, target_px // width)))
    jobs = [(flat, width, height, rows[j::cores], max_depth) for j in range(cores) if len(rows[j::cores])]
    n = width * len(rows)
    times = []
    with mp.get_context("fork").Pool(cores) as pool:
        for step in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pool.map(_rows_job, jobs, chunksize=1)
            dt = time.perf_counter() - t0
            if step >= args.warmup:
                times.append(dt)
    sec = float(np.mean(times))
    value = n / sec / 1e6
    sample = (f"{len(rows)} evenly spaced rows of {width}x{height} ({n} pixels) per step, interleaved over {len(jobs)} processes; "
              "oracle/numpy_port.py = the reference's whole-array NumPy algorithm on the flattened scene")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "ms_per_frame_extrapolated": width * height / (value * 1e6) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": label, "width": width, "height": height, "spheres": flat.n_spheres},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
def run_b200_arm(args) -> None:
    import torch
    import torch.distributed as dist

    from python_ray_tracer_b200.infrastructure.cuda import abi, affinity, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bound = affinity.bind_to_device(local) if world > 1 else None  # pinned buffers next to this rank's GPU
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = abi.load()
    stream = torch.cuda.current_stream(local)
    sptr = int(stream.cuda_stream)
    mode = abi.RT_MODE_FAST if args.precision == "fast" else abi.RT_MODE_EXACT
    banded = args.workload == "c4"  # one frame over all ranks (N = 1 included: same output format)

    total = args.warmup + args.steps
    frames = [workload(args.workload, rank, s, world) for s in range(total)]
    label, _, width, height, max_depth = frames[0]
    # one frame over all ranks: interleaved strips (sharding.StripShard) for the device path,
    # contiguous bands for the host-buffer path
    shard = sharding.StripShard(height, world, rank) if banded else None
    row0, row1 = sharding.band_rows(height, world, rank) if banded else (0, height)
    n = shard.n_rows * width if banded else (row1 - row0) * width
    cam = abi.make_camera(frames[0][1].eye, width, height)

    # scene tables of every step resident in HBM before the timed region
    tables = []
    for _, flat, _, _, _ in frames:
        env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, max_depth=max_depth,
                           mode=mode, cutoff=args.cutoff if mode == abi.RT_MODE_FAST else 0.0,
                           spheres_host=flat.spheres, point_lights_host=flat.point_lights)
        tables.append((torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev), env, flat))
    # sphere tree (rt_accel_build_host) for scenes past the measured break-even; built and
    # uploaded before the timed region for the device path (scene-resident data), rebuilt
    # inside the timed region by the host entry point of the e2e path
    use_tree = args.accel == "on" or (args.accel == "auto" and frames[0][1].n_spheres >= abi.RT_ACCEL_AUTO_SPHERES)
    blobs = {}
    if use_tree:
        for _, _, env, flat in tables:
            key = flat.key()
            if key not in blobs:
                capacity = int(lib.rt_accel_bytes(flat.n_spheres))
                blob = np.empty(capacity, dtype=np.uint8)
                used = C.c_int64(0)
                abi.check(lib.rt_accel_build_host(flat.spheres.ctypes.data, flat.n_spheres, C.byref(env), blob.ctypes.data,
                                                  capacity, C.byref(used)), "rt_accel_build_host")
                blobs[key] = torch.from_numpy(blob[: used.value]).to(dev)
            env.accel_dev, env.accel_bytes = blobs[key].data_ptr(), blobs[key].numel()
    planes = torch.empty((3, n), dtype=torch.float32, device=dev)
    rgba8 = torch.empty((n, 4), dtype=torch.uint8, device=dev)
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    peer = None
    peer_layout = None
    if banded and args.exchange == "peer":
        # rank 0 owns the whole-frame 8-bit buffer; the other ranks map it and their kernels
        # store into it over NVLink while they render (no gather, no permutation).  Up to four
        # writers rank 0's port keeps up with 32-byte row segments of 8x4 patches; from eight on
        # the frame goes out as 96-byte rows of 32x1 patches of 3-byte pixels, the patches taken
        # from both ends of the shard (profiles/r01/c4_store_layouts.txt).
        peer_layout = args.peer_layout if args.peer_layout != "auto" else ("rgb8-wide" if world >= 8 else "rgba8")
        env_flags = abi.RT_FLAG_IMAGE_ORDER
        if peer_layout == "rgb8-wide":
            env_flags |= abi.RT_FLAG_WIDE_STORES | abi.RT_FLAG_RGB8 | abi.RT_FLAG_EVEN_STORES
        peer = sharding.PeerFrame(height, width, rank, world, local, dst=0, channels=3 if peer_layout == "rgb8-wide" else 4)
        for i, (sph, lig, env, flat_) in enumerate(tables):
            env.flags |= env_flags

    def launch(step: int, with_stats: bool = False) -> None:
        spheres, lights, env, flat = tables[step]
        if peer is not None:
            # the exchange step is inside the kernel: stores go to rank 0's framebuffer
            abi.check(lib.rt_render_strips_dev(local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                               C.byref(cam), shard.strip_rows, rank, world, None, abi.RT_DTYPE_F32, 0,
                                               peer.ptr, stats.data_ptr() if with_stats else None, sptr),
                      "rt_render_strips_dev")
        elif banded:
            abi.check(lib.rt_render_strips_dev(local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                               C.byref(cam), shard.strip_rows, rank, world, planes.data_ptr(),
                                               abi.RT_DTYPE_F32, n, rgba8.data_ptr(),
                                               stats.data_ptr() if with_stats else None, sptr), "rt_render_strips_dev")
            # the one exchange step of the path: 8-bit shards -> rank 0, permuted to image order
            sharding.gather_frame(rgba8.view(shard.n_rows, width, 4), shard, dst=0)
        else:
            abi.check(lib.rt_render_rows_dev(local, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                                             C.byref(cam), row0, row1, planes.data_ptr(), abi.RT_DTYPE_F32, n,
                                             rgba8.data_ptr(), stats.data_ptr() if with_stats else None, sptr),
                      "rt_render_rows_dev")

    def barrier() -> None:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(local)

    for s in range(args.warmup):
        launch(s)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    events = []
    wall0 = time.perf_counter()
    for s in range(args.warmup, total):
        flush.fill_(s & 0xFF)  # evict the previous frame's lines from L2 (not timed)
        if banded and world > 1:
            # one frame over all ranks: start it together, after every rank's flush (a rank
            # still flushing would compete with the others' peer stores into rank 0)
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        launch(s)
        e1.record(stream)
        events.append((e0, e1))
    sampler.sample()
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.finish()
    step_ms = np.array([a.elapsed_time(b) for a, b in events])
    mean_ms = torch.tensor([float(step_ms.mean())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(mean_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(mean_ms.item())
    pixels_per_step = width * height if banded else width * height * world
    value = pixels_per_step / (ms_per_step * 1e-3) / 1e6

    # ---- work counters of one step (untimed) for the arithmetic roofline
    stats.zero_()
    launch(total - 1, with_stats=True)
    torch.cuda.synchronize(local)
    traced, hits, deepest, tests = (int(v) for v in stats.cpu().tolist())
    flop = tests * FLOP_PER_SPHERE_TEST + hits * FLOP_PER_HIT

    # ---- end to end through the host-buffer C-ABI entry point, pinned host outputs.
    # The step's result is what raytrace_scene returns: the linear RGB planes.  A second
    # pass also brings back the 8-bit image save_image would write (reported beside it).
    if peer is not None:
        torch.cuda.synchronize(local)
        peer.close()
        peer = None
        for _, _, env, _ in tables:
            env.flags &= ~(abi.RT_FLAG_IMAGE_ORDER | abi.RT_FLAG_WIDE_STORES | abi.RT_FLAG_RGB8 | abi.RT_FLAG_EVEN_STORES)
    for _, _, env, _ in tables:  # the host entry point builds its own tree (or none: --accel off)
        env.accel_dev, env.accel_bytes = None, 0
        if not use_tree:
            env.flags |= abi.RT_FLAG_NO_ACCEL
    n_host = (row1 - row0) * width  # host path: this rank's contiguous band
    rgb_host = torch.empty((3, n_host), dtype=torch.float32).pin_memory()
    rgb8_host = torch.empty((n_host, 3), dtype=torch.uint8).pin_memory()

    def e2e_pass(with_rgb8: bool, with_planes: bool = True) -> float:
        times = []
        for s in range(total):
            _, _, env, flat = tables[s]
            t0 = time.perf_counter()
            abi.check(lib.rt_render_frame_host(local, flat.spheres.ctypes.data, flat.n_spheres, flat.point_lights.ctypes.data,
                                               C.byref(env), C.byref(cam), row0, row1,
                                               rgb_host.data_ptr() if with_planes else None, abi.RT_DTYPE_F32,
                                               rgb8_host.data_ptr() if with_rgb8 else None, None), "rt_render_frame_host")
            # the result is read on the host inside the timed region
            _ = (float(rgb_host[0, n_host // 2]) + float(rgb_host[2, n_host - 1])) if with_planes else int(rgb8_host[n_host // 2, 1])
            if s >= args.warmup:
                times.append(time.perf_counter() - t0)
        ms = torch.tensor([float(np.mean(times)) * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    e2e_ms_value = e2e_pass(False)
    e2e_rgb8_ms = e2e_pass(True)
    e2e_rgb8_only_ms = e2e_pass(True, with_planes=False)
    e2e_value = pixels_per_step / (e2e_ms_value * 1e-3) / 1e6
    flat = tables[-1][3]
    h2d = flat.spheres.nbytes + flat.point_lights.nbytes
    d2h = rgb_host.numel() * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines of the trace kernel (rank 0's launch)
    kernel_ms = float(step_ms.mean())
    peaks = measured_peaks()
    bytes_per_pixel = (3.0 if peer_layout == "rgb8-wide" else 4.0) if (banded and args.exchange == "peer") else BYTES_PER_PIXEL
    achieved_gbs = bytes_per_pixel * n / (kernel_ms * 1e-3) / 1e9
    fp64 = C.c_double(0.0)
    fp32 = C.c_double(0.0)
    abi.check(lib.rt_measure_fma_peak(local, abi.RT_DTYPE_F64, C.byref(fp64)), "rt_measure_fma_peak f64")
    abi.check(lib.rt_measure_fma_peak(local, abi.RT_DTYPE_F32, C.byref(fp32)), "rt_measure_fma_peak f32")
    achieved_tflops = flop / (kernel_ms * 1e-3) / 1e12

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_single_core(frames[0][1], width, height, max_depth)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if banded else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": label, "width": width, "height": height, "spheres": frames[0][1].n_spheres,
                   "max_depth": max_depth or None, "precision": args.precision, "cutoff": args.cutoff,
                   "accel": ("sphere BVH (rt_accel_build_host): built before the timed region for `value`, inside it for `e2e`"
                             if use_tree else "none: every ray tests every sphere"),
                   "output": (("whole-frame %s on rank 0 (fused quantisation), no linear planes" % ("RGB8, 32x1 patches" if peer_layout == "rgb8-wide" else "RGBA8, 8x4 patches")) if banded and args.exchange == "peer" else "planar f32 linear RGB + RGBA8 (fused quantisation)"),
                   "parallelism": (("interleaved 16-row strips x%d, kernels store the 8-bit frame into rank 0's framebuffer over NVLink peer memory (no gather)" % world)
                                   if args.exchange == "peer" else ("interleaved 16-row strips x%d + NCCL gather of RGBA8 to rank 0 + row permutation" % world)) if banded else ("frame-sharded x%d, no collective" % world),
                   "l2": "512 MiB buffer written between timed steps (L2 flush); each step timed by its own CUDA event pair"},
        "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms_value, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "d2h_gb_per_s_per_gpu": d2h / (e2e_ms_value * 1e-3) / 1e9,
                "host_cpus_bound": len(bound) if bound else None,
                "api": "rt_render_frame_host: scene upload + kernel + D2H of the f32 linear planes into pinned host memory, row chunks pipelined (kernel of chunk i+1 under the copy of chunk i, small first chunk, one strided copy per chunk)",
                "with_rgb8": {"ms_per_step": e2e_rgb8_ms, "value": pixels_per_step / (e2e_rgb8_ms * 1e-3) / 1e6,
                              "d2h_bytes_per_step": d2h + rgb8_host.numel()},
                # what render_image_pipeline needs on the host: only the 8-bit image save_image writes
                "rgb8_only": {"ms_per_step": e2e_rgb8_only_ms, "value": pixels_per_step / (e2e_rgb8_only_ms * 1e-3) / 1e6,
                              "d2h_bytes_per_step": rgb8_host.numel()}},
        "gpu_launches": args.steps,
        # The binding roofline of this path is the FP64 pipe (no contraction, 16 B written per
        # pixel): algorithmic flops / kernel time against the FP64 FMA rate measured on this GPU
        # in this run.  The HBM reading (MEASURED_PEAKS.json denominator) is kept beside it.
        "roofline": {"bound": "fp64-alu", "achieved": achieved_tflops, "peak": fp64.value / 1e12, "unit": "TFLOP/s",
                     "frac": achieved_tflops / (fp64.value / 1e12) if fp64.value else None,
                     "traffic": NCU_DRAM_BYTES_PER_LAUNCH.get(args.workload),
                     "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of one launch under ncu --set full "
                                     "(profiles/): the frame stays in the 126 MB L2 while the kernel runs, so DRAM sees "
                                     "almost nothing of the 33 MB the launch writes",
                     "peak_source": "rt_measure_fma_peak (register-resident FP64 FMA chains), same GPU, same run; "
                                    "MEASURED_PEAKS.json holds no ALU peak",
                     "fp32_peak": fp32.value / 1e12, "kernel": "render_rows_kernel", "kernel_ms": kernel_ms,
                     "algorithmic_flop_per_step": flop,
                     "counters": {"rays_traced": traced, "hits_shaded": hits, "generations": deepest, "sphere_tests": tests}},
        "roofline_hbm": {"bound": "hbm", "achieved": achieved_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved_gbs / peaks["hbm_gbs"], "peak_source": peaks["source"],
                         "algorithmic_bytes_per_step": bytes_per_pixel * n},
        "roofline_alu": {"frac": achieved_tflops / (fp64.value / 1e12) if fp64.value else None,
                         "note": "same as roofline (kept for the sweep scripts)"},
        "clocks": clocks,
        "wall_s_timed_region": wall,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


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
                        help="c4 with --exchange peer: how the frame is stored into rank 0's buffer (auto: rgb8-wide from 8 ranks on)")
    parser.add_argument("--exchange", choices=["peer", "nccl"], default="peer",
                        help="c4 with N > 1: kernels store into rank 0's framebuffer over NVLink (peer) or "
                             "render locally and gather with NCCL (nccl)")
    args = parser.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
