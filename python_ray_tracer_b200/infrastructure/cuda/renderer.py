"""``CudaRenderer``: the B200 backend of the reference's ``Renderer`` port.

Drop-in for ``NumpyRenderer`` (reference ``ray_tracer/infrastructure/numpy/base.py:90-151``)
under ``render_image_pipeline(scene, path, renderer)`` (``application.py:43-52``).  The
three port methods keep their names and argument meaning:

* ``get_ray_directions(camera)``  (``base.py:123-141``) returns a ``CameraRays`` handle;
  the directions are computed inside the trace kernel from (row, col) and only
  materialised if somebody reads ``.x/.y/.z``;
* ``raytrace_scene(origin, directions, scene)``  (``base.py:91-121``) flattens the scene
  into the sphere SoA of ``include/rt_abi.h`` and launches the trace kernel through the
  C ABI; explicit NumPy origin/direction arrays are accepted too;
* ``save_image(color, camera, path)``  (``base.py:143-151``) quantises on the device
  (``uint8(255*clip(v,0,1))``, truncating) and writes the PNG on the host.

torch is used for device buffers and streams only.  There is no CPU fallback: if
``librt_b200.so`` is missing or no CUDA device is visible the calls raise.
"""

from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Any, Dict, Optional, Tuple

import numpy as np

from ...application import Renderer
from ...domain import Camera, Scene3D
from ..flatten import DefaultMaterial, FlatScene, flatten_scene
from ..png import write_png
from ..values import NumpyRGBColor, NumpyVector3D, as_float_triple
from . import abi

_PRECISIONS = {"exact": abi.RT_MODE_EXACT, "fast": abi.RT_MODE_FAST}


def _torch():
    import torch  # deferred: describing scenes must not need torch

    return torch


def _require_cuda(device: int):
    torch = _torch()
    if not torch.cuda.is_available():
        raise RuntimeError("CudaRenderer needs a CUDA device; there is no CPU fallback")
    if device >= torch.cuda.device_count():
        raise RuntimeError(f"cuda:{device} does not exist ({torch.cuda.device_count()} visible)")
    return torch


class CameraRays(NumpyVector3D):
    """Handle returned by ``get_ray_directions``: rows ``[row0, row1)`` of a camera.

    Reading ``x``, ``y`` or ``z`` materialises the unit directions (float64, flat,
    row-major) through ``rt_ray_directions_dev``.
    """

    __slots__ = ("camera", "row0", "row1", "_renderer", "_planes")

    def __init__(self, renderer: "CudaRenderer", camera: Camera, row0: int, row1: int) -> None:
        self.camera = camera
        self.row0 = int(row0)
        self.row1 = int(row1)
        self._renderer = renderer
        self._planes = None

    def _materialise(self) -> np.ndarray:
        if self._planes is None:
            self._planes = self._renderer._ray_directions(self.camera, self.row0, self.row1)
        return self._planes

    x = property(lambda self: self._materialise()[0])
    y = property(lambda self: self._materialise()[1])
    z = property(lambda self: self._materialise()[2])

    def __len__(self) -> int:
        return (self.row1 - self.row0) * int(self.camera.width)


class DeviceColor(NumpyRGBColor):
    """Colour result living on the GPU: planar linear RGB ``[3, n]`` and/or the 8-bit image.

    ``x``/``y``/``z`` are NumPy views of one device-to-host copy, made on first access,
    so the object reads like the reference's ``NumpyRGBColor`` of flat arrays
    (``base.py:119-121``).  When ``raytrace_scene`` was handed the renderer's own
    ``CameraRays`` it renders only the 8-bit image ``save_image`` needs
    (``rgba8``, here ``[n, 3]``) and keeps a recipe for the linear planes: they are rendered
    if and when somebody reads them (``render_image_pipeline`` never does,
    ``application.py:43-52``).
    """

    __slots__ = ("_planes", "rgba8", "stats", "device", "_host", "_render_planes")

    def __init__(self, planes, rgba8=None, stats=None, device: int = 0, render_planes=None) -> None:
        self._planes = planes  # torch tensor [3, n] float32/float64 on the device, or None (lazy)
        self.rgba8 = rgba8  # torch tensor [n, 4] or [n, 3] uint8 on the device, or None
        self.stats = stats  # torch tensor [4] int64 on the device, or None
        self.device = device
        self._host = None
        self._render_planes = render_planes

    @property
    def planes(self):
        if self._planes is None:
            if self._render_planes is None:
                raise RuntimeError("this colour holds no linear planes")
            self._planes = self._render_planes()
        return self._planes

    def numpy(self) -> np.ndarray:
        if self._host is None:
            self._host = self.planes.cpu().numpy()
            # the copy synchronised with the launch: did its kernel report anything? (rt_device_status)
            abi.check(abi.load().rt_device_status(self.device), "rt_device_status")
        return self._host

    x = property(lambda self: self.numpy()[0])
    y = property(lambda self: self.numpy()[1])
    z = property(lambda self: self.numpy()[2])

    def counters(self) -> Optional[Dict[str, int]]:
        """Work counters of the launch (``collect_stats=True``)."""
        if self.stats is None:
            return None
        traced, hits, deepest, tests = (int(v) for v in self.stats.cpu().tolist())
        return {"rays_traced": traced, "hits_shaded": hits, "generations": deepest, "sphere_tests": tests}


class _Nvtx:
    """NVTX range around one stage (upload, kernel, download, encode): shows up in Nsight timelines."""

    def __init__(self, name: str) -> None:
        self.name = name

    def __enter__(self):
        try:
            _torch().cuda.nvtx.range_push(self.name)
            self.on = True
        except Exception:  # no CUDA runtime: describing scenes must still work
            self.on = False
        return self

    def __exit__(self, *exc):
        if self.on:
            _torch().cuda.nvtx.range_pop()
        return False


class CudaRenderer(Renderer):
    """``Renderer`` port on one B200 (``cuda:device``).

    The five gains are the README constructor's (``README.md:50-56``); they are the
    material of every shape that carries a colour instead of a ``NumpyShader``
    (SURVEY.md section 8b).  Keyword extras:

    ``precision``  ``"exact"``: all-FP64 in the reference's operation order, no FMA
                   contraction.  ``"fast"`` (default): FP64 with restructured arithmetic.
    ``max_depth``  generations of rays traced (primary = generation 0); ``None`` = no cap,
                   like the reference (``shader.py:143-161``).
    ``cutoff``     fast mode only: stop a bounce chain once what it can still add is
                   provably below this many linear units (0 = never).
    ``all_point_lights``  extension (``main.py:45`` TODO): shade with every point light.
    ``output_dtype``  ``"float64"`` (default, the reference's dtype) or ``"float32"``.
    ``collect_stats``  count rays / hits / generations / sphere tests per launch.
    ``png_compress_level``  ``None`` (default) = the library's own deflate (``rt_png_deflate_rows``, row
                   bands on ``png_threads`` threads); an integer = that zlib level (6 = Pillow's
                   default, byte-for-byte the kind of file the reference writes, ~30x slower).
                   Decoded pixels do not depend on it.
    ``png_threads``  threads ``save_image`` deflates with (default: up to 8 of the host's cores).
    ``png_encoder``  ``"device"`` (default): the file is made on the GPU (``rt_png_encode_bands_dev`` + ``rt_png_pack_dev``) and only it
                   crosses PCIe; ``"host"``: the 8-bit image is copied to the host and deflated there
                   (``png_compress_level`` / ``png_threads``).  Decoded pixels are the same.
    ``accel``      ``"auto"`` (default): scenes of at least ``ACCEL_MIN_SPHERES`` shapes get a
                   bounding-volume tree (``rt_accel_build_host``) so rays skip sphere tests
                   that provably miss; ``True`` / ``False`` force it on or off.  Results do
                   not depend on it.
    """

    ACCEL_MIN_SPHERES = abi.RT_ACCEL_AUTO_SPHERES  # measured break-even (profiles/r01/accel_sweep.txt)

    def __init__(
        self,
        reflection_gain: float = 0.8,
        specular_gain: float = 1.0,
        specular_roughness: float = 0.5,
        iridescence_gain: float = 0.05,
        diffuse_gain: float = 1.0,
        *,
        precision: str = "fast",
        max_depth: Optional[int] = None,
        cutoff: float = 1e-7,
        all_point_lights: bool = False,
        output_dtype: str = "float64",
        collect_stats: bool = False,
        device: int = 0,
        png_compress_level: Optional[int] = None,
        png_threads: Optional[int] = None,
        png_encoder: str = "device",
        accel: Any = "auto",
    ) -> None:
        if precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}, got {precision!r}")
        if output_dtype not in ("float32", "float64"):
            raise ValueError("output_dtype must be 'float32' or 'float64'")
        if max_depth is not None and int(max_depth) < 1:
            raise ValueError("max_depth must be >= 1 (generations traced) or None")
        if not cutoff >= 0.0:
            raise ValueError("cutoff must be >= 0")
        self.default_material = DefaultMaterial(
            reflection_gain=reflection_gain,
            specular_gain=specular_gain,
            specular_roughness=specular_roughness,
            iridescence_gain=iridescence_gain,
            diffuse_gain=diffuse_gain,
        )
        self.precision = precision
        self.max_depth = None if max_depth is None else int(max_depth)
        self.cutoff = float(cutoff)
        self.all_point_lights = bool(all_point_lights)
        self.output_dtype = output_dtype
        self.collect_stats = bool(collect_stats)
        self.device = int(device)
        self.png_compress_level = None if png_compress_level is None else int(png_compress_level)
        if png_threads is None:
            import os

            try:
                png_threads = len(os.sched_getaffinity(0))
            except AttributeError:
                png_threads = os.cpu_count() or 1
            png_threads = min(8, png_threads)
        self.png_threads = max(1, int(png_threads))
        if png_encoder not in ("device", "host"):
            raise ValueError("png_encoder must be 'device' or 'host'")
        # a zlib level asks for a zlib file: that one is made on the host
        self.png_encoder = "host" if self.png_compress_level is not None else png_encoder
        self._png_device = None
        self._png_pool = None
        self._rgb8_dev = None     # reused across frames: the 8-bit image on the device ...
        self._rgb8_pinned = None  # ... and its pinned landing zone on the host
        self.last_timing: Dict[str, float] = {}
        if accel not in ("auto", True, False):
            raise ValueError("accel must be 'auto', True or False")
        self.accel = accel
        self._accel_dev = None
        self._scene_key: Optional[bytes] = None
        self._scene_dev: Optional[Tuple[Any, Any]] = None
        self._unused_dev = None  # what camera launches of kernel-parameter scenes are handed as table pointers

    # ------------------------------------------------------------------ helpers
    def _wants_accel(self, flat: FlatScene) -> bool:
        return self.accel is True or (self.accel == "auto" and flat.n_spheres >= self.ACCEL_MIN_SPHERES)

    def _env(self, flat: FlatScene, extra_flags: int = 0) -> abi.RtEnv:
        env = self._plain_env(flat, extra_flags)
        if self._accel_dev is not None and self._scene_key == flat.key():
            env.accel_dev = self._accel_dev.data_ptr()
            env.accel_bytes = self._accel_dev.numel()
        return env

    def _plain_env(self, flat: FlatScene, extra_flags: int = 0) -> abi.RtEnv:
        return abi.make_env(
            flat.eye,
            len(flat.point_lights),
            flat.dome_intensity,
            flat.dome_color,
            max_depth=self.max_depth or 0,
            mode=_PRECISIONS[self.precision],
            flags=(abi.RT_FLAG_ALL_POINT_LIGHTS if self.all_point_lights else abi.RT_FLAG_NONE) | extra_flags,
            cutoff=self.cutoff if self.precision == "fast" else 0.0,
            spheres_host=flat.spheres,
            point_lights_host=flat.point_lights,
        )

    def flatten(self, scene: Scene3D) -> FlatScene:
        return flatten_scene(scene, self.default_material)

    def upload_scene(self, flat: FlatScene):
        """Device copies of the sphere SoA and the point-light table (cached by content)."""
        torch = _require_cuda(self.device)
        key = flat.key()
        if key == self._scene_key and self._scene_dev is not None:
            return self._scene_dev
        with _Nvtx("rt: scene upload"):
            dev = torch.device("cuda", self.device)
            spheres = torch.from_numpy(flat.spheres).to(dev, non_blocking=False)
            lights = torch.from_numpy(flat.point_lights).to(dev, non_blocking=False)
            self._accel_dev = None
            if self._wants_accel(flat):
                lib = abi.load()
                capacity = int(lib.rt_accel_bytes(flat.n_spheres))
                blob = np.empty(capacity, dtype=np.uint8)
                used = C.c_int64(0)
                env = self._plain_env(flat)
                abi.check(lib.rt_accel_build_host(flat.spheres.ctypes.data, flat.n_spheres, C.byref(env),
                                                  blob.ctypes.data, capacity, C.byref(used)), "rt_accel_build_host")
                self._accel_dev = torch.from_numpy(blob[: used.value]).to(dev, non_blocking=False)
            self._scene_key, self._scene_dev = key, (spheres, lights)
        return self._scene_dev

    def _scene_pointers(self, flat: FlatScene, env: abi.RtEnv):
        """Device pointers of the sphere SoA and the light table for a CAMERA launch with ``env``.  A scene the
        library sends as a kernel parameter (``rt_scene_travels_as_parameter``: the README and ``main.py`` scenes)
        is never read from device memory, so nothing is uploaded for it -- and a scene that changes every frame
        (an animation) costs no copy and no wait on the stream."""
        if abi.load().rt_scene_travels_as_parameter(C.byref(env), flat.n_spheres):
            if self._unused_dev is None:
                torch = _require_cuda(self.device)
                self._unused_dev = torch.zeros(256, dtype=torch.uint8, device=torch.device("cuda", self.device))
            return self._unused_dev.data_ptr(), self._unused_dev.data_ptr()
        spheres, lights = self.upload_scene(flat)
        return spheres.data_ptr(), lights.data_ptr()

    def _stream(self) -> int:
        return int(_torch().cuda.current_stream(self.device).cuda_stream)

    def _torch_dtype(self):
        torch = _torch()
        return torch.float64 if self.output_dtype == "float64" else torch.float32

    def _abi_dtype(self) -> int:
        return abi.RT_DTYPE_F64 if self.output_dtype == "float64" else abi.RT_DTYPE_F32

    def _new_stats(self):
        if not self.collect_stats:
            return None
        torch = _torch()
        return torch.zeros(abi.RT_STATS_WORDS, dtype=torch.int64, device=torch.device("cuda", self.device))

    def _ray_directions(self, camera: Camera, row0: int, row1: int) -> np.ndarray:
        torch = _require_cuda(self.device)
        lib = abi.load()
        n = (row1 - row0) * int(camera.width)
        out = torch.empty((3, n), dtype=torch.float64, device=torch.device("cuda", self.device))
        cam = abi.make_camera(as_float_triple(camera.position, "camera position"), camera.width, camera.height)
        abi.check(
            lib.rt_ray_directions_dev(self.device, C.byref(cam), row0, row1, out.data_ptr(), n, self._stream()),
            "rt_ray_directions_dev",
        )
        return out.cpu().numpy()

    # ------------------------------------------------------------------ the port
    def get_ray_directions(self, camera: Camera, rows: Optional[Tuple[int, int]] = None) -> CameraRays:
        """Primary rays of ``camera`` (``base.py:123-141``); ``rows=(row0, row1)`` selects
        a row band (the multi-GPU sharding unit)."""
        width, height = int(camera.width), int(camera.height)
        if width <= 0 or height <= 0:
            raise ValueError(f"camera size {width}x{height}")
        row0, row1 = (0, height) if rows is None else (int(rows[0]), int(rows[1]))
        if not 0 <= row0 <= row1 <= height:
            raise ValueError(f"rows {rows} outside 0..{height}")
        return CameraRays(self, camera, row0, row1)

    def render_rows(self, flat: FlatScene, camera: Camera, row0: int, row1: int, *, planes=None,
                    rgba8=None, want_rgba8: bool = False, flags: int = 0, no_planes: bool = False) -> DeviceColor:
        """Launch the camera-ray kernel for rows ``[row0, row1)`` into (optionally
        caller-supplied) device buffers.  ``flags``: launch options of ``include/rt_abi.h``
        (``RT_FLAG_WIDE_STORES``: 32x1 patches; ``RT_FLAG_RGB8``: the 8-bit image has 3 bytes per
        pixel) -- they change how the frame is stored, never its values.  Asynchronous on
        torch's current stream."""
        torch = _require_cuda(self.device)
        lib = abi.load()
        dev = torch.device("cuda", self.device)
        n = (row1 - row0) * int(camera.width)
        if planes is None and not no_planes:
            planes = torch.empty((3, n), dtype=self._torch_dtype(), device=dev)
        if rgba8 is None and want_rgba8:
            rgba8 = torch.empty((n, 3 if flags & abi.RT_FLAG_RGB8 else 4), dtype=torch.uint8, device=dev)
        stats = self._new_stats()
        if self._wants_accel(flat):
            self.upload_scene(flat)  # (builds and uploads the tree that _env names)
        env = self._env(flat, flags)
        spheres_ptr, lights_ptr = self._scene_pointers(flat, env)
        cam = abi.make_camera(as_float_triple(camera.position, "camera position"), camera.width, camera.height)
        abi.check(
            lib.rt_render_rows_dev(
                self.device, spheres_ptr, flat.n_spheres, lights_ptr, C.byref(env), C.byref(cam),
                row0, row1, None if planes is None else planes.data_ptr(), self._abi_dtype(),
                max(n, 1) if planes is None else planes.stride(0),
                None if rgba8 is None else rgba8.data_ptr(), None if stats is None else stats.data_ptr(),
                self._stream(),
            ),
            "rt_render_rows_dev",
        )
        return DeviceColor(planes, rgba8, stats, self.device)

    def render_strips(self, flat: FlatScene, camera: Camera, shard: Any, *, planes=None, rgba8=None,
                      want_rgba8: bool = False, flags: int = 0) -> DeviceColor:
        """Launch the camera-ray kernel for an interleaved shard of the frame
        (``sharding.StripShard``: strips ``rank, rank + world, ...`` of ``strip_rows`` rows)
        through ``rt_render_strips_dev``.  The result holds the shard's rows packed in strip
        order.  Asynchronous on torch's current stream."""
        torch = _require_cuda(self.device)
        lib = abi.load()
        dev = torch.device("cuda", self.device)
        if int(camera.height) != shard.height:
            raise ValueError(f"shard is for {shard.height} rows, camera has {camera.height}")
        rows = int(lib.rt_strip_shard_rows(int(camera.height), shard.strip_rows, shard.rank, shard.world))
        if rows != shard.n_rows:
            raise RuntimeError(f"host and library disagree on the shard size: {shard.n_rows} vs {rows}")
        n = rows * int(camera.width)
        if planes is None:
            planes = torch.empty((3, n), dtype=self._torch_dtype(), device=dev)
        if rgba8 is None and want_rgba8:
            rgba8 = torch.empty((n, 3 if flags & abi.RT_FLAG_RGB8 else 4), dtype=torch.uint8, device=dev)
        stats = self._new_stats()
        if self._wants_accel(flat):
            self.upload_scene(flat)
        env = self._env(flat, flags)
        spheres_ptr, lights_ptr = self._scene_pointers(flat, env)
        cam = abi.make_camera(as_float_triple(camera.position, "camera position"), camera.width, camera.height)
        abi.check(
            lib.rt_render_strips_dev(
                self.device, spheres_ptr, flat.n_spheres, lights_ptr, C.byref(env), C.byref(cam),
                shard.strip_rows, shard.rank, shard.world, planes.data_ptr(), self._abi_dtype(), max(planes.stride(0), 1),
                None if rgba8 is None else rgba8.data_ptr(), None if stats is None else stats.data_ptr(),
                self._stream(),
            ),
            "rt_render_strips_dev",
        )
        return DeviceColor(planes, rgba8, stats, self.device)

    def render_strips_into(self, flat: FlatScene, camera: Camera, shard: Any, frame_rgba8_ptr: int, *,
                           flags: int = 0) -> None:
        """Render this rank's strips straight into a whole-frame RGBA8 buffer at
        ``frame_rgba8_ptr`` (``RT_FLAG_IMAGE_ORDER``): local memory or a peer GPU's
        framebuffer (``sharding.PeerFrame``).  No linear planes are written.  ``flags`` as in
        ``render_rows`` (a ``PeerFrame(channels=3)`` needs ``RT_FLAG_RGB8``; peer frames of many
        writers want ``RT_FLAG_WIDE_STORES``).  Asynchronous on torch's current stream."""
        _require_cuda(self.device)
        lib = abi.load()
        spheres, lights = self.upload_scene(flat)
        env = self._env(flat, abi.RT_FLAG_IMAGE_ORDER | flags)
        cam = abi.make_camera(as_float_triple(camera.position, "camera position"), camera.width, camera.height)
        abi.check(
            lib.rt_render_strips_dev(
                self.device, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env), C.byref(cam),
                shard.strip_rows, shard.rank, shard.world, None, abi.RT_DTYPE_F32, 0, frame_rgba8_ptr, None,
                self._stream(),
            ),
            "rt_render_strips_dev",
        )

    def render_frame_sharded(self, scene: Scene3D, rank: int, world: int, *, strip_rows: Optional[int] = None,
                             dst: int = 0, group: Any = None):
        """One frame over ``world`` ranks (one process per GPU, ``torch.distributed``
        initialised by the caller): this rank renders its strips, the 8-bit shards are
        gathered to rank ``dst``.  Returns the ``[height, width, 4]`` RGBA8 device tensor in
        image order on ``dst`` and ``None`` elsewhere."""
        from . import sharding

        flat = self.flatten(scene)
        camera = scene.camera
        width = int(camera.width)
        shard = sharding.StripShard(int(camera.height), world, rank, strip_rows or sharding.DEFAULT_STRIP_ROWS)
        color = self.render_strips(flat, camera, shard, want_rgba8=True)
        local = color.rgba8.view(shard.n_rows, width, 4)
        return sharding.gather_frame(local, shard, dst=dst, group=group)

    def raytrace_scene(self, ray_origin: Any, normalized_ray_direction: Any, scene: Scene3D) -> DeviceColor:
        """Colour of every ray (``base.py:91-121`` with every ``NumpyShader.create`` it
        triggers, ``shader.py:63-112``).

        Handed the renderer's own ``CameraRays`` (what ``render_image_pipeline`` does,
        ``application.py:48-50``) it launches the camera kernel for the 8-bit image only -- into a
        device buffer reused across frames -- and renders the linear planes the first time
        ``.x/.y/.z`` is read; explicit arrays go through ``trace_rays``."""
        flat = self.flatten(scene)
        rays = normalized_ray_direction
        if isinstance(rays, CameraRays):
            origin = as_float_triple(ray_origin, "ray origin")
            # the camera kernel takes "towards the camera" (shader.py:76) from the ray itself: camera rays
            # that do not start at scene.camera.position are traced as explicit rays
            if origin == as_float_triple(rays.camera.position, "camera position") == tuple(flat.eye):
                return self._render_camera_lazy(flat, rays)
        return self.trace_rays(flat, ray_origin, rays)

    def _render_camera_lazy(self, flat: FlatScene, rays: "CameraRays") -> DeviceColor:
        torch = _require_cuda(self.device)
        camera, row0, row1 = rays.camera, rays.row0, rays.row1
        n = (row1 - row0) * int(camera.width)
        if self._rgb8_dev is None or self._rgb8_dev.shape[0] != n:
            self._rgb8_dev = torch.empty((n, 3), dtype=torch.uint8, device=torch.device("cuda", self.device))
        with _Nvtx("rt: trace kernel (8-bit image)"):
            color = self.render_rows(flat, camera, row0, row1, rgba8=self._rgb8_dev, flags=abi.RT_FLAG_RGB8, no_planes=True)
        color._render_planes = lambda: self.render_rows(flat, camera, row0, row1).planes
        return color

    def trace_rays(self, flat: FlatScene, ray_origin: Any, directions: Any) -> DeviceColor:
        """Explicit ray batch: ``directions`` has flat ``x, y, z`` arrays; ``ray_origin`` is
        a scalar triple (broadcast, as the reference's primary rays) or arrays."""
        torch = _require_cuda(self.device)
        lib = abi.load()
        dev = torch.device("cuda", self.device)
        d = np.stack([np.asarray(c, dtype=np.float64).reshape(-1) for c in (directions.x, directions.y, directions.z)])
        n = d.shape[1]
        o_parts = [np.asarray(c, dtype=np.float64) for c in (ray_origin.x, ray_origin.y, ray_origin.z)]
        origin0 = None
        origins_dev = None
        if all(p.ndim == 0 for p in o_parts):
            origin0 = (C.c_double * 3)(*[float(p) for p in o_parts])
        else:
            o = np.stack([np.broadcast_to(p.reshape(-1) if p.ndim else p, (n,)) for p in o_parts])
            origins_dev = torch.from_numpy(np.ascontiguousarray(o)).to(dev)
        d_dev = torch.from_numpy(np.ascontiguousarray(d)).to(dev)
        spheres, lights = self.upload_scene(flat)
        planes = torch.empty((3, n), dtype=self._torch_dtype(), device=dev)
        stats = self._new_stats()
        env = self._env(flat)
        abi.check(
            lib.rt_trace_rays_dev(
                self.device, spheres.data_ptr(), flat.n_spheres, lights.data_ptr(), C.byref(env),
                None if origins_dev is None else origins_dev.data_ptr(), origin0, d_dev.data_ptr(), n,
                planes.data_ptr(), self._abi_dtype(), max(n, 1), None if stats is None else stats.data_ptr(),
                self._stream(),
            ),
            "rt_trace_rays_dev",
        )
        torch.cuda.current_stream(self.device).synchronize()  # d_dev/origins_dev die with this frame
        return DeviceColor(planes, None, stats, self.device)

    def quantize(self, color: Any, channels: int = 3):
        """``uint8(255*clip(v,0,1))`` truncating (``base.py:143-151``) on the device;
        returns a ``[n, channels]`` uint8 device tensor."""
        torch = _require_cuda(self.device)
        lib = abi.load()
        dev = torch.device("cuda", self.device)
        if isinstance(color, DeviceColor):
            planes = color.planes
        else:  # a host colour (e.g. the reference's own NumpyRGBColor of arrays)
            host = np.stack([np.asarray(c, dtype=np.float64).reshape(-1) for c in (color.x, color.y, color.z)])
            planes = torch.from_numpy(np.ascontiguousarray(host)).to(dev)
        n = planes.shape[1]
        out = torch.empty((n, channels), dtype=torch.uint8, device=dev)
        dtype = abi.RT_DTYPE_F64 if planes.dtype == torch.float64 else abi.RT_DTYPE_F32
        abi.check(
            lib.rt_quantize_dev(self.device, planes.data_ptr(), dtype, planes.stride(0), n, out.data_ptr(), channels,
                                self._stream()),
            "rt_quantize_dev",
        )
        return out

    def save_image(self, color: Any, camera: Camera, output_path: Path) -> None:
        """Quantise and write the PNG (``base.py:143-151``).  A frame in which no ray hit
        anything is written black (the reference raises on it).

        The 8-bit image comes from the trace kernel's fused quantisation when ``color`` carries it,
        through ``rt_quantize_dev`` otherwise; it lands in a pinned buffer reused across frames and is
        deflated in row bands on ``png_threads`` threads.  ``last_timing`` holds the stage times (ms)."""
        import time

        torch = _require_cuda(self.device)
        width, height = int(camera.width), int(camera.height)
        t0 = time.perf_counter()
        if isinstance(color, DeviceColor) and color.rgba8 is not None:
            dev8 = color.rgba8
        else:
            dev8 = self.quantize(color, 3)
        if dev8.shape[0] != width * height:
            raise ValueError(f"colour has {dev8.shape[0]} pixels, camera is {width}x{height}")
        channels = int(dev8.shape[1])
        if self.png_encoder == "device" and 1 + 3 * width <= 65000:
            from ..png import DevicePngEncoder

            with _Nvtx("rt: PNG encode on the device + file to host"):
                enc = self._png_device
                if enc is None or (enc.width, enc.height) != (width, height):
                    enc = self._png_device = DevicePngEncoder(width, height, self.device)
                stream = torch.cuda.current_stream(self.device)
                enc.enqueue(dev8, int(stream.cuda_stream))
                stream.synchronize()  # the trace kernel, the encoder and the file's way to the host
                abi.check(abi.load().rt_device_status(self.device), "rt_device_status")
                data = enc.finish()
            t1 = time.perf_counter()
            with _Nvtx("rt: PNG write"), open(output_path, "wb") as f:
                f.write(data)
            t2 = time.perf_counter()
            self.last_timing = {"kernel_and_d2h_ms": (t1 - t0) * 1e3, "png_ms": (t2 - t1) * 1e3, "png_bytes": len(data),
                                "encoder": "device (rt_png_encode_bands_dev + rt_png_pack_dev); kernel_and_d2h_ms includes the encode"}
            return
        with _Nvtx("rt: 8-bit image to host"):
            if self._rgb8_pinned is None or tuple(self._rgb8_pinned.shape) != (width * height, channels):
                self._rgb8_pinned = torch.empty((width * height, channels), dtype=torch.uint8).pin_memory()
            self._rgb8_pinned.copy_(dev8, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()  # the kernel and the copy
            abi.check(abi.load().rt_device_status(self.device), "rt_device_status")
        t1 = time.perf_counter()
        with _Nvtx("rt: PNG encode + write"):
            if self._png_pool is None and self.png_threads > 1:
                from concurrent.futures import ThreadPoolExecutor

                self._png_pool = ThreadPoolExecutor(self.png_threads)
            pixels = self._rgb8_pinned.numpy().reshape(height, width, channels)
            if channels == 4 and self.png_compress_level is not None:
                pixels = pixels[:, :, :3]
            write_png(Path(output_path), pixels, self.png_compress_level, self.png_threads, self._png_pool)
        t2 = time.perf_counter()
        self.last_timing = {"kernel_and_d2h_ms": (t1 - t0) * 1e3, "png_ms": (t2 - t1) * 1e3,
                            "encoder": "rt_png_deflate_rows x%d threads" % self.png_threads if self.png_compress_level is None
                            else "zlib level %d x%d threads" % (self.png_compress_level, self.png_threads)}


def sphere_distance(sphere: Any, ray_origin: Any, normalized_ray_direction: Any, *, device: int = 0,
                    precision: str = "exact") -> np.ndarray:
    """``NumpySphere.intersect`` (``shape.py:28-51``) on the GPU: distance per ray,
    ``FARAWAY`` on a miss."""
    torch = _require_cuda(device)
    lib = abi.load()
    dev = torch.device("cuda", device)
    d = np.stack([np.asarray(c, dtype=np.float64).reshape(-1)
                  for c in (normalized_ray_direction.x, normalized_ray_direction.y, normalized_ray_direction.z)])
    n = d.shape[1]
    o_parts = [np.asarray(c, dtype=np.float64) for c in (ray_origin.x, ray_origin.y, ray_origin.z)]
    origin0, origins_dev = None, None
    if all(p.ndim == 0 for p in o_parts):
        origin0 = (C.c_double * 3)(*[float(p) for p in o_parts])
    else:
        o = np.stack([np.broadcast_to(p.reshape(-1) if p.ndim else p, (n,)) for p in o_parts])
        origins_dev = torch.from_numpy(np.ascontiguousarray(o)).to(dev)
    d_dev = torch.from_numpy(np.ascontiguousarray(d)).to(dev)
    out = torch.empty(n, dtype=torch.float64, device=dev)
    centre = (C.c_double * 3)(*as_float_triple(sphere.center, "sphere center"))
    abi.check(
        lib.rt_sphere_distance_dev(device, None if origins_dev is None else origins_dev.data_ptr(), origin0,
                                   d_dev.data_ptr(), n, centre, float(sphere.radius), _PRECISIONS[precision],
                                   out.data_ptr(), int(torch.cuda.current_stream(device).cuda_stream)),
        "rt_sphere_distance_dev",
    )
    return out.cpu().numpy()
