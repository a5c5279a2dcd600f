"""ctypes view of ``include/rt_abi.h`` and the loader of ``librt_b200.so``.

The structures and prototypes below are a 1:1 transcription of the header.  The
library is the only compute path: if it is missing or fails to load, ``load()``
raises -- there is no CPU fallback.
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional, Sequence

RT_ABI_VERSION = 3
RT_FARAWAY = 1.0e39
RT_MAX_DOME_LIGHTS = 8
RT_STATS_WORDS = 4
RT_SPHERE_FIELDS = 12

RT_MODE_EXACT = 0
RT_MODE_FAST = 1
RT_FLAG_NONE = 0
RT_FLAG_ALL_POINT_LIGHTS = 1
RT_FLAG_IMAGE_ORDER = 2
RT_FLAG_NO_ACCEL = 4
RT_FLAG_ACCEL_FAST_BUILD = 8
RT_FLAG_WIDE_STORES = 16
RT_FLAG_RGB8 = 32
RT_FLAG_EVEN_STORES = 64
RT_FLAG_GENERIC_KERNEL = 128
RT_SMALL_SCENE_SPHERES = 4
RT_ACCEL_AUTO_SPHERES = 32
RT_IPC_HANDLE_BYTES = 64
RT_PNG_STATIC_SCRATCH_BYTES = 8 << 20
RT_PNG_ONE_LAUNCH_BANDS = 2048
RT_DTYPE_F32 = 0
RT_DTYPE_F64 = 1

RT_OK = 0
STATUS_NAMES = {
    0: "RT_OK",
    -1: "RT_ERR_BAD_ARG",
    -2: "RT_ERR_CUDA",
    -3: "RT_ERR_NO_DEVICE",
    -4: "RT_ERR_UNSUPPORTED",
}


class RtEnv(C.Structure):
    _fields_ = [
        ("eye", C.c_double * 3),
        ("n_point_lights", C.c_int32),
        ("n_dome_lights", C.c_int32),
        ("dome_intensity", C.c_double * RT_MAX_DOME_LIGHTS),
        ("dome_color", C.c_double * 3),
        ("max_depth", C.c_int32),
        ("mode", C.c_int32),
        ("flags", C.c_int32),
        ("reserved", C.c_int32),
        ("cutoff", C.c_double),
        ("accel_dev", C.c_void_p),
        ("accel_bytes", C.c_int64),
        ("spheres_host", C.c_void_p),
        ("point_lights_host", C.c_void_p),
    ]


class RtCamera(C.Structure):
    _fields_ = [
        ("position", C.c_double * 3),
        ("width", C.c_int32),
        ("height", C.c_int32),
    ]


def make_env(
    eye: Sequence[float],
    n_point_lights: int,
    dome_intensity: Sequence[float],
    dome_color: Sequence[float],
    max_depth: int = 0,
    mode: int = RT_MODE_EXACT,
    flags: int = RT_FLAG_NONE,
    cutoff: float = 0.0,
    spheres_host=None,
    point_lights_host=None,
) -> RtEnv:
    """``spheres_host`` / ``point_lights_host``: the caller's C-contiguous float64 NumPy copies of the
    tables the call's device pointers hold (``rt_env.spheres_host``); the caller keeps them alive
    for the duration of the call."""
    env = RtEnv()
    env.eye[:] = [float(v) for v in eye]
    env.n_point_lights = int(n_point_lights)
    if len(dome_intensity) > RT_MAX_DOME_LIGHTS:
        raise ValueError("more dome lights than RT_MAX_DOME_LIGHTS; fold them on the host")
    env.n_dome_lights = len(dome_intensity)
    for i, value in enumerate(dome_intensity):
        env.dome_intensity[i] = float(value)
    env.dome_color[:] = [float(v) for v in dome_color]
    env.max_depth = int(max_depth or 0)
    env.mode = int(mode)
    env.flags = int(flags)
    env.cutoff = float(cutoff)
    if spheres_host is not None:
        env.spheres_host = spheres_host.ctypes.data
    if point_lights_host is not None:
        env.point_lights_host = point_lights_host.ctypes.data
    return env


def make_camera(position: Sequence[float], width: int, height: int) -> RtCamera:
    cam = RtCamera()
    cam.position[:] = [float(v) for v in position]
    cam.width = int(width)
    cam.height = int(height)
    return cam


# name -> (restype, argtypes); every symbol include/rt_abi.h declares.
_P = C.c_void_p
PROTOTYPES = {
    "rt_abi_version": (C.c_int, []),
    "rt_last_error": (C.c_char_p, []),
    "rt_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "rt_release_cache": (C.c_int, []),
    "rt_device_status": (C.c_int, [C.c_int]),
    "rt_measure_fma_peak": (C.c_int, [C.c_int, C.c_int32, C.POINTER(C.c_double)]),
    "rt_accel_bytes": (C.c_int64, [C.c_int32]),
    "rt_accel_build_host": (C.c_int, [_P, C.c_int32, C.POINTER(RtEnv), _P, C.c_int64, C.POINTER(C.c_int64)]),
    "rt_render_rows_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, _P, C.POINTER(RtEnv), C.POINTER(RtCamera), C.c_int32, C.c_int32,
         _P, C.c_int32, C.c_int64, _P, _P, _P],
    ),
    "rt_strip_shard_rows": (C.c_int64, [C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "rt_render_strips_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, _P, C.POINTER(RtEnv), C.POINTER(RtCamera), C.c_int32, C.c_int32, C.c_int32,
         _P, C.c_int32, C.c_int64, _P, _P, _P],
    ),
    "rt_peer_buffer_create": (C.c_int, [C.c_int, C.c_uint64, C.POINTER(C.c_void_p), C.c_char_p]),
    "rt_peer_buffer_open": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]),
    "rt_peer_buffer_close": (C.c_int, [C.c_int, _P]),
    "rt_peer_buffer_destroy": (C.c_int, [C.c_int, _P]),
    "rt_trace_rays_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, _P, C.POINTER(RtEnv), _P, C.POINTER(C.c_double), _P, C.c_int64,
         _P, C.c_int32, C.c_int64, _P, _P],
    ),
    "rt_ray_directions_dev": (
        C.c_int,
        [C.c_int, C.POINTER(RtCamera), C.c_int32, C.c_int32, _P, C.c_int64, _P],
    ),
    "rt_sphere_distance_dev": (
        C.c_int,
        [C.c_int, _P, C.POINTER(C.c_double), _P, C.c_int64, C.POINTER(C.c_double), C.c_double,
         C.c_int32, _P, _P],
    ),
    "rt_quantize_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, C.c_int64, C.c_int64, _P, C.c_int32, _P],
    ),
    "rt_render_frame_host": (
        C.c_int,
        [C.c_int, _P, C.c_int32, _P, C.POINTER(RtEnv), C.POINTER(RtCamera), C.c_int32, C.c_int32,
         _P, C.c_int32, _P, _P],
    ),
    "rt_trace_rays_host": (
        C.c_int,
        [C.c_int, _P, C.c_int32, _P, C.POINTER(RtEnv), _P, C.POINTER(C.c_double), _P, C.c_int64,
         _P, C.c_int32, _P],
    ),
    "rt_scene_travels_as_parameter": (C.c_int, [C.POINTER(RtEnv), C.c_int32]),
    "rt_png_rows_per_band": (C.c_int32, [C.c_int32, C.c_int32]),
    "rt_png_scratch_bytes": (C.c_int64, [C.c_int32, C.c_int32]),
    "rt_png_file_bound": (C.c_int64, [C.c_int32, C.c_int32]),
    "rt_png_encode_bands_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, C.c_int64, _P],
    ),
    "rt_png_pack_dev": (C.c_int, [C.c_int, C.c_int32, C.c_int32, _P, _P, C.c_int64, _P, _P]),
    "rt_render_png_host": (
        C.c_int,
        [C.c_int, _P, C.c_int32, _P, C.POINTER(RtEnv), C.POINTER(RtCamera), _P, C.c_int64, C.POINTER(C.c_int64)],
    ),
    "rt_png_encode_file_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, C.c_int64, C.c_int32, C.c_int32, _P, C.c_int64, _P, C.c_int64, _P, _P],
    ),
    "rt_png_encode_static_dev": (
        C.c_int,
        [C.c_int, _P, C.c_int32, C.c_int64, C.c_int32, C.c_int32, _P, C.c_int64, _P, _P],
    ),
    "rt_png_deflate_bound": (C.c_int64, [C.c_int32, C.c_int32]),
    "rt_png_deflate_rows": (
        C.c_int,
        [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int32, _P, C.c_int64,
         C.POINTER(C.c_int64), C.POINTER(C.c_uint32), C.POINTER(C.c_int64)],
    ),
}

LIBRARY_NAME = "librt_b200.so"
_lib: Optional[C.CDLL] = None


def library_path() -> Path:
    override = os.environ.get("RT_B200_LIBRARY")
    if override:
        return Path(override)
    return Path(__file__).resolve().parents[2] / "csrc" / LIBRARY_NAME


def load() -> C.CDLL:
    """Load ``librt_b200.so`` and bind every prototype; raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not path.exists():
        raise RuntimeError(
            f"{path} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C python_ray_tracer_b200/csrc`). There is no CPU fallback."
        )
    lib = C.CDLL(str(path))
    for name, (restype, argtypes) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here means header and library disagree
        fn.restype = restype
        fn.argtypes = argtypes
    got = lib.rt_abi_version()
    if got != RT_ABI_VERSION:
        raise RuntimeError(f"{path}: ABI version {got}, expected {RT_ABI_VERSION}")
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    """Raise for a non-zero ``rt_status`` (``ValueError`` for bad arguments)."""
    if status == RT_OK:
        return
    lib = load()
    detail = lib.rt_last_error()
    text = detail.decode("utf-8", "replace") if detail else ""
    name = STATUS_NAMES.get(status, str(status))
    if status == -1:
        raise ValueError(f"{what}: {name}: {text}")
    raise RuntimeError(f"{what}: {name}: {text}")
