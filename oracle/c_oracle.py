"""ctypes wrapper of ``librt_oracle.so`` (``rt_oracle.c``).  TEST INFRASTRUCTURE ONLY.

See ``oracle/__init__.py`` for who may import this.  ``build()`` compiles the C
restatement with the recipe in ``oracle/Makefile``.
"""

from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path
from typing import Optional, Sequence, Tuple

import numpy as np

from python_ray_tracer_b200.infrastructure.cuda.abi import (
    RT_FLAG_ALL_POINT_LIGHTS,
    RT_FLAG_NONE,
    RT_STATS_WORDS,
    RtCamera,
    RtEnv,
    make_camera,
    make_env,
)

HERE = Path(__file__).resolve().parent
LIBRARY = HERE / "librt_oracle.so"
_lib: Optional[C.CDLL] = None

_P = C.c_void_p


def build(force: bool = False) -> Path:
    """Compile ``rt_oracle.c`` if the library is missing or older than its sources."""
    sources = [HERE / "rt_oracle.c", HERE.parent / "include" / "rt_abi.h"]
    stale = not LIBRARY.exists() or any(s.stat().st_mtime > LIBRARY.stat().st_mtime for s in sources)
    if force or stale:
        subprocess.run(["make", "-C", str(HERE), "-B", "librt_oracle.so"], check=True, capture_output=True)
    return LIBRARY


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIBRARY.exists():
            build()
        lib = C.CDLL(str(LIBRARY))
        lib.orc_sphere_distance.restype = C.c_double
        lib.orc_sphere_distance.argtypes = [_P, _P, _P, C.c_double]
        lib.orc_ray_directions.restype = C.c_int
        lib.orc_ray_directions.argtypes = [C.POINTER(RtCamera), _P, C.c_int64, C.c_int64, _P]
        lib.orc_trace_rays.restype = C.c_int
        lib.orc_trace_rays.argtypes = [_P, C.c_int32, _P, C.POINTER(RtEnv), _P, _P, _P, C.c_int64,
                                       _P, _P, C.c_int]
        lib.orc_render_pixels.restype = C.c_int
        lib.orc_render_pixels.argtypes = [_P, C.c_int32, _P, C.POINTER(RtEnv), C.POINTER(RtCamera),
                                          _P, C.c_int64, C.c_int64, _P, _P, C.c_int]
        lib.orc_quantize.restype = C.c_int
        lib.orc_quantize.argtypes = [_P, C.c_int64, _P]
        lib.orc_max_threads.restype = C.c_int
        lib.orc_max_threads.argtypes = []
        _lib = lib
    return _lib


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(_P)


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def env_of(flat, max_depth: int = 0, all_point_lights: bool = False) -> RtEnv:
    return make_env(
        flat.eye,
        len(flat.point_lights),
        flat.dome_intensity,
        flat.dome_color,
        max_depth=max_depth,
        flags=RT_FLAG_ALL_POINT_LIGHTS if all_point_lights else RT_FLAG_NONE,
    )


def max_threads() -> int:
    return int(load().orc_max_threads())


def sphere_distance(origin: Sequence[float], direction: Sequence[float], centre: Sequence[float],
                    radius: float) -> float:
    o, d, c = _f64(origin), _f64(direction), _f64(centre)
    return float(load().orc_sphere_distance(_ptr(o), _ptr(d), _ptr(c), float(radius)))


def ray_directions(cam_position, width: int, height: int, pixels: Optional[np.ndarray] = None) -> np.ndarray:
    cam = make_camera(cam_position, width, height)
    if pixels is None:
        n, px = width * height, None
    else:
        px = np.ascontiguousarray(pixels, dtype=np.int64)
        n = px.size
    out = np.empty((3, n), dtype=np.float64)
    rc = load().orc_ray_directions(C.byref(cam), _ptr(px), 0, n, _ptr(out))
    assert rc == 0, rc
    return out


def render(flat, width: int, height: int, pixels: Optional[np.ndarray] = None, first: int = 0,
           count: Optional[int] = None, max_depth: int = 0, all_point_lights: bool = False,
           threads: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Linear RGB float64 ``[3, n]`` and the 4 work counters for camera pixels.

    ``pixels`` (flat indices) selects a subset; otherwise ``first``/``count`` select a
    contiguous run (default: the whole frame)."""
    cam = make_camera(flat.eye, width, height)
    env = env_of(flat, max_depth, all_point_lights)
    spheres, lights = _f64(flat.spheres), _f64(flat.point_lights)
    if pixels is not None:
        px = np.ascontiguousarray(pixels, dtype=np.int64)
        n = px.size
    else:
        px = None
        n = width * height - first if count is None else count
    rgb = np.empty((3, n), dtype=np.float64)
    stats = np.zeros(RT_STATS_WORDS, dtype=np.uint64)
    rc = load().orc_render_pixels(_ptr(spheres), spheres.shape[1], _ptr(lights), C.byref(env),
                                  C.byref(cam), _ptr(px), first, n, _ptr(rgb), _ptr(stats), threads)
    assert rc == 0, rc
    return rgb, stats


def trace_rays(flat, origins, directions: np.ndarray, max_depth: int = 0,
               all_point_lights: bool = False, threads: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """``origins``: ``[3, n]`` array or a 3-sequence broadcast to every ray."""
    env = env_of(flat, max_depth, all_point_lights)
    spheres, lights = _f64(flat.spheres), _f64(flat.point_lights)
    d = _f64(directions)
    n = d.shape[1]
    o = _f64(origins)
    o_arr, o0 = (o, None) if o.ndim == 2 else (None, o)
    rgb = np.empty((3, n), dtype=np.float64)
    stats = np.zeros(RT_STATS_WORDS, dtype=np.uint64)
    rc = load().orc_trace_rays(_ptr(spheres), spheres.shape[1], _ptr(lights), C.byref(env),
                               _ptr(o_arr), _ptr(o0), _ptr(d), n, _ptr(rgb), _ptr(stats), threads)
    assert rc == 0, rc
    return rgb, stats


def quantize(rgb: np.ndarray) -> np.ndarray:
    """``[3, n]`` float64 -> ``[n, 3]`` uint8, truncating (base.py:143-151)."""
    rgb = _f64(rgb)
    n = rgb.shape[1]
    out = np.empty((n, 3), dtype=np.uint8)
    rc = load().orc_quantize(_ptr(rgb), n, _ptr(out))
    assert rc == 0, rc
    return out
