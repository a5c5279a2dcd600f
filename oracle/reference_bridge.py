"""Imports the UNMODIFIED reference from ``/root/reference``.  TEST INFRASTRUCTURE ONLY.

Usable only where the reference tree is mounted (the build container); the GPU box
does not have it, so nothing in ``-m gpu`` tests, ``smoke()`` or ``bench.py`` may
depend on this module at run time.  It is used by ``tests/golden/make_golden.py`` to
write the committed fixtures and by the container-only tests that compare the
restatements with the reference directly.

No reference code is copied: the reference's classes are imported and driven through
their own public API.  ``build_scene`` rebuilds a reference ``Scene3D`` from a
flattened table; ``Instrumented`` is a ``NumpyRenderer`` subclass that counts rays and
applies the depth cap the reference itself lacks.
"""

from __future__ import annotations

import os
import sys
from pathlib import Path
from types import SimpleNamespace
from typing import Optional, Tuple

import numpy as np

REFERENCE_ROOT = Path(os.environ.get("RT_REFERENCE_ROOT", "/root/reference"))


def available() -> bool:
    return (REFERENCE_ROOT / "ray_tracer" / "infrastructure" / "numpy" / "base.py").is_file()


def modules() -> SimpleNamespace:
    """The reference's modules, imported by full path (its package __init__s are empty)."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    root = str(REFERENCE_ROOT)
    if root not in sys.path:
        sys.path.insert(0, root)
    from ray_tracer import application, domain  # type: ignore
    from ray_tracer.infrastructure.numpy import base, shader, shape  # type: ignore

    return SimpleNamespace(application=application, domain=domain, base=base, shader=shader, shape=shape)


def build_scene(flat, width: int, height: int):
    """A reference ``Scene3D`` with one ``NumpySphere(NumpyShader)`` per SoA column.

    Point lights keep their order and come first; dome lights follow (only their
    relative order matters to ``shader.py:239-242``)."""
    m = modules()
    V, RGB = m.base.NumpyVector3D, m.base.NumpyRGBColor
    f = flat.spheres
    shapes = []
    for s in range(f.shape[1]):
        if f[11, s] != 0.0:
            texture = m.shader.TextureChecker()
        else:
            texture = m.shader.Texture(RGB(float(f[8, s]), float(f[9, s]), float(f[10, s])))
        material = m.shader.NumpyShader(
            reflection_gain=0.0,  # stored, never read (shader.py:45)
            specular_gain=float(f[5, s]),
            specular_roughness=float(f[6, s]),
            iridescence_gain=float(f[7, s]),
            diffuse_gain=float(f[4, s]),
            diffuse_color=texture,
        )
        shapes.append(m.shape.NumpySphere(V(float(f[0, s]), float(f[1, s]), float(f[2, s])), float(f[3, s]), material))
    lights = [m.domain.PointLight(V(*(float(v) for v in row))) for row in flat.point_lights]
    for i, inten in enumerate(flat.dome_intensity):
        last = i == len(flat.dome_intensity) - 1
        colour = RGB(*flat.dome_color) if last else RGB(1.0, 1.0, 1.0)
        lights.append(m.domain.DomeLight(float(inten), colour))
    camera = m.domain.Camera(V(*(float(v) for v in flat.eye)), int(width), int(height))
    return m.domain.Scene3D(shapes, lights, camera)


def instrumented_renderer(max_depth: int = 0):
    """``NumpyRenderer`` subclass: counts traced rays / generations and returns black for
    generations >= ``max_depth`` (0 = uncapped, the reference's own behaviour)."""
    m = modules()

    class Instrumented(m.base.NumpyRenderer):
        def __init__(self) -> None:
            self.generation = 0
            self.traced = 0
            self.deepest = 0

        def raytrace_scene(self, ray_origin, normalized_ray_direction, scene):
            if max_depth > 0 and self.generation >= max_depth:
                return m.base.NumpyRGBColor(0, 0, 0)
            self.traced += int(np.size(normalized_ray_direction.x))
            self.deepest = max(self.deepest, self.generation + 1)
            self.generation += 1
            try:
                return super().raytrace_scene(ray_origin, normalized_ray_direction, scene)
            finally:
                self.generation -= 1

    return Instrumented()


def render(scene, pixels: Optional[np.ndarray] = None, max_depth: int = 0) -> Tuple[np.ndarray, Tuple[int, int]]:
    """Linear RGB float64 ``[3, n]`` of ``scene`` through the reference's own
    ``get_ray_directions`` + ``raytrace_scene``; ``pixels`` slices the direction arrays
    (bit-identical to the full frame)."""
    m = modules()
    renderer = instrumented_renderer(max_depth)
    dirs = renderer.get_ray_directions(scene.camera)
    if pixels is not None:
        dirs = m.base.NumpyVector3D(dirs.x[pixels], dirs.y[pixels], dirs.z[pixels])
    n = int(np.size(dirs.x))
    color = renderer.raytrace_scene(scene.camera.position, dirs, scene)
    rgb = np.stack([np.broadcast_to(np.asarray(c, dtype=np.float64), (n,)) for c in color.components()])
    return np.ascontiguousarray(rgb), (renderer.traced, renderer.deepest)


def save_png(scene, rgb: np.ndarray, path) -> None:
    """The reference's own ``save_image`` (base.py:143-151) on a full-frame result."""
    m = modules()
    color = m.base.NumpyRGBColor(rgb[0], rgb[1], rgb[2])
    m.base.NumpyRenderer().save_image(color, scene.camera, path)
