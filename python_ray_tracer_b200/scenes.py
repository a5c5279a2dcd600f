"""Builders for the scenes BASELINE.json names.

They produce ordinary ``Scene3D`` objects through the public classes, exactly as a
user script would; ``bench.py``, the tests and the golden-fixture generator share them
so every leg renders the same thing.
"""

from __future__ import annotations

import math
from typing import Tuple

import numpy as np

from .domain import Camera, DomeLight, PointLight, Scene3D
from .infrastructure.materials import NumpyShader, Texture, TextureChecker
from .infrastructure.shapes import CheckeredSphere, NumpySphere
from .infrastructure.values import NumpyRGBColor, NumpyVector3D

# README.md:50-56 -- the gains the README example hands to its renderer.
README_GAINS = dict(
    reflection_gain=0.8,
    specular_gain=1.0,
    specular_roughness=0.5,
    iridescence_gain=0.05,
    diffuse_gain=1.0,
)


def main_scene(width: int = 960, height: int = 540) -> Scene3D:
    """The scene of the reference's ``main.py:13-51`` (golden scene G0: its 8-bit
    render at 960x540 is the reference's committed ``render.png``)."""
    return Scene3D(
        [
            NumpySphere(
                NumpyVector3D(0.55, 0.5, 3),
                1.0,
                NumpyShader(0.0, 0, 0.01, 0, 0.0, Texture(NumpyRGBColor(1, 1, 1))),
            ),
            NumpySphere(
                NumpyVector3D(-0.45, 0.1, 1),
                0.4,
                NumpyShader(0, 1, 0.1, 0.0, 0.0, Texture(NumpyRGBColor(1, 0, 0))),
            ),
            NumpySphere(
                NumpyVector3D(0, -99999.5, 0),
                99999,
                NumpyShader(0.0, 0.1, 0.5, 0.0, 1.0, TextureChecker()),
            ),
        ],
        [PointLight(NumpyVector3D(-2, 1, 2)), DomeLight(0.1, NumpyRGBColor(1, 1, 1))],
        Camera(NumpyVector3D(0, 0.2, -2), width, height),
    )


def readme_scene(width: int = 1920, height: int = 1080) -> Scene3D:
    """The README example (``README.md:58-84``) through the README surface: shapes carry
    colours and take the renderer's default gains (C1 at 960x540, C2 at 1080p, C4 at 8K)."""
    return Scene3D(
        [
            NumpySphere(NumpyVector3D(0.55, 0.5, 3), 1.0, NumpyRGBColor(1, 0, 1)),
            NumpySphere(NumpyVector3D(-0.45, 0.1, 1), 0.4, NumpyRGBColor(0.5, 0.5, 0.5)),
            CheckeredSphere(NumpyVector3D(0, -99999.5, 0), 99999, NumpyRGBColor(1, 1, 1), 1.0),
        ],
        [PointLight(NumpyVector3D(5, 10, -10)), DomeLight(0.1, NumpyRGBColor(1, 1, 1))],
        Camera(NumpyVector3D(0, 0.2, -2), width, height),
    )


def readme_scene_head_api(width: int = 1920, height: int = 1080) -> Scene3D:
    """The same README scene written against the HEAD constructor surface (every shape
    carries its own ``NumpyShader``); must flatten to the same table as ``readme_scene``."""

    def shader(texture: Texture) -> NumpyShader:
        return NumpyShader(diffuse_color=texture, **README_GAINS)

    return Scene3D(
        [
            NumpySphere(NumpyVector3D(0.55, 0.5, 3), 1.0, shader(Texture(NumpyRGBColor(1, 0, 1)))),
            NumpySphere(
                NumpyVector3D(-0.45, 0.1, 1), 0.4, shader(Texture(NumpyRGBColor(0.5, 0.5, 0.5)))
            ),
            NumpySphere(NumpyVector3D(0, -99999.5, 0), 99999, shader(TextureChecker())),
        ],
        [PointLight(NumpyVector3D(5, 10, -10)), DomeLight(0.1, NumpyRGBColor(1, 1, 1))],
        Camera(NumpyVector3D(0, 0.2, -2), width, height),
    )


def random_spheres_scene(
    width: int = 3840, height: int = 2160, n_spheres: int = 1024, seed: int = 1234
) -> Scene3D:
    """C3: ``n_spheres`` random spheres resting on the README checker ground, four point
    lights (the reference shades with the first only) and the README dome light.

    Draw order per sphere: radius U(0.08, 0.35), x U(-8, 8),
    z U(0, 24), colour U(0.1, 1.0)^3; centre ``(x, radius - 0.5, z)``.
    """
    rng = np.random.default_rng(seed)
    shapes = []
    for _ in range(n_spheres):
        radius = float(rng.uniform(0.08, 0.35))
        x = float(rng.uniform(-8.0, 8.0))
        z = float(rng.uniform(0.0, 24.0))
        r, g, b = (float(v) for v in rng.uniform(0.1, 1.0, size=3))
        shapes.append(NumpySphere(NumpyVector3D(x, radius - 0.5, z), radius, NumpyRGBColor(r, g, b)))
    shapes.append(CheckeredSphere(NumpyVector3D(0, -99999.5, 0), 99999, NumpyRGBColor(1, 1, 1), 1.0))
    lights = [
        PointLight(NumpyVector3D(5, 10, -10)),
        PointLight(NumpyVector3D(-5, 10, -10)),
        PointLight(NumpyVector3D(5, 10, 10)),
        PointLight(NumpyVector3D(-5, 10, 10)),
        DomeLight(0.1, NumpyRGBColor(1, 1, 1)),
    ]
    return Scene3D(shapes, lights, Camera(NumpyVector3D(0, 0.2, -2), width, height))


def _rotate_about_pivot(x: float, z: float, theta: float, pivot: Tuple[float, float]) -> Tuple[float, float]:
    px, pz = pivot
    dx, dz = x - px, z - pz
    c, s = math.cos(theta), math.sin(theta)
    return px + c * dx + s * dz, pz - s * dx + c * dz


def turntable_scene(frame: int, n_frames: int = 256, width: int = 1920, height: int = 1080) -> Scene3D:
    """C5 frame ``frame`` of ``n_frames``: the README scene with the two small spheres and
    the point light rotated by ``2*pi*frame/n_frames`` about the vertical axis through
    ``(x=0, z=2)``.  The reference camera has no orientation (``base.py:130-141``), so the
    turntable turns the scene; ground, checker and camera stay fixed."""
    theta = 2.0 * math.pi * frame / n_frames
    pivot = (0.0, 2.0)
    x0, z0 = _rotate_about_pivot(0.55, 3.0, theta, pivot)
    x1, z1 = _rotate_about_pivot(-0.45, 1.0, theta, pivot)
    lx, lz = _rotate_about_pivot(5.0, -10.0, theta, pivot)
    return Scene3D(
        [
            NumpySphere(NumpyVector3D(x0, 0.5, z0), 1.0, NumpyRGBColor(1, 0, 1)),
            NumpySphere(NumpyVector3D(x1, 0.1, z1), 0.4, NumpyRGBColor(0.5, 0.5, 0.5)),
            CheckeredSphere(NumpyVector3D(0, -99999.5, 0), 99999, NumpyRGBColor(1, 1, 1), 1.0),
        ],
        [PointLight(NumpyVector3D(lx, 10, lz)), DomeLight(0.1, NumpyRGBColor(1, 1, 1))],
        Camera(NumpyVector3D(0, 0.2, -2), width, height),
    )
