"""Scene flattener: ``Scene3D`` objects -> the sphere SoA and light table of ``rt_abi.h``.

Inputs are the domain objects of ``python_ray_tracer_b200.domain`` or, by duck typing,
the reference's own (``ray_tracer/domain.py:14-59``, ``shape.py:17-26``,
``shader.py:13-54``), so one scene description can feed both backends.

Mapping rules:
  * ``shape.shader`` is a shader object (has the five gains)  -> use it;
  * ``shape.shader`` is a colour (has x, y, z)                -> renderer default gains
    with ``Texture(colour)``, or ``TextureChecker`` for a ``CheckeredSphere``;
  * ``lights[0]`` must be a point light (``shader.py:75``); dome lights anywhere in
    the list contribute (``shader.py:239-242``); further point lights are carried in
    the table and only used by the opt-in all-lights extension.
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

from .materials import TEXTURE_CHECKER, TEXTURE_CONSTANT
from .values import as_float_triple

# Row indices of the sphere SoA; must match rt_sphere_field in include/rt_abi.h.
F_CX, F_CY, F_CZ, F_RADIUS = 0, 1, 2, 3
F_DIFFUSE_GAIN, F_SPECULAR_GAIN, F_SPECULAR_ROUGHNESS, F_IRIDESCENCE_GAIN = 4, 5, 6, 7
F_COLOR_R, F_COLOR_G, F_COLOR_B, F_TEXTURE = 8, 9, 10, 11
SPHERE_FIELDS = 12
MAX_DOME_LIGHTS = 8

_GAIN_NAMES = ("diffuse_gain", "specular_gain", "specular_roughness", "iridescence_gain")


@dataclass
class DefaultMaterial:
    """Renderer-level gains of the README constructor (``README.md:50-56``)."""

    reflection_gain: float = 0.8
    specular_gain: float = 1.0
    specular_roughness: float = 0.5
    iridescence_gain: float = 0.05
    diffuse_gain: float = 1.0


@dataclass
class FlatScene:
    """Host copy of everything the kernels read, in ``rt_abi.h`` layout."""

    spheres: np.ndarray  # float64 [SPHERE_FIELDS, n_spheres], C-contiguous
    point_lights: np.ndarray  # float64 [n_point_lights, 3]
    eye: Tuple[float, float, float]
    dome_intensity: List[float] = field(default_factory=list)
    dome_color: Tuple[float, float, float] = (1.0, 1.0, 1.0)

    @property
    def n_spheres(self) -> int:
        return int(self.spheres.shape[1])

    def key(self) -> tuple:
        """Content key used to skip re-uploading an unchanged scene: a 64-bit multilinear
        fingerprint of the tables (one vectorised pass, no copy of the table) plus the small
        values themselves."""
        cached = self.__dict__.get("_key")
        if cached is None:
            cached = self.__dict__["_key"] = self._make_key()
        return cached

    def _make_key(self) -> tuple:
        return (
            self.spheres.shape,
            _fingerprint(self.spheres),
            self.point_lights.shape,
            _fingerprint(self.point_lights),
            tuple(self.eye),
            tuple(self.dome_color),
            tuple(self.dome_intensity),
        )


_WEIGHTS: Dict[int, np.ndarray] = {}


def _fingerprint(table: np.ndarray) -> int:
    """sum_i bits_i * w_i mod 2^64 with fixed odd random w_i: changes with any single entry and,
    with probability 1 - 2^-63, with any set of entries."""
    bits = np.ascontiguousarray(table, dtype=np.float64).reshape(-1).view(np.uint64)
    w = _WEIGHTS.get(bits.size)
    if w is None:
        w = _WEIGHTS[bits.size] = np.random.default_rng(0x5EED).integers(0, 2**63, bits.size, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    with np.errstate(over="ignore"):
        return int((bits * w).sum(dtype=np.uint64))


def _is_shader(obj: Any) -> bool:
    return all(hasattr(obj, name) for name in _GAIN_NAMES) and hasattr(obj, "diffuse_color")


def _is_checker(texture: Any) -> bool:
    kind = getattr(texture, "kind", None)
    if kind is not None:
        return kind == TEXTURE_CHECKER
    # the reference's classes carry no tag: go by class name through the MRO
    return any(cls.__name__ == "TextureChecker" for cls in type(texture).__mro__)


def _is_dome_light(light: Any) -> bool:
    return hasattr(light, "intensity") and hasattr(light, "color")


def _material_row(shape: Any, default: DefaultMaterial) -> Dict[int, float]:
    shader = getattr(shape, "shader", None)
    if shader is None:
        raise ValueError(f"shape {shape!r} has no shader or colour")
    if _is_shader(shader):
        gains, texture = shader, shader.diffuse_color
    elif all(hasattr(shader, axis) for axis in "xyz"):
        gains = default
        texture = shape.material_texture() if hasattr(shape, "material_texture") else None
        if texture is None:
            raise ValueError("a colour material needs a shape that names its texture")
    else:
        raise ValueError(
            f"shape material must be a NumpyShader or a colour, got {type(shader).__name__}"
        )
    checker = _is_checker(texture)
    colour = (1.0, 1.0, 1.0) if checker else as_float_triple(texture.color, "texture colour")
    return {
        F_DIFFUSE_GAIN: float(gains.diffuse_gain),
        F_SPECULAR_GAIN: float(gains.specular_gain),
        F_SPECULAR_ROUGHNESS: float(gains.specular_roughness),
        F_IRIDESCENCE_GAIN: float(gains.iridescence_gain),
        F_COLOR_R: colour[0],
        F_COLOR_G: colour[1],
        F_COLOR_B: colour[2],
        F_TEXTURE: TEXTURE_CHECKER if checker else TEXTURE_CONSTANT,
    }


def flatten_scene(scene: Any, default: Optional[DefaultMaterial] = None) -> FlatScene:
    """Flatten ``scene`` (shapes, lights, camera position); raises ``ValueError`` on a
    scene the reference could not shade either."""
    default = default or DefaultMaterial()
    shapes = list(scene.shapes)
    if not shapes:
        raise ValueError("scene has no shapes")  # reduce() of an empty list at base.py:98
    spheres = np.zeros((SPHERE_FIELDS, len(shapes)), dtype=np.float64)
    for s, shape in enumerate(shapes):
        centre = getattr(shape, "center", None)
        if centre is None:
            centre = getattr(shape, "position", None)
        if centre is None or not hasattr(shape, "radius"):
            raise ValueError(f"shape {s} is not a sphere (needs center and radius)")
        cx, cy, cz = as_float_triple(centre, f"shape {s} center")
        radius = float(shape.radius)
        if radius == 0.0:
            raise ValueError(f"shape {s} has zero radius")  # 1.0 / radius at shader.py:74
        spheres[F_CX, s], spheres[F_CY, s], spheres[F_CZ, s] = cx, cy, cz
        spheres[F_RADIUS, s] = radius
        for row, value in _material_row(shape, default).items():
            spheres[row, s] = value

    lights = list(scene.lights)
    if not lights or _is_dome_light(lights[0]) or not hasattr(lights[0], "position"):
        # the reference raises AttributeError at shader.py:75 for the same scenes
        raise ValueError("scene.lights[0] must be a PointLight")
    points = [as_float_triple(l.position, "light position") for l in lights if not _is_dome_light(l)]
    domes = [l for l in lights if _is_dome_light(l)]
    dome_intensity = [float(l.intensity) for l in domes]
    if len(dome_intensity) > MAX_DOME_LIGHTS:  # fold the excess into the last slot
        head, tail = dome_intensity[: MAX_DOME_LIGHTS - 1], dome_intensity[MAX_DOME_LIGHTS - 1 :]
        dome_intensity = head + [float(np.sum(tail))]
    dome_color = as_float_triple(domes[-1].color, "dome colour") if domes else (1.0, 1.0, 1.0)

    return FlatScene(
        spheres=np.ascontiguousarray(spheres),
        point_lights=np.ascontiguousarray(np.asarray(points, dtype=np.float64).reshape(-1, 3)),
        eye=as_float_triple(scene.camera.position, "camera position"),
        dome_intensity=dome_intensity,
        dome_color=dome_color,
    )
