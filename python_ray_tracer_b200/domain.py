"""Scene description value types consumed by the scene flattener.

Host-side mirror of the reference's domain layer (``ray_tracer/domain.py:5-59``): the
same class names, constructor argument order and attribute names, so a scene written
for the reference builds unchanged against this package.  These are plain data
holders; none of the ray/sphere arithmetic lives here (it is in ``csrc/``).
"""

from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Any, Sequence


class Vector3D:
    """A triple ``x, y, z`` (reference ``domain.py:5-7``)."""

    __slots__ = ("x", "y", "z")

    def __init__(self, x: Any, y: Any, z: Any) -> None:
        self.x = x
        self.y = y
        self.z = z

    def __repr__(self) -> str:  # pragma: no cover - cosmetic
        return f"{type(self).__name__}({self.x!r}, {self.y!r}, {self.z!r})"


class RGBColor(Vector3D):
    """A colour stored in the ``x, y, z`` slots (reference ``domain.py:10-11``)."""


@dataclass
class Camera:
    """Eye position and image size in pixels (reference ``domain.py:14-23``).

    The image plane is the fixed world plane ``z = 0``; there is no orientation
    or field of view (``base.py:130-141``).
    """

    position: Vector3D
    width: int
    height: int


@dataclass
class PointLight:
    """Omnidirectional light at ``position`` (reference ``domain.py:26-30``)."""

    position: Vector3D


@dataclass
class DomeLight:
    """Sky light: ``intensity * max(N.y, 0)`` tinted by ``color`` (``domain.py:33-40``)."""

    intensity: float
    color: RGBColor


class Shape(ABC):
    """Intersectable surface (reference ``domain.py:43-50``)."""

    @abstractmethod
    def intersect(self, ray_origin: Vector3D, normalized_ray_direction: Vector3D) -> Any:
        ...

    @abstractmethod
    def diffusecolor(self, intersection_point: Vector3D) -> RGBColor:
        ...


@dataclass
class Scene3D:
    """Shapes, lights and the camera of one frame (reference ``domain.py:53-59``)."""

    shapes: Sequence[Shape]
    lights: Sequence[Any]
    camera: Camera
