"""Sphere shapes accepted by the scene flattener.

``NumpySphere`` mirrors the reference's constructor at HEAD
(``ray_tracer/infrastructure/numpy/shape.py:16-26``: ``center, radius, shader``) and
also takes the README's form (``README.md:60-77``: ``center, radius, colour``), in
which the material is the renderer's default gains with that colour.
``CheckeredSphere`` is the README's checker-ground shape (``README.md:66-76``).
"""

from __future__ import annotations

from typing import Any

from ..domain import Shape
from .materials import NumpyShader, Texture, TextureChecker
from .values import NumpyRGBColor


class NumpySphere(Shape):
    """Sphere ``(center, radius)`` with a ``NumpyShader`` or a plain colour."""

    def __init__(self, center: Any, radius: float, shader: Any) -> None:
        self.center = center
        self.position = center  # alias kept by the reference (shape.py:24)
        self.radius = radius
        self.shader = shader

    def material_texture(self) -> Texture:
        """Texture used when ``shader`` is a colour rather than a ``NumpyShader``."""
        return Texture(self.shader)

    def intersect(self, ray_origin: Any, normalized_ray_direction: Any) -> Any:
        """Distance along each ray to this sphere, ``FARAWAY`` on a miss.

        Replaces ``NumpySphere.intersect`` (``shape.py:28-51``); evaluated on the GPU
        through ``rt_sphere_distance_host``.
        """
        from .cuda.renderer import sphere_distance

        return sphere_distance(self, ray_origin, normalized_ray_direction)

    def diffusecolor(self, intersection_point: Any) -> NumpyRGBColor:
        """Constant colour of the shape's texture (the reference's version is dead code)."""
        if isinstance(self.shader, NumpyShader):
            return self.shader.diffuse_color.color
        return self.shader


class CheckeredSphere(NumpySphere):
    """Sphere textured with the XZ checker (``README.md:66-76``).

    The README passes a colour and a fourth number; the checker ignores the colour
    (``shader.py:29-32``) and nothing at HEAD reads the number, so both are accepted
    and ignored.
    """

    def __init__(self, center: Any, radius: float, color: Any = None, *_ignored: Any) -> None:
        super().__init__(center, radius, color if color is not None else NumpyRGBColor(1, 1, 1))

    def material_texture(self) -> Texture:
        return TextureChecker(self.shader)
