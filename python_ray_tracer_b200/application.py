"""The renderer plug-in port and the three-call pipeline that drives it.

Host-side mirror of the reference's application layer
(``ray_tracer/application.py:7-52``).  ``Renderer`` keeps the reference's three
abstract methods with the same names and argument meaning, and
``render_image_pipeline`` makes the same three calls in the same order, so
``CudaRenderer`` plugs in exactly where ``NumpyRenderer`` does.
"""

from __future__ import annotations

from abc import ABC, abstractmethod
from pathlib import Path
from typing import Any

from .domain import Camera, RGBColor, Scene3D, Vector3D


class Renderer(ABC):
    """Port implemented by a rendering backend (reference ``application.py:7-32``)."""

    @abstractmethod
    def raytrace_scene(
        self,
        ray_origin: Vector3D,
        normalized_ray_direction: Any,
        scene: Scene3D,
    ) -> RGBColor:
        """Colour of every ray ``(ray_origin, normalized_ray_direction)`` in ``scene``."""

    @abstractmethod
    def get_ray_directions(self, camera: Camera) -> Any:
        """One unit direction per pixel, row-major from the top-left pixel."""

    @abstractmethod
    def save_image(self, color: RGBColor, camera: Camera, output_path: Path) -> None:
        """Quantise ``color`` to 8 bits per channel and write it to ``output_path``."""


class Shader(ABC):
    """Port of a per-shape material (reference ``application.py:35-40``)."""

    @abstractmethod
    def create(self, *args: Any, **kwargs: Any) -> Vector3D:
        ...


def render_image_pipeline(scene: Scene3D, output_path: Path, render_service: Renderer) -> None:
    """Render ``scene`` to ``output_path`` (reference ``application.py:43-52``).

    The intermediate values only ever flow from the renderer back into the same
    renderer, which is what lets ``CudaRenderer`` hand over light-weight handles
    (a camera descriptor, a device-backed colour) instead of materialised arrays.
    """
    directions = render_service.get_ray_directions(scene.camera)
    color = render_service.raytrace_scene(scene.camera.position, directions, scene)
    render_service.save_image(color, scene.camera, output_path)
