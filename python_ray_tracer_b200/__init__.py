"""B200-native backend for tristanGIANDO/python_ray_tracer scenes.

One hot path -- ``NumpyRenderer.raytrace_scene`` and everything it calls -- rebuilt as
hand-written sm_100a CUDA behind the reference's ``Renderer`` port: ``CudaRenderer``
plugs into ``render_image_pipeline(scene, path, renderer)`` where ``NumpyRenderer`` did.
"""

from .application import Renderer, Shader, render_image_pipeline
from .domain import Camera, DomeLight, PointLight, RGBColor, Scene3D, Shape, Vector3D

__all__ = [
    "Renderer",
    "Shader",
    "render_image_pipeline",
    "Camera",
    "DomeLight",
    "PointLight",
    "RGBColor",
    "Scene3D",
    "Shape",
    "Vector3D",
]
__version__ = "0.1.0"
