"""Names the README imports from ``ray_tracer.infrastructure`` (``README.md:40-48``),
with ``CudaRenderer`` in the place of ``NumpyRenderer``.  ``CudaRenderer`` is resolved
lazily so that describing a scene does not load the CUDA library."""

from .materials import NumpyShader, Texture, TextureChecker
from .shapes import CheckeredSphere, NumpySphere
from .values import NumpyRGBColor, NumpyVector3D, NumpyVectorArray3D

FARAWAY = 1.0e39  # base.py:12

__all__ = [
    "CheckeredSphere",
    "CudaRenderer",
    "FARAWAY",
    "NumpyRGBColor",
    "NumpyShader",
    "NumpySphere",
    "NumpyVector3D",
    "NumpyVectorArray3D",
    "Texture",
    "TextureChecker",
]


def __getattr__(name):
    if name == "CudaRenderer":
        from .cuda.renderer import CudaRenderer

        return CudaRenderer
    raise AttributeError(name)
