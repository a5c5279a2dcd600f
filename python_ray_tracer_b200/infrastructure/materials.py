"""Material description: textures and the per-shape shader parameter block.

Mirrors the constructor surface of the reference's ``Texture``, ``TextureChecker``
and ``NumpyShader`` (``ray_tracer/infrastructure/numpy/shader.py:13-54``).  Here they
are parameter holders only: the flattener (``flatten.py``) copies their numbers into
the sphere SoA and the shading itself (``shader.py:63-320``) runs in the CUDA kernel.
"""

from __future__ import annotations

from typing import Any, Optional

from ..application import Shader
from .values import NumpyRGBColor

# Row values of RT_F_TEXTURE in include/rt_abi.h.
TEXTURE_CONSTANT = 0.0
TEXTURE_CHECKER = 1.0


class Texture:
    """Constant colour (reference ``shader.py:13-19``); white when no colour is given."""

    kind = TEXTURE_CONSTANT

    def __init__(self, color: Optional[Any] = None) -> None:
        self.color = NumpyRGBColor(1, 1, 1) if color is None else color


class TextureChecker(Texture):
    """World-space XZ checker, always white/black (reference ``shader.py:22-32``).

    The colour argument is accepted and ignored, as the reference's ``get_color`` does.
    """

    kind = TEXTURE_CHECKER


class NumpyShader(Shader):
    """The five gains plus the diffuse texture of one shape (``shader.py:36-54``).

    ``reflection_gain`` is stored and never read by the reference (the reflected colour
    is weighted by ``0.5 * specular_gain``, ``shader.py:106``); it is kept for API
    compatibility and does not reach the device.
    """

    def __init__(
        self,
        reflection_gain: float,
        specular_gain: float,
        specular_roughness: float,
        iridescence_gain: float,
        diffuse_gain: float,
        diffuse_color: Texture,
    ) -> None:
        self.reflection_gain = reflection_gain
        self.specular_gain = specular_gain
        self.specular_roughness = specular_roughness
        self.iridescence_gain = iridescence_gain
        self.diffuse_gain = diffuse_gain
        self.diffuse_color = diffuse_color

    def create(self, *args: Any, **kwargs: Any):
        raise NotImplementedError(
            "NumpyShader.create is evaluated inside the CUDA trace kernel; "
            "render through CudaRenderer.raytrace_scene"
        )
