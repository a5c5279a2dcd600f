"""Host-side value types: 3-vectors and colours that may hold scalars or flat arrays.

Mirrors the public surface of the reference's ``NumpyVector3D`` family
(``ray_tracer/infrastructure/numpy/base.py:15-87``): the same class names, the same
``x``/``y``/``z`` attributes and the same operator set with the same meaning
(``abs(v)`` is the SQUARED length, ``norm`` multiplies by a zero-safe reciprocal of
the length).  They describe scenes and carry results across the boundary; the
per-ray arithmetic of the renderer itself runs in the CUDA kernels (``csrc/``),
never through these operators.
"""

from __future__ import annotations

from typing import Any, Tuple

import numpy as np

from ..domain import Vector3D


def _is_triple(value: Any) -> bool:
    return all(hasattr(value, axis) for axis in ("x", "y", "z"))


class NumpyVector3D(Vector3D):
    """Structure-of-arrays 3-vector (reference ``base.py:28-79``)."""

    __slots__ = ()

    def components(self) -> Tuple[Any, Any, Any]:
        return self.x, self.y, self.z

    def dot(self, other: "NumpyVector3D") -> Any:
        # base.py:34-35: three products summed left to right.
        return (self.x * other.x) + (self.y * other.y) + (self.z * other.z)

    def __abs__(self) -> Any:
        # base.py:37-38: squared length, not the length.
        return self.dot(self)

    def _zip(self, other: Any, op) -> "NumpyVector3D":
        if _is_triple(other):
            return NumpyVector3D(op(self.x, other.x), op(self.y, other.y), op(self.z, other.z))
        return NumpyVector3D(op(self.x, other), op(self.y, other), op(self.z, other))

    def __mul__(self, other: Any) -> "NumpyVector3D":
        return self._zip(other, lambda a, b: a * b)

    def __add__(self, other: Any) -> "NumpyVector3D":
        return self._zip(other, lambda a, b: a + b)

    def __sub__(self, other: Any) -> "NumpyVector3D":
        return self._zip(other, lambda a, b: a - b)

    def __truediv__(self, other: Any) -> "NumpyVector3D":
        return self._zip(other, lambda a, b: a / b)

    def __neg__(self) -> "NumpyVector3D":
        return NumpyVector3D(-self.x, -self.y, -self.z)

    def norm(self) -> "NumpyVector3D":
        # base.py:61-64: v * (1 / |v|) with |v| == 0 replaced by 1.
        length = np.sqrt(self.dot(self))
        return self * (1.0 / np.where(length == 0, 1, length))

    def extract(self, cond: Any) -> "NumpyVector3D":
        # base.py:66-71: keep the entries where cond holds; scalars pass through.
        if np.isscalar(cond):
            return self
        picked = [c if np.isscalar(c) else np.extract(cond, c) for c in self.components()]
        return NumpyVectorArray3D(*picked)

    def place(self, cond: np.ndarray) -> "NumpyVector3D":
        # base.py:73-79: scatter back into zero-filled arrays shaped like cond.
        planes = []
        for comp in self.components():
            plane = np.zeros(cond.shape)
            np.place(plane, cond, comp)
            planes.append(plane)
        return NumpyVector3D(*planes)


class NumpyVectorArray3D(NumpyVector3D):
    """Marker subclass for compacted ray batches (reference ``base.py:82-83``)."""

    __slots__ = ()


class NumpyRGBColor(NumpyVector3D):
    """Colour with channels in ``x, y, z`` (reference ``base.py:86-87``)."""

    __slots__ = ()


def as_float_triple(value: Any, what: str) -> Tuple[float, float, float]:
    """``(x, y, z)`` of a scalar-valued vector as Python floats, or ``ValueError``."""
    if not _is_triple(value):
        raise ValueError(f"{what}: expected an object with x, y, z, got {type(value).__name__}")
    try:
        return float(value.x), float(value.y), float(value.z)
    except (TypeError, ValueError) as exc:
        raise ValueError(f"{what}: components must be scalars") from exc
