"""NumPy restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module, as the checker or as the timed CPU
baseline.  The product never imports it.

It evaluates the same whole-array algorithm as the reference's ``NumpyRenderer``
(``ray_tracer/infrastructure/numpy/base.py:91-151``, ``shape.py:28-51``,
``shader.py:63-320``): one full-size ``intersect`` pass per shape, a mask +
``np.extract`` compaction per shape with hits, ``create`` on the compacted rays
including a recursive trace of the mirror rays, ``np.place`` back.  Every ufunc is
applied in the reference's order, so on the same NumPy build the result is
bit-identical to the reference's (pinned by ``tests/golden/make_golden.py`` in the
build container; the reference itself cannot travel to the GPU box).  It works on the
flattened scene (``rt_abi.h`` layout) instead of the reference's object graph; that
is the only structural difference, and it is why its timing stands in for the
reference's in ``bench.py`` (kind "port").
"""

from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

FARAWAY = 1.0e39  # base.py:12

Vec = Tuple[object, object, object]  # three scalars or three flat float64 arrays


def _dot(a: Vec, b: Vec):
    return (a[0] * b[0]) + (a[1] * b[1]) + (a[2] * b[2])  # base.py:34-35


def _sub(a: Vec, b: Vec) -> Vec:
    return (a[0] - b[0], a[1] - b[1], a[2] - b[2])


def _add(a: Vec, b: Vec) -> Vec:
    return (a[0] + b[0], a[1] + b[1], a[2] + b[2])


def _mul(a: Vec, k) -> Vec:
    return (a[0] * k, a[1] * k, a[2] * k)


def _norm(a: Vec) -> Vec:
    mag = np.sqrt(_dot(a, a))  # base.py:61-64
    return _mul(a, 1.0 / np.where(mag == 0, 1, mag))


def _pick(cond, v: Vec) -> Vec:
    return tuple(c if np.isscalar(c) else np.extract(cond, c) for c in v)  # base.py:15-25,66-71


def intersect(o: Vec, d: Vec, c: Vec, r: float):
    """``NumpySphere.intersect`` (shape.py:28-51)."""
    b = 2 * _dot(d, _sub(o, c))
    cc = _dot(c, c) + _dot(o, o) - 2 * _dot(c, o) - (r * r)
    disc = (b**2) - (4 * cc)
    sq = np.sqrt(np.maximum(0, disc))
    h0 = (-b - sq) / 2
    h1 = (-b + sq) / 2
    h = np.where((h0 > 0) & (h0 < h1), h0, h1)
    return np.where((disc > 0) & (h > 0), h, FARAWAY)


class NumpyPort:
    """Holds one flattened scene and traces ray batches through it."""

    def __init__(self, flat, max_depth: int = 0, all_point_lights: bool = False) -> None:
        f = flat.spheres
        self.n = f.shape[1]
        self.centres = [(float(f[0, s]), float(f[1, s]), float(f[2, s])) for s in range(self.n)]
        self.radius = [float(f[3, s]) for s in range(self.n)]
        self.diffuse_gain = [float(f[4, s]) for s in range(self.n)]
        self.specular_gain = [float(f[5, s]) for s in range(self.n)]
        self.roughness = [float(f[6, s]) for s in range(self.n)]
        self.irid_gain = [float(f[7, s]) for s in range(self.n)]
        self.colour = [(float(f[8, s]), float(f[9, s]), float(f[10, s])) for s in range(self.n)]
        self.checker = [bool(f[11, s] != 0.0) for s in range(self.n)]
        self.lights = [tuple(float(v) for v in row) for row in flat.point_lights]
        self.eye = tuple(float(v) for v in flat.eye)
        self.dome_intensity = list(flat.dome_intensity)
        self.dome_color = tuple(flat.dome_color)
        self.max_depth = int(max_depth or 0)
        self.all_point_lights = bool(all_point_lights)
        self.traced = 0
        self.hits = 0
        self.deepest = 0

    # -- base.py:91-121 ---------------------------------------------------------------
    def trace(self, o: Vec, d: Vec, generation: int = 0) -> Vec:
        if self.max_depth > 0 and generation >= self.max_depth:
            return (0, 0, 0)
        self.traced += int(np.size(d[0]))
        self.deepest = max(self.deepest, generation + 1)
        distances = [intersect(o, d, self.centres[s], self.radius[s]) for s in range(self.n)]
        nearest = distances[0]
        for dist in distances[1:]:
            nearest = np.minimum(nearest, dist)
        color: Vec = (0, 0, 0)
        for s, dist in enumerate(distances):
            hit = (nearest != FARAWAY) & (dist == nearest)
            if np.any(hit):
                self.hits += int(np.count_nonzero(hit))
                shaded = self._create(s, _pick(hit, o), _pick(hit, d), np.extract(hit, dist), generation)
                placed = []
                for comp in shaded:
                    plane = np.zeros(hit.shape)
                    np.place(plane, hit, comp)
                    placed.append(plane)
                color = _add(color, tuple(placed))
        return color

    # -- shader.py:114-128 ------------------------------------------------------------
    def _lit(self, s: int, nudged: Vec, to_light: Vec):
        dist = [intersect(nudged, to_light, self.centres[j], self.radius[j]) for j in range(self.n)]
        nearest = dist[0]
        for u in dist[1:]:
            nearest = np.minimum(nearest, u)
        return dist[s] == nearest

    # -- shader.py:13-32 --------------------------------------------------------------
    def _texture(self, s: int, m: Vec) -> Vec:
        if self.checker[s]:
            on = ((m[0] * 2).astype(int) % 2) == ((m[2] * 2).astype(int) % 2)
            return (1 * on, 1 * on, 1 * on)
        return self.colour[s]

    # -- shader.py:246-320 ------------------------------------------------------------
    def _specular(self, s: int, n: Vec, to_light: Vec, to_eye: Vec):
        eps = 1e-8
        gain = self.specular_gain[s]
        L = _norm(to_light)
        V = _norm(to_eye)
        H = _norm(_add(L, V))
        ndv = np.clip(_dot(n, V), 0, 1)
        ndh = np.clip(_dot(n, H), 0, 1)
        vdh = np.clip(_dot(V, H), 0, 1)
        ndl = np.clip(_dot(n, L), 0, 1)
        f0 = ((1.5 - 1) / (1.5 + 1)) ** 2
        fres = f0 + (1 - f0) * (1 - vdh) ** 5
        alpha = self.roughness[s] ** 2
        den = ndh**2 * (alpha**2 - 1) + 1
        dist = (alpha**2) / (np.pi * (den**2 + eps))

        def g1(x: Vec):
            q = np.clip(_dot(n, x), 0, 1)
            return 2 * q / (q + np.sqrt(alpha**2 + (1 - alpha**2) * (q**2)) + eps)

        geo = g1(L) * g1(V)
        base = (fres * dist * geo) / (4 * ndv + eps)
        glint = (1 - ndv) ** 2.5
        glint = glint * ndl
        total = base + gain * glint
        return np.where(ndv <= 0, 0, total)

    # -- shader.py:186-232 ------------------------------------------------------------
    def _iridescence(self, s: int, n: Vec, to_eye: Vec) -> Vec:
        weight, thick, film_ior = 0.1, 0.3, 1.4
        va = np.clip(_dot(n, to_eye), 0.0, 1.0)
        af = np.abs(va - 0.5) * 2.0
        phase = af * np.pi * thick * 10.0
        ip = np.sin(phase)
        hs = (film_ior - 1.0) / 2.0
        r = ip * hs + (1.0 - hs) * (1.0 - ip)
        g = ip * (1.0 - hs) + hs * (1.0 - ip)
        b = 0.5 + 0.5 * ip
        gain = self.irid_gain[s]
        return (r * weight * gain, g * weight * gain, b * weight * gain)

    # -- shader.py:63-112 -------------------------------------------------------------
    def _create(self, s: int, o: Vec, d: Vec, dist, generation: int) -> Vec:
        m = _add(o, _mul(d, dist))
        n = _mul(_sub(m, self.centres[s]), 1.0 / self.radius[s])
        to_eye = _norm(_sub(self.eye, m))
        nudged = _add(m, _mul(n, 0.0001))
        tex = self._texture(s, m)
        lights = self.lights if self.all_point_lights else self.lights[:1]

        color: Vec = (0.004, 0.004, 0.004)
        spec_sum = 0.0
        lit0 = None
        for li, lp in enumerate(lights):
            to_light = _norm(_sub(lp, m))
            lit = self._lit(s, nudged, to_light)
            if li == 0:
                lit0 = lit
            lambert = np.maximum(_dot(n, to_light), 0)
            color = _add(color, _mul(_mul(_mul(tex, lambert), lit), self.diffuse_gain[s]))
            spec = self._specular(s, n, to_light, to_eye)
            spec_sum = spec if len(lights) == 1 else spec_sum + spec * lit

        rd = _norm(_sub(d, _mul(_mul(n, 2), _dot(d, n))))
        refl = self.trace(nudged, rd, generation + 1)
        refl = tuple(np.array(c) for c in refl)

        ny = np.maximum(_dot(n, (0, 1, 0)), 0)
        dome = 0.0
        for inten in self.dome_intensity:
            dome = dome + inten * ny
        color = _add(color, _mul(self.dome_color, dome))

        gain = self.specular_gain[s]
        if len(lights) == 1:
            color = _add(color, tuple((spec_sum + c * 0.5) * gain * lit0 for c in refl))
        else:
            color = _add(color, tuple((spec_sum + c * 0.5 * lit0) * gain for c in refl))
        return _add(color, self._iridescence(s, n, to_eye))


def ray_directions(cam_position, width: int, height: int, pixels: Optional[np.ndarray] = None) -> Vec:
    """``NumpyRenderer.get_ray_directions`` (base.py:123-141), optionally on a subset."""
    aspect = float(width) / height
    x = np.tile(np.linspace(-1, 1, width), height)
    y = np.repeat(np.linspace(1 / aspect + 0.25, -1 / aspect + 0.25, height), width)
    if pixels is not None:
        x, y = x[pixels], y[pixels]
    return _norm((x - cam_position[0], y - cam_position[1], 0 - cam_position[2]))


def render(flat, width: int, height: int, pixels: Optional[np.ndarray] = None, max_depth: int = 0,
           all_point_lights: bool = False):
    """Linear RGB as float64 ``[3, n]`` plus ``(traced, hits, deepest)``."""
    port = NumpyPort(flat, max_depth=max_depth, all_point_lights=all_point_lights)
    d = ray_directions(flat.eye, width, height, pixels)
    color = port.trace(tuple(flat.eye), d)
    n = int(np.size(d[0]))
    rgb = np.stack([np.broadcast_to(np.asarray(c, dtype=np.float64), (n,)) for c in color])
    return rgb, (port.traced, port.hits, port.deepest)


def quantize(rgb: np.ndarray) -> np.ndarray:
    """``save_image`` arithmetic (base.py:143-151): ``[3, n]`` -> ``[n, 3]`` uint8."""
    return (255 * np.clip(rgb, 0, 1)).astype(np.uint8).T.copy()
