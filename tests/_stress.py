"""Seeded stress scenes for the parity suites: configurations the README scenes never reach
(camera inside a sphere, overlapping shapes, a light inside a shape or under the ground,
zero / negative / large gains, roughness 0, more than eight dome lights, tiny and huge radii).
Each returns a ``FlatScene`` (the tables of include/rt_abi.h) plus the depth cap to use;
``tests/golden/make_stress.py`` renders them with the unmodified reference."""

from __future__ import annotations

import numpy as np

from python_ray_tracer_b200.infrastructure.flatten import FlatScene

WIDTH, HEIGHT = 48, 27
GROUND = (0.0, -99999.5, 0.0, 99999.0)


def _table(rows) -> np.ndarray:
    """rows: (cx, cy, cz, r, diffuse, specular, roughness, iridescence, (r, g, b) or None=checker)."""
    t = np.zeros((12, len(rows)))
    for i, (cx, cy, cz, r, dg, sg, rough, ig, colour) in enumerate(rows):
        t[0:4, i] = (cx, cy, cz, r)
        t[4:8, i] = (dg, sg, rough, ig)
        t[8:11, i] = (1.0, 1.0, 1.0) if colour is None else colour
        t[11, i] = 1.0 if colour is None else 0.0
    return np.ascontiguousarray(t)


def _flat(rows, lights, eye=(0.0, 0.2, -2.0), domes=(0.1,), dome_color=(1.0, 1.0, 1.0)) -> FlatScene:
    domes = [float(v) for v in domes]
    if len(domes) > 8:  # what flatten_scene does with more than RT_MAX_DOME_LIGHTS dome lights
        domes = domes[:7] + [float(np.sum(domes[7:]))]
    return FlatScene(_table(rows), np.ascontiguousarray(np.asarray(lights, dtype=np.float64).reshape(-1, 3)),
                     tuple(float(v) for v in eye), [float(v) for v in domes], tuple(float(v) for v in dome_color))


def stress_case(index: int):
    """(name, FlatScene, max_depth) of stress case ``index`` in ``range(N_STRESS)``."""
    rng = np.random.default_rng(9000 + index)
    ground = (*GROUND, 1.0, 1.0, 0.5, 0.05, None)
    light = [(5.0, 10.0, -10.0)]
    if index == 0:  # camera inside a big sphere: every ray hits, chains never end by a miss
        rows = [(0.0, 0.2, -2.0, 5.0, 0.8, 0.9, 0.4, 0.2, (0.9, 0.4, 0.2)), (0.5, 0.3, 0.5, 0.6, 1.0, 1.0, 0.5, 0.05, (0.2, 0.9, 0.4))]
        return "stress_inside", _flat(rows, [(0.5, 1.0, -1.0)]), 6
    if index == 1:  # intersecting spheres with different materials
        rows = [(0.0, 0.3, 2.0, 0.8, 1.0, 1.0, 0.3, 0.0, (1, 0, 0)), (0.6, 0.4, 2.2, 0.7, 0.5, 0.2, 0.9, 0.5, (0, 0, 1)),
                (-0.5, 0.0, 1.6, 0.5, 0.0, 1.0, 0.05, 0.0, (1, 1, 1)), ground]
        return "stress_overlap", _flat(rows, light), 0
    if index == 2:  # a sphere behind the camera, the light inside another one
        rows = [(0.0, 0.2, -6.0, 2.0, 1.0, 1.0, 0.5, 0.05, (1, 1, 0)), (1.0, 1.0, 3.0, 1.2, 1.0, 1.0, 0.5, 0.05, (0, 1, 1)), ground]
        return "stress_light_inside", _flat(rows, [(1.0, 1.2, 3.1)]), 0
    if index == 3:  # mirror-like: roughness 0 and 0.01, specular gain 2 (bounce weight 1.0)
        rows = [(0.6, 0.4, 2.5, 0.9, 0.2, 2.0, 0.0, 0.0, (0.8, 0.8, 0.8)), (-0.7, 0.2, 1.8, 0.7, 0.2, 2.0, 0.01, 0.0, (0.8, 0.2, 0.2)),
                (*GROUND, 1.0, 2.0, 0.3, 0.0, None)]
        return "stress_mirror", _flat(rows, light), 12
    if index == 4:  # diffuse only, no dome light
        rows = [(0.4, 0.5, 3.0, 1.0, 1.0, 0.0, 0.5, 0.0, (1, 0, 1)), (-0.5, 0.1, 1.0, 0.4, 0.7, 0.0, 0.5, 0.0, (0.5, 0.5, 0.5)),
                (*GROUND, 1.0, 0.0, 0.5, 0.0, None)]
        return "stress_diffuse_only", _flat(rows, light, domes=()), 0
    if index == 5:  # ten dome lights (more than RT_MAX_DOME_LIGHTS), coloured; colour of the last one counts
        rows = [(0.55, 0.5, 3.0, 1.0, 1.0, 1.0, 0.5, 0.05, (1, 0, 1)), ground]
        return "stress_domes", _flat(rows, light, domes=tuple(0.01 * (k + 1) for k in range(10)), dome_color=(0.3, 0.5, 1.0)), 0
    if index == 6:  # tiny spheres near the camera, a huge one far away
        rows = [(0.02 * k - 0.05, 0.2 + 0.01 * k, -1.9, 1e-3 * (k + 1), 1.0, 1.0, 0.5, 0.05, tuple(rng.uniform(0.2, 1, 3))) for k in range(5)]
        rows += [(0.0, 0.0, 2.0e6, 1.0e6, 1.0, 0.5, 0.5, 0.05, (0.3, 0.6, 0.9)), ground]
        return "stress_scales", _flat(rows, light), 0
    if index == 7:  # negative specular gain (negative bounce weight), gains above 1
        rows = [(0.55, 0.5, 3.0, 1.0, 1.5, -0.5, 0.5, 0.3, (1, 0.5, 0)), (-0.45, 0.1, 1.0, 0.4, 2.0, 1.2, 0.7, 1.0, (0.2, 0.5, 0.9)), ground]
        return "stress_signed_gains", _flat(rows, light), 10
    if index == 8:  # checker on every shape, strong iridescence
        rows = [(0.55, 0.5, 3.0, 1.0, 1.0, 1.0, 0.5, 1.0, None), (-0.45, 0.1, 1.0, 0.4, 1.0, 0.6, 0.2, 1.0, None), (*GROUND, 1.0, 1.0, 0.5, 1.0, None)]
        return "stress_checkers", _flat(rows, light), 0
    if index == 9:  # twenty random spheres, uncapped
        rows = []
        for _ in range(20):
            rad = rng.uniform(0.1, 0.5)
            rows.append((rng.uniform(-3, 3), rad - 0.5, rng.uniform(0, 8), rad, rng.uniform(0, 1), rng.uniform(0, 1), rng.uniform(0.05, 1),
                         rng.uniform(0, 0.3), tuple(rng.uniform(0.1, 1, 3))))
        rows.append(ground)
        return "stress_random20", _flat(rows, [(5.0, 10.0, -10.0), (-5.0, 10.0, 10.0)]), 0
    if index == 10:  # the light is under the ground: everything is shadowed
        rows = [(0.55, 0.5, 3.0, 1.0, 1.0, 1.0, 0.5, 0.05, (1, 0, 1)), ground]
        return "stress_dark", _flat(rows, [(0.0, -5.0, 0.0)]), 0
    if index == 11:  # camera off axis and high, light very far, three point lights (the first shades)
        rows = [(0.55, 0.5, 3.0, 1.0, 1.0, 1.0, 0.5, 0.05, (1, 0, 1)), (-0.45, 0.1, 1.0, 0.4, 1.0, 1.0, 0.5, 0.05, (0.5, 0.5, 0.5)), ground]
        return "stress_far_light", _flat(rows, [(1.0e6, 2.0e6, -1.5e6), (0, 5, 0), (3, 3, 3)], eye=(1.5, 2.5, -4.0)), 0
    raise IndexError(index)


N_STRESS = 12
STRESS_NAMES = [stress_case(i)[0] for i in range(N_STRESS)]
