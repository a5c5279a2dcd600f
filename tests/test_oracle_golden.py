"""Pins the oracle (oracle/rt_oracle.c and oracle/numpy_port.py) to what the reference
computed: every committed golden fixture, including the main.py scene whose 8-bit image
is the reference's own render.png."""

import hashlib

import numpy as np
import pytest

from _cases import FULL_FRAME_CASES, GOLDEN, SUBSET_CASES, load_case, manifest
from oracle import numpy_port
from python_ray_tracer_b200.infrastructure.png import read_png

# The restatements follow the reference's operation order; what may differ is libm-grade
# functions (pow, sin) in the last place, amplified by at most the specular peak.
ORACLE_TOL = 1e-11


def _have(name):
    return (GOLDEN / f"{name}.npz").exists()


@pytest.mark.parametrize("name", FULL_FRAME_CASES)
def test_c_oracle_full_frames(c_oracle, name):
    case = load_case(name)
    rgb, stats = c_oracle.render(case.flat, case.width, case.height, max_depth=case.max_depth)
    assert np.abs(rgb - case.rgb).max() <= ORACLE_TOL * max(1.0, np.abs(case.rgb).max())
    assert int(stats[0]) == case.traced          # rays traced over all generations
    assert int(stats[2]) == case.deepest         # deepest generation reached


@pytest.mark.parametrize("name", SUBSET_CASES)
def test_c_oracle_subsets(c_oracle, name):
    if not _have(name):
        pytest.skip(f"{name} not generated")
    case = load_case(name)
    rgb, _ = c_oracle.render(case.flat, case.width, case.height, pixels=case.pixels, max_depth=case.max_depth)
    assert np.abs(rgb - case.rgb).max() <= ORACLE_TOL * max(1.0, np.abs(case.rgb).max())


@pytest.mark.parametrize("name", ["main_160x90", "readme_160x90", "readme_37x23", "readme_1x7", "readme_9x1",
                                  "c3_small_48_192x108_depth8"])
def test_numpy_port_is_bit_identical(name):
    case = load_case(name)
    rgb, (traced, _hits, deepest) = numpy_port.render(case.flat, case.width, case.height, max_depth=case.max_depth)
    assert np.array_equal(rgb, case.rgb)  # same ufuncs, same order, same NumPy build family
    assert traced == case.traced and deepest == case.deepest


def test_golden_g0_render_png(c_oracle):
    """The reference's only committed result: render.png == main.py scene at 960x540."""
    man = manifest()
    want = read_png(GOLDEN / "g0_main_960x540.png")
    assert hashlib.sha256(want.tobytes()).hexdigest() == man["g0_main_960x540_rgb8_sha256"]
    assert man["g0_main_960x540_rgb8_sha256"] == "cfe744300c99933f6815e1ee77978f894a9cb775cd293b0223dbc6fc18745291"
    case = load_case("g0_main_960x540_subset")
    rgb, stats = c_oracle.render(case.flat, 960, 540)
    assert int(stats[0]) == man["g0_main_960x540_traced"]
    got = c_oracle.quantize(rgb).reshape(540, 960, 3)
    # truncating quantisation can flip a byte where 255*v sits on an integer; libm-level
    # differences are ~1e-16, so at most a handful of bytes may differ by one
    diff = np.abs(got.astype(np.int16) - want.astype(np.int16))
    assert diff.max() <= 1 and np.count_nonzero(diff) <= 8
    assert np.abs(rgb[:, case.pixels] - case.rgb).max() <= ORACLE_TOL * 10


def test_golden_c1_readme_png(c_oracle):
    man = manifest()
    case = load_case("c1_readme_960x540_subset")
    rgb, stats = c_oracle.render(case.flat, 960, 540)
    assert int(stats[0]) == man["c1_readme_960x540_traced"] == 1060815  # SURVEY.md 8c pin
    assert int(stats[2]) == man["c1_readme_960x540_deepest"] == 74
    assert abs(rgb.max() - man["c1_readme_960x540_max"]) < 1e-12
    want = read_png(GOLDEN / "c1_readme_960x540.png")
    got = c_oracle.quantize(rgb).reshape(540, 960, 3)
    diff = np.abs(got.astype(np.int16) - want.astype(np.int16))
    assert diff.max() <= 1 and np.count_nonzero(diff) <= 8


def test_intersect_known_answers(c_oracle):
    z = np.load(GOLDEN / "intersect_vectors.npz")
    assert float(z["head_on"]) == 2.0  # reference tests/test_objects.py:6-11 re-expressed
    assert c_oracle.sphere_distance((0, 0, -3), (0, 0, 1), (0, 0, 0), 1.0) == 2.0
    assert c_oracle.sphere_distance((0, 0, -3), (0, 1, 0), (0, 0, 0), 1.0) == 1.0e39  # miss -> FARAWAY
    o, d = z["origins"], z["directions"]
    for centre, radius, want in zip(z["centres"], z["radii"], z["distances"]):
        got = np.array([c_oracle.sphere_distance(o[:, i], d[:, i], centre, float(radius)) for i in range(0, o.shape[1], 8)])
        assert np.array_equal(got, want[::8])  # IEEE basic operations only: bit-exact
        port = numpy_port.intersect(tuple(o), tuple(d), tuple(float(v) for v in centre), float(radius))
        assert np.array_equal(port, want)


def test_ray_directions(c_oracle):
    z = np.load(GOLDEN / "ray_directions.npz")
    for key in z.files:
        w, h = (int(v) for v in key[2:].split("x"))
        got = c_oracle.ray_directions((0.0, 0.2, -2.0), w, h)
        port = np.stack([np.broadcast_to(np.asarray(c, dtype=np.float64), (w * h,))
                         for c in numpy_port.ray_directions((0.0, 0.2, -2.0), w, h)])
        if w * h > 1024:
            got, port = got[:, ::97], port[:, ::97]
        assert np.array_equal(port, z[key])
        assert np.array_equal(got, z[key])


def test_quantize_truncates(c_oracle):
    rgb = np.array([[0.0, 1.0, 2.0, -1.0, 0.999999, 0.5, 1.0 / 255.0]] * 3)
    want = np.array([0, 255, 255, 0, 254, 127, 1], dtype=np.uint8)
    assert np.array_equal(c_oracle.quantize(rgb)[:, 0], want)
    assert np.array_equal(numpy_port.quantize(rgb)[:, 0], want)
