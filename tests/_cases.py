"""Shared helpers of the test-suite: golden-fixture loading and the parity metric."""

from __future__ import annotations

import json
from pathlib import Path
from types import SimpleNamespace

import numpy as np

from python_ray_tracer_b200.infrastructure.flatten import FlatScene

GOLDEN = Path(__file__).resolve().parent / "golden"

# north_star: "linear float RGB within 1e-4 max-abs"
LINEAR_TOL = 1e-4
# north_star: "the 8-bit PNG within 1 LSB on at least 99.9% of pixels"
PNG_LSB = 1
PNG_FRACTION = 0.999

FULL_FRAME_CASES = [
    "main_160x90", "readme_160x90", "turntable37_160x90", "readme_1x7", "readme_9x1", "readme_37x23",
    "readme_2x2", "c3_small_48_192x108", "c3_small_48_192x108_depth8", "c3_small_48_192x108_depth1",
]
STRESS_CASES = [
    "stress_inside", "stress_overlap", "stress_light_inside", "stress_mirror", "stress_diffuse_only", "stress_domes",
    "stress_scales", "stress_signed_gains", "stress_checkers", "stress_random20", "stress_dark", "stress_far_light",
]
FULL_FRAME_CASES += [f"{name}_48x27" for name in STRESS_CASES]
SUBSET_CASES = [
    "g0_main_960x540_subset", "c1_readme_960x540_subset", "c2_readme_1080p_subset", "c4_readme_8k_subset",
    "c3_1024_4k_subset_depth8",
]


def manifest() -> dict:
    return json.loads((GOLDEN / "manifest.json").read_text())


def load_case(name: str) -> SimpleNamespace:
    """A golden case: the flattened scene the reference rendered, and what it returned."""
    z = np.load(GOLDEN / f"{name}.npz")
    flat = FlatScene(
        spheres=np.ascontiguousarray(z["spheres"]),
        point_lights=np.ascontiguousarray(z["point_lights"]),
        eye=tuple(float(v) for v in z["eye"]),
        dome_intensity=[float(v) for v in z["dome_intensity"]],
        dome_color=tuple(float(v) for v in z["dome_color"]),
    )
    return SimpleNamespace(
        name=name, flat=flat, width=int(z["width"]), height=int(z["height"]), rgb=z["rgb"],
        traced=int(z["traced"]), deepest=int(z["deepest"]), max_depth=int(z["max_depth"]),
        pixels=z["pixels"] if "pixels" in z.files else None,
    )


def parity_report(got: np.ndarray, want: np.ndarray) -> dict:
    """The three readings of "1e-4 max-abs" on unclipped linear values:
    absolute where |want| <= 1, relative above 1, and absolute after clip to [0, 1]."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, (got.shape, want.shape)
    if got.size == 0:
        return dict(max_abs_low=0.0, max_rel_high=0.0, max_abs_clipped=0.0, max_abs=0.0)
    diff = np.abs(got - want)
    low = np.abs(want) <= 1.0
    return dict(
        max_abs_low=float(diff[low].max()) if low.any() else 0.0,
        max_rel_high=float((diff[~low] / np.abs(want[~low])).max()) if (~low).any() else 0.0,
        max_abs_clipped=float(np.abs(np.clip(got, 0, 1) - np.clip(want, 0, 1)).max()),
        max_abs=float(diff.max()),
    )


def assert_linear_parity(got: np.ndarray, want: np.ndarray, tol: float = LINEAR_TOL, what: str = "") -> dict:
    rep = parity_report(got, want)
    assert rep["max_abs_low"] <= tol, (what, rep)
    assert rep["max_rel_high"] <= tol, (what, rep)
    assert rep["max_abs_clipped"] <= tol, (what, rep)
    return rep


def assert_png_parity(got8: np.ndarray, want8: np.ndarray, what: str = "") -> float:
    """All three channels of a pixel within 1 LSB on >= 99.9 % of pixels."""
    got8 = np.asarray(got8).reshape(-1, 3).astype(np.int16)
    want8 = np.asarray(want8).reshape(-1, 3).astype(np.int16)
    assert got8.shape == want8.shape, (got8.shape, want8.shape)
    ok = (np.abs(got8 - want8) <= PNG_LSB).all(axis=1)
    frac = float(ok.mean()) if ok.size else 1.0
    assert frac >= PNG_FRACTION, (what, frac)
    return frac
