"""Writes the committed golden fixtures by running the UNMODIFIED reference.

Run in the build container (where ``/root/reference`` is mounted):

    python tests/golden/make_golden.py [--heavy]

Everything here goes through ``oracle/reference_bridge.py``: the reference's own
``NumpyRenderer`` / ``NumpySphere`` / ``NumpyShader`` classes are imported from
``/root/reference`` and driven through their public API.  The fixtures are the inputs
(flattened scene tables, pixel subsets) together with the reference's outputs, so the
tests on the GPU box -- which has no reference tree -- compare against what the
reference actually computed.  ``--heavy`` also regenerates the 1025-sphere subset
(minutes of NumPy time); without it that file is left as committed.
"""

from __future__ import annotations

import argparse
import hashlib
import json
import sys
import time
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parents[1]
sys.path.insert(0, str(ROOT))

from oracle import reference_bridge as ref  # noqa: E402
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402
from python_ray_tracer_b200.infrastructure.png import read_png, write_png  # noqa: E402

README_DEFAULT = DefaultMaterial(**scenes.README_GAINS)


def flat_fields(flat) -> dict:
    return dict(
        spheres=flat.spheres,
        point_lights=flat.point_lights,
        eye=np.asarray(flat.eye),
        dome_intensity=np.asarray(flat.dome_intensity, dtype=np.float64),
        dome_color=np.asarray(flat.dome_color),
    )


def save_case(name: str, flat, width: int, height: int, rgb: np.ndarray, counters, pixels=None,
              max_depth: int = 0, **extra) -> None:
    fields = flat_fields(flat)
    fields.update(
        width=np.int64(width), height=np.int64(height), rgb=rgb, traced=np.int64(counters[0]),
        deepest=np.int64(counters[1]), max_depth=np.int64(max_depth), **extra,
    )
    if pixels is not None:
        fields["pixels"] = np.asarray(pixels, dtype=np.int64)
    np.savez_compressed(HERE / f"{name}.npz", **fields)
    print(f"  {name}.npz: {rgb.shape[1]} rays, traced {counters[0]}, deepest {counters[1]}")


def reference_render(flat, width, height, pixels=None, max_depth=0):
    scene = ref.build_scene(flat, width, height)
    t0 = time.perf_counter()
    rgb, counters = ref.render(scene, pixels=pixels, max_depth=max_depth)
    print(f"  reference: {width}x{height} n={rgb.shape[1]} in {time.perf_counter() - t0:.2f}s")
    return scene, rgb, counters


def main() -> None:
    parser = argparse.ArgumentParser()
    parser.add_argument("--heavy", action="store_true")
    args = parser.parse_args()
    if not ref.available():
        raise SystemExit("reference tree not mounted; fixtures can only be made in the build container")
    m = ref.modules()
    manifest = {}

    # G0 -- main.py scene, 960x540: the reference's committed render.png pins this one.
    print("G0 main.py scene 960x540")
    flat = flatten_scene(scenes.main_scene(960, 540))
    scene, rgb, counters = reference_render(flat, 960, 540)
    ref.save_png(scene, rgb, HERE / "_tmp.png")
    ours = read_png(HERE / "_tmp.png")
    committed = read_png(ref.REFERENCE_ROOT / "render.png")
    (HERE / "_tmp.png").unlink()
    assert np.array_equal(ours, committed), "HEAD reference no longer reproduces its own render.png"
    manifest["g0_main_960x540_rgb8_sha256"] = hashlib.sha256(ours.tobytes()).hexdigest()
    manifest["g0_main_960x540_traced"] = int(counters[0])
    write_png(HERE / "g0_main_960x540.png", ours, 9)  # the reference's 8-bit output (decoded pixels == render.png)
    # linear values on a fixed subset of that frame
    pixels = np.sort(np.random.default_rng(11).choice(960 * 540, 8192, replace=False))
    save_case("g0_main_960x540_subset", flat, 960, 540, rgb[:, pixels], counters, pixels=pixels)

    # C1 -- README scene 960x540 (README surface -> HEAD API mapping of SURVEY.md 8b)
    print("C1 README scene 960x540")
    flat = flatten_scene(scenes.readme_scene(960, 540), README_DEFAULT)
    flat_head = flatten_scene(scenes.readme_scene_head_api(960, 540))
    assert np.array_equal(flat.spheres, flat_head.spheres)
    scene, rgb, counters = reference_render(flat, 960, 540)
    ref.save_png(scene, rgb, HERE / "_tmp.png")
    write_png(HERE / "c1_readme_960x540.png", read_png(HERE / "_tmp.png"), 9)
    (HERE / "_tmp.png").unlink()
    manifest["c1_readme_960x540_traced"] = int(counters[0])
    manifest["c1_readme_960x540_deepest"] = int(counters[1])
    manifest["c1_readme_960x540_max"] = float(rgb.max())
    pixels = np.sort(np.random.default_rng(12).choice(960 * 540, 8192, replace=False))
    save_case("c1_readme_960x540_subset", flat, 960, 540, rgb[:, pixels], counters, pixels=pixels)

    # small full frames (cheap enough for every CPU test run)
    print("small full frames")
    for name, builder, default in (
        ("main_160x90", scenes.main_scene, None),
        ("readme_160x90", scenes.readme_scene, README_DEFAULT),
    ):
        flat = flatten_scene(builder(160, 90), default)
        _, rgb, counters = reference_render(flat, 160, 90)
        save_case(name, flat, 160, 90, rgb, counters)
    flat = flatten_scene(scenes.turntable_scene(37, 256, 160, 90), README_DEFAULT)
    _, rgb, counters = reference_render(flat, 160, 90)
    save_case("turntable37_160x90", flat, 160, 90, rgb, counters)
    # odd sizes: 1-pixel-wide / 1-pixel-high / non-multiple-of-tile frames (linspace edge cases)
    for w, h in ((1, 7), (9, 1), (37, 23), (2, 2)):
        flat = flatten_scene(scenes.readme_scene(w, h), README_DEFAULT)
        _, rgb, counters = reference_render(flat, w, h)
        save_case(f"readme_{w}x{h}", flat, w, h, rgb, counters)

    # C2 / C4 -- README scene at 1080p and 8K: subsets (rows 0, H/2, H-1 + random pixels)
    for tag, (w, h), k in (("c2_readme_1080p", (1920, 1080), 8192), ("c4_readme_8k", (7680, 4320), 65536)):
        print(tag)
        flat = flatten_scene(scenes.readme_scene(w, h), README_DEFAULT)
        rng = np.random.default_rng(13)
        rows = np.concatenate([np.arange(w) + r * w for r in (0, h // 2, h - 1)])
        pixels = np.unique(np.concatenate([rows[:: max(1, w // 512)], rng.choice(w * h, k, replace=False)]))
        _, rgb, counters = reference_render(flat, w, h, pixels=pixels)
        save_case(f"{tag}_subset", flat, w, h, rgb, counters, pixels=pixels)

    # C3-small -- 48 random spheres + ground, uncapped and capped at 8 generations
    print("C3-small (48 spheres)")
    flat = flatten_scene(scenes.random_spheres_scene(192, 108, n_spheres=48, seed=1234), README_DEFAULT)
    _, rgb, counters = reference_render(flat, 192, 108)
    save_case("c3_small_48_192x108", flat, 192, 108, rgb, counters)
    _, rgb, counters = reference_render(flat, 192, 108, max_depth=8)
    save_case("c3_small_48_192x108_depth8", flat, 192, 108, rgb, counters, max_depth=8)
    _, rgb, counters = reference_render(flat, 192, 108, max_depth=1)
    save_case("c3_small_48_192x108_depth1", flat, 192, 108, rgb, counters, max_depth=1)

    # C3 -- 1024 random spheres + ground at 4K, depth 8, the 4096-pixel subset of SURVEY.md 8d (heavy)
    if args.heavy:
        print("C3 (1025 shapes) 4K subset, depth 8 -- this takes minutes")
        flat = flatten_scene(scenes.random_spheres_scene(3840, 2160), README_DEFAULT)
        pixels = np.sort(np.random.default_rng(4321).choice(3840 * 2160, 4096, replace=False))
        _, rgb, counters = reference_render(flat, 3840, 2160, pixels=pixels, max_depth=8)
        save_case("c3_1024_4k_subset_depth8", flat, 3840, 2160, rgb, counters, pixels=pixels, max_depth=8)

    # known answers for NumpySphere.intersect (shape.py:28-51) and get_ray_directions
    print("intersect / ray-direction vectors")
    V = m.base.NumpyVector3D
    rng = np.random.default_rng(21)
    n = 4096
    o = rng.normal(size=(3, n)) * 3.0
    d = rng.normal(size=(3, n))
    d /= np.sqrt((d * d).sum(axis=0))
    cases = []
    for centre, radius in (((0.0, 0.0, 0.0), 1.0), ((0.55, 0.5, 3.0), 1.0), ((0.0, -99999.5, 0.0), 99999.0)):
        sphere = m.shape.NumpySphere(V(*centre), radius, None)
        t = sphere.intersect(V(o[0], o[1], o[2]), V(d[0], d[1], d[2]))
        cases.append((centre, radius, np.asarray(t, dtype=np.float64)))
    head_on = m.shape.NumpySphere(V(0.0, 0.0, 0.0), 1.0, None).intersect(V(0.0, 0.0, -3.0), V(0.0, 0.0, 1.0))
    np.savez_compressed(
        HERE / "intersect_vectors.npz", origins=o, directions=d,
        centres=np.asarray([c[0] for c in cases]), radii=np.asarray([c[1] for c in cases]),
        distances=np.stack([c[2] for c in cases]), head_on=np.float64(head_on),
    )
    dirs = {}
    for w, h in ((1, 1), (1, 5), (7, 1), (16, 9), (33, 17), (960, 540)):
        cam = m.domain.Camera(V(0.0, 0.2, -2.0), w, h)
        r = m.base.NumpyRenderer().get_ray_directions(cam)
        planes = np.stack([np.broadcast_to(np.asarray(c, dtype=np.float64), (w * h,)) for c in r.components()])
        dirs[f"d_{w}x{h}"] = planes if w * h <= 1024 else planes[:, :: 97]
    np.savez_compressed(HERE / "ray_directions.npz", **dirs)

    (HERE / "manifest.json").write_text(json.dumps(manifest, indent=1, sort_keys=True) + "\n")
    print("manifest:", manifest)


if __name__ == "__main__":
    main()
