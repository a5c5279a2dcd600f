"""Regenerate only the two at-size subsets at the sizes SURVEY.md 8d names, with the UNMODIFIED
reference (build container only): c4 = README scene at 7680x4320, rows {0, H/2, H-1} (every 15th
pixel) + 65 536 random pixels; c3 = 1025 shapes at 3840x2160, depth 8, 4 096 random pixels.
Same code path as make_golden.py (its helpers are imported); takes a few minutes.

    python tests/golden/make_at_size.py
"""
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
import make_golden as mg  # noqa: E402


def main() -> None:
    if not mg.ref.available():
        raise SystemExit("reference tree not mounted; fixtures can only be made in the build container")
    w, h = 7680, 4320
    flat = mg.flatten_scene(mg.scenes.readme_scene(w, h), mg.README_DEFAULT)
    rng = np.random.default_rng(13)
    rows = np.concatenate([np.arange(w) + r * w for r in (0, h // 2, h - 1)])
    pixels = np.unique(np.concatenate([rows[:: max(1, w // 512)], rng.choice(w * h, 65536, replace=False)]))
    _, rgb, counters = mg.reference_render(flat, w, h, pixels=pixels)
    mg.save_case("c4_readme_8k_subset", flat, w, h, rgb, counters, pixels=pixels)
    print("c4:", pixels.size, "pixels")
    flat = mg.flatten_scene(mg.scenes.random_spheres_scene(3840, 2160), mg.README_DEFAULT)
    pixels = np.sort(np.random.default_rng(4321).choice(3840 * 2160, 4096, replace=False))
    _, rgb, counters = mg.reference_render(flat, 3840, 2160, pixels=pixels, max_depth=8)
    mg.save_case("c3_1024_4k_subset_depth8", flat, 3840, 2160, rgb, counters, pixels=pixels, max_depth=8)
    print("c3:", pixels.size, "pixels")


if __name__ == "__main__":
    main()
