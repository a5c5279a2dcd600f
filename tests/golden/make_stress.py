"""Renders the stress scenes of tests/_stress.py with the UNMODIFIED reference and writes
tests/golden/stress_*.npz (same layout as make_golden.py's cases).  Build container only:

    python tests/golden/make_stress.py
"""

from __future__ import annotations

import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parents[1]))
sys.path.insert(0, str(HERE.parent))

from make_golden import reference_render, save_case  # noqa: E402
from _stress import HEIGHT, N_STRESS, WIDTH, stress_case  # noqa: E402

from oracle import reference_bridge as ref  # noqa: E402


def main() -> None:
    if not ref.available():
        raise SystemExit("reference tree not mounted; fixtures can only be made in the build container")
    for i in range(N_STRESS):
        name, flat, max_depth = stress_case(i)
        print(name)
        _, rgb, counters = reference_render(flat, WIDTH, HEIGHT, max_depth=max_depth)
        save_case(f"{name}_{WIDTH}x{HEIGHT}", flat, WIDTH, HEIGHT, rgb, counters, max_depth=max_depth)
    tie_case()


def tie_case():
    """A shape listed twice (known deviation): the reference lets EVERY shape at the nearest
    distance contribute and recurse (base.py:102-119), so a duplicated sphere is shaded twice; the
    kernels shade the lowest index once.  The fixture holds the reference's render of the scene as
    given (`rgb`) and of the scene with the duplicate removed (`rgb_single`)."""
    import numpy as np

    from _stress import GROUND, _flat

    ball = (-0.45, 0.1, 1.0, 0.4, 1.0, 1.0, 0.5, 0.05, (0.5, 0.5, 0.5))
    rest = [(0.55, 0.5, 3.0, 1.0, 1.0, 1.0, 0.5, 0.05, (1, 0, 1)), (*GROUND, 1.0, 1.0, 0.5, 0.05, None)]
    light = [(5.0, 10.0, -10.0)]
    twice, once = _flat([ball, ball] + rest, light), _flat([ball] + rest, light)
    _, rgb, counters = reference_render(twice, WIDTH, HEIGHT, max_depth=4)
    _, rgb_single, _ = reference_render(once, WIDTH, HEIGHT, max_depth=4)
    save_case(f"tie_duplicate_{WIDTH}x{HEIGHT}", twice, WIDTH, HEIGHT, rgb, counters, max_depth=4)
    path = HERE / f"tie_duplicate_{WIDTH}x{HEIGHT}.npz"
    z = dict(np.load(path))
    z["rgb_single"] = rgb_single
    np.savez_compressed(path, **z)
    print("tie_duplicate: pixels where the reference adds the duplicate:", int((np.abs(rgb - rgb_single) > 1e-12).any(axis=0).sum()))


if __name__ == "__main__":
    if "--tie-only" in sys.argv:
        tie_case()
        raise SystemExit(0)
    main()
