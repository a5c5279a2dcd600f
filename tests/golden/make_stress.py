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


if __name__ == "__main__":
    main()
