"""The launches tools/sanitize.sh puts under compute-sanitizer: every kernel family once, small frames."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure import CudaRenderer, NumpyVector3D  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi  # noqa: E402

W, H = 160, 90
small = scenes.readme_scene(W, H)
mid = scenes.random_spheres_scene(W, H, n_spheres=48, seed=1234)
n = 0
for precision in ("fast", "exact"):
    for scene, accel in ((small, False), (small, True), (mid, False), (mid, True)):
        r = CudaRenderer(**scenes.README_GAINS, precision=precision, accel=accel, max_depth=6, collect_stats=True)
        flat = r.flatten(scene)
        for flags in (0, abi.RT_FLAG_GENERIC_KERNEL, abi.RT_FLAG_WIDE_STORES | abi.RT_FLAG_RGB8):
            c = r.render_rows(flat, scene.camera, 0, H, want_rgba8=True, flags=flags)
            assert np.isfinite(c.numpy()).all()
            n += 1
        d = r.get_ray_directions(scene.camera)
        c = r.trace_rays(flat, scene.camera.position, NumpyVector3D(d.x, d.y, d.z))
        assert np.isfinite(c.numpy()).all()
        n += 1
print("launch groups under the sanitizer:", n)
