import sys, ctypes as C
from pathlib import Path; R = str(Path(__file__).resolve().parents[1]); sys.path.insert(0, R); sys.path.insert(0, R + "/tests")
import numpy as np
from _cases import load_case
from test_host_logic import _parse_accel
from python_ray_tracer_b200.infrastructure import CudaRenderer, NumpyVector3D
from python_ray_tracer_b200.infrastructure.cuda import abi
from python_ray_tracer_b200 import Camera

case = load_case("c3_small_48_192x108_depth1")
cam = Camera(NumpyVector3D(*case.flat.eye), case.width, case.height)
for prec in ("exact", "fast"):
    a = CudaRenderer(precision=prec, accel=False, max_depth=1, collect_stats=True).render_rows(case.flat, cam, 0, case.height)
    b = CudaRenderer(precision=prec, accel=True, max_depth=1, collect_stats=True).render_rows(case.flat, cam, 0, case.height)
    A, B = a.numpy(), b.numpy()
    bad = np.flatnonzero((A != B).any(axis=0))
    print(prec, "differing pixels", bad.size, "of", A.shape[1], a.counters(), b.counters())
    print(bad[:20], [(int(i) // case.width, int(i) % case.width) for i in bad[:10]])
    for i in bad[:5]:
        print(i, A[:, i], B[:, i])
