"""CPU checker for the hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import anything under ``oracle/``, as the checker or as
the timed CPU baseline.  The product (``python_ray_tracer_b200``) never does.

Parity pinning: ``rt_oracle.c`` and ``numpy_port.py`` are validated against the
unmodified reference imported from ``/root/reference`` in the build container
(``tests/golden/make_golden.py``), and against the committed fixtures it wrote --
among them the ``main.py`` scene whose 8-bit image equals the reference's own
``render.png``.
"""
