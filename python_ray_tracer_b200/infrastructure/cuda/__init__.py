"""CUDA backend: the ``Renderer`` port over ``librt_b200.so`` (``csrc/``)."""
