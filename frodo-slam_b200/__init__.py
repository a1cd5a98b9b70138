"""frodo-slam_b200: B200-native (sm_100a) implementation of the
``mast3r_slam_backends`` operator surface of jiaming-ai/frodo-slam
(iter_proj, refine_matches, gauss_newton_rays/points/calib).

Layout: ``csrc/`` hand-written CUDA + the C-ABI (``libfrodo_b200.so``),
``_lib.py`` ctypes loader, ``backends.py`` the reference-facing operator
surface, ``dist.py`` edge-sharded multi-GPU Gauss-Newton, ``synth.py`` seeded
synthetic inputs.  There is no CPU fallback: ops raise if the CUDA library is
missing.
"""
__version__ = "0.1.0"
