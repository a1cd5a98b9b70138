"""Replacement for the reference's mast3r_slam/backend/__init__.py (which JIT-compiles
gn.cpp / gn_kernels.cu / matching_kernels.cu, backend/__init__.py:9-29): copy this
file over it so that ``from mast3r_slam.backend import mast3r_slam_backends``
(matching.py:5) resolves to the B200 implementation.  See INTEGRATION.md."""
import mast3r_slam_backends  # noqa: F401  (the frodo-slam_b200 drop-in on sys.path)
