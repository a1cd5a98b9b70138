"""Drop-in replacement for the reference's pybind extension ``mast3r_slam_backends``
(mast3r_slam/backend/src/gn.cpp:152-160; imported as a top-level module by
global_opt.py:11 and local_mapping.py:9).  Backed by frodo-slam_b200 (sm_100a CUDA
behind a C ABI); there is no CPU fallback."""
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _root not in sys.path:
    sys.path.insert(0, _root)

from frodo_slam_b200.backends import (  # noqa: E402,F401
    gauss_newton_calib,
    gauss_newton_points,
    gauss_newton_rays,
    gauss_newton_rays_odom,
    iter_proj,
    refine_matches,
)
