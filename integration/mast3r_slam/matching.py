"""Replacement for the reference's mast3r_slam/matching.py: copy this file over it so that
``from mast3r_slam.matching import match`` (tracker.py, global_opt.py) runs the whole matching
pipeline -- prologue, iter_proj, occlusion test, fp16 casts, refine_matches, epilogue -- as one
C-ABI call into libfrodo_b200.  Same signatures as matching.py:8-95.  See INTEGRATION.md."""
from mast3r_slam.config import config
from frodo_slam_b200 import matching as _m

pixel_to_lin = _m.pixel_to_lin
lin_to_pixel = _m.lin_to_pixel
prep_for_iter_proj = _m.prep_for_iter_proj


def match(X11, X21, D11, D21, idx_1_to_2_init=None):
    return _m.match(X11, X21, D11, D21, idx_1_to_2_init, cfg=config["matching"])


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None):
    return _m.match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init, cfg=config["matching"])
