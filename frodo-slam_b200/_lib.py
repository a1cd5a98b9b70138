"""ctypes loader for libfrodo_b200.so (the C ABI declared in include/frodo_b200.h).

The library is the only compute path of this package: if it is missing or fails
to load, importing the operators raises -- there is no CPU or PyTorch fallback.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfrodo_b200.so")

c_int, c_float, c_void_p, c_size_t = ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_size_t


class GnParams(ctypes.Structure):
    """struct frodo_gn_params (include/frodo_b200.h)."""
    _fields_ = [("mode", c_int), ("sigma_a", c_float), ("sigma_b", c_float),
                ("C_thresh", c_float), ("Q_thresh", c_float), ("height", c_int),
                ("width", c_int), ("pixel_border", c_int), ("z_eps", c_float)]


GN_RAYS, GN_POINTS, GN_CALIB = 0, 1, 2


class MatchParams(ctypes.Structure):
    """struct frodo_match_params (include/frodo_b200.h)."""
    _fields_ = [("max_iter", c_int), ("lambda_init", c_float), ("convergence_thresh", c_float),
                ("dist_thresh", c_float), ("radius", c_int), ("dilation_max", c_int),
                ("ft_matching_threshold", c_float)]

class TrackParams(ctypes.Structure):
    """struct frodo_track_params (include/frodo_b200.h)."""
    _fields_ = [("mode", c_int), ("sigma_a", c_float), ("sigma_b", c_float), ("huber", c_float),
                ("max_iters", c_int), ("rel_error", c_float), ("delta_norm", c_float),
                ("height", c_int), ("width", c_int), ("pixel_border", c_int), ("z_eps", c_float),
                ("raw_calib", c_int)]


class GnPeer(ctypes.Structure):
    """struct frodo_gn_peer (include/frodo_b200.h)."""
    _fields_ = [("peer_buffers", c_void_p), ("local_buffer", c_void_p), ("world", c_int),
                ("rank", c_int), ("max_edges", c_int), ("epoch0", ctypes.c_uint32),
                ("arrivals0", ctypes.c_uint32 * 2)]


class OdomTerms(ctypes.Structure):
    """struct frodo_gn_odom_terms (include/frodo_b200.h)."""
    _fields_ = [("odom_ii_row", c_void_p), ("odom_jj_row", c_void_p), ("odom_rank", c_void_p),
                ("Tij", c_void_p), ("tij_stride", c_int), ("E_odom", c_int), ("s_bar", c_void_p),
                ("sigma_odom_t", c_float), ("sigma_odom_r", c_float),
                ("sigma_scale_prior", c_float), ("blocks", c_void_p), ("blocks_bytes", c_size_t)]


# name -> (restype, argtypes); kept in sync with include/frodo_b200.h (tests check this)
SIGNATURES = {
    "frodo_version": (c_int, []),
    "frodo_last_error": (ctypes.c_char_p, []),
    "frodo_iter_proj": (c_int, [c_void_p] * 5 + [c_int] * 5 + [c_float, c_float, c_void_p]),
    "frodo_refine_matches_f16": (c_int, [c_void_p] * 5 + [c_int] * 7 + [c_void_p]),
    "frodo_set_refine_patch": (c_int, [c_int]),
    "frodo_refine_matches_f64": (c_int, [c_void_p] * 5 + [c_int] * 7 + [c_void_p]),
    "frodo_refine_matches_f32": (c_int, [c_void_p] * 5 + [c_int] * 7 + [c_void_p]),
    "frodo_match_prep": (c_int, [c_void_p] * 6 + [c_int] * 3 + [c_void_p]),
    "frodo_match_occlusion": (c_int, [c_void_p] * 4 + [c_float, c_void_p, c_void_p] + [c_int] * 3 +
                              [c_void_p]),
    "frodo_convert_f32_f16": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "frodo_constrain_points_to_ray": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                              c_void_p]),
    "frodo_match_finish": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_float, c_void_p,
                                   c_void_p] + [c_int] * 3 + [c_void_p]),
    "frodo_match_workspace_bytes": (c_size_t, [c_int] * 4),
    "frodo_match": (c_int, [ctypes.POINTER(MatchParams)] + [c_void_p] * 7 + [c_int] * 4 +
                    [c_void_p, c_size_t, c_void_p]),
    "frodo_match_quality": (c_int, [c_void_p] * 6 + [c_float, c_float] + [c_void_p] * 4 +
                            [c_int] * 3 + [c_void_p]),
    "frodo_gn_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "frodo_gn_chol_stamps_offset": (c_size_t, [c_int]),
    "frodo_gn_workspace_init": (c_int, [c_void_p, c_size_t, c_int, c_int, c_int, c_void_p]),
    "frodo_gn_edge_index": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    "frodo_gn_pack_points": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "frodo_gn_workspace_packed_offset": (c_size_t, [c_int, c_int, c_int]),
    "frodo_gn_packed_bytes": (c_size_t, [c_int, c_int]),
    "frodo_gn_accumulate": (c_int, [ctypes.POINTER(GnParams)] + [c_void_p] * 10 + [c_int] * 4 +
                            [c_void_p] * 4 + [c_size_t, c_void_p]),
    "frodo_gn_solve_retract": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float,
                                       c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "frodo_gauss_newton": (c_int, [ctypes.POINTER(GnParams)] + [c_void_p] * 9 + [c_int] * 5 +
                           [c_float, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "frodo_gn_odom_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "frodo_gauss_newton_rays_odom": (c_int, [ctypes.POINTER(GnParams)] + [c_void_p] * 11 +
                                     [c_int, c_int, c_void_p, c_float, c_float, c_float] +
                                     [c_int] * 5 + [c_float, c_void_p, c_void_p, c_void_p,
                                                    c_size_t, c_void_p]),
    "frodo_gn_pack_slot": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "frodo_gn_arena_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "frodo_gauss_newton_arena": (c_int, [ctypes.POINTER(GnParams)] + [c_void_p] * 13 + [c_int] * 6 +
                                 [c_float, ctypes.POINTER(OdomTerms), c_void_p, c_void_p, c_void_p,
                                  c_size_t, c_void_p]),
    "frodo_gn_odom_blocks_bytes": (c_size_t, [c_int]),
    "frodo_gn_add_odom": (c_int, [ctypes.POINTER(OdomTerms), c_void_p, c_void_p, c_int, c_int,
                                  c_void_p, c_void_p, c_void_p]),
    "frodo_gn_peer_buffer_bytes": (c_size_t, [c_int]),
    "frodo_gauss_newton_sharded": (c_int, [ctypes.POINTER(GnParams)] + [c_void_p] * 6 +
                                   [c_int] * 3 + [c_void_p] * 3 + [c_int] * 4 + [c_float] +
                                   [ctypes.POINTER(GnPeer), ctypes.POINTER(OdomTerms)] +
                                   [c_void_p] * 4 + [c_size_t, c_void_p]),
    "frodo_gn_peer_allreduce": (c_int, [c_void_p, c_void_p, c_size_t, c_int, c_int, c_int,
                                        ctypes.c_uint32, c_void_p, c_size_t, c_void_p, c_void_p]),
    "frodo_track_workspace_bytes": (c_size_t, [c_int]),
    "frodo_track_pose_static_max_points": (c_int, []),
    "frodo_track_pose_static": (c_int, [ctypes.POINTER(TrackParams)] + [c_void_p] * 10 + [c_int] +
                                [c_void_p] * 5),
    "frodo_track_pose": (c_int, [ctypes.POINTER(TrackParams)] + [c_void_p] * 10 + [c_int] +
                         [c_void_p] * 5 + [c_size_t, c_void_p]),
}

_lib = None


class FrodoError(RuntimeError):
    pass


def load():
    """Load the CUDA library (once).  Raises FrodoError if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FrodoError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        msg = load().frodo_last_error().decode("utf-8", "replace")
        raise FrodoError(f"libfrodo_b200 error {rc}: {msg}")
