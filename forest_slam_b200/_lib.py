"""ctypes binding of the C ABI declared in include/forestslam_b200.h.

There is NO CPU fallback: if the shared library is missing or a symbol cannot be bound, importing the
operators raises.  The library is built in-tree by ``__graft_entry__.build()``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# FORESTSLAM_B200_LIB: another build of the same library (tuning experiments: scratch/build_variant.sh)
LIB_PATH = os.environ.get("FORESTSLAM_B200_LIB") or os.path.join(_HERE, "libforestslam_b200.so")

c_i32p = C.c_void_p
c_f32p = C.c_void_p
c_f64p = C.c_void_p
c_u8p = C.c_void_p
ptr = C.c_void_p


class FSCamera(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("dist", C.c_double * 5)]


# name -> (restype, argtypes); mirrors include/forestslam_b200.h one to one
SIGNATURES = {
    "fs_version": (C.c_int, []),
    "fs_last_error_string": (C.c_char_p, []),
    "fs_status_string": (C.c_char_p, [C.c_int]),
    "fs_device_sm_count": (C.c_int, []),
    "fs_launch_count": (C.c_longlong, []),
    "fs_hamming_match_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_hamming_match": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_int, C.c_float,
                                   ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_hamming_operand_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_hamming_prepare": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr]),
    "fs_hamming_match_tc_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_hamming_match_tc": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_int, C.c_float,
                                      ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_float_prepare": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr]),
    "fs_float_match_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_float_match_tc_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_float_operand_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_float_match_split3_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fs_float_match": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_float, C.c_float,
                                 C.c_float, ptr, ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_float_match_knn": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_float, C.c_float,
                                     C.c_float, ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_float_top3_keys": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_int, ptr, ptr, C.c_size_t, ptr]),
    "fs_float_match_split3": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_float, C.c_float,
                                        C.c_float, ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_float_match_fp32": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_float, C.c_float,
                                      C.c_float, ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_float_scores": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, C.c_float, ptr, ptr]),
    "fs_sinkhorn_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "fs_sinkhorn_match": (C.c_int, [ptr, C.c_int, C.c_int, C.c_int, ptr, ptr, C.c_float, C.c_int, C.c_float,
                                    C.c_int, ptr, ptr, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_gather_matches": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, ptr, ptr, ptr, ptr, ptr,
                                    ptr]),
    "fs_backproject_stereo": (C.c_int, [ptr, ptr, ptr, C.c_int64, C.c_int, C.c_int, C.POINTER(FSCamera),
                                        C.c_double, C.c_double, C.c_double, ptr, ptr, ptr]),
    "fs_backproject_filter": (C.c_int, [ptr, ptr, ptr, ptr, C.c_int, C.c_int, C.POINTER(FSCamera),
                                        C.c_double, C.c_double, C.c_double, ptr, ptr, ptr, ptr]),
    "fs_backproject_filter_bank": (C.c_int, [ptr, ptr, ptr, C.c_int, C.c_int, ptr, C.c_int, C.c_int, C.c_int, ptr, ptr,
                                             C.POINTER(FSCamera), C.c_double, C.c_double, C.c_double, ptr, ptr, ptr,
                                             ptr]),
    "fs_filter_matches_distance": (C.c_int, [ptr, ptr, C.c_int, C.c_int, C.c_int, ptr, ptr]),
    "fs_pack_records": (C.c_int, [ptr, ptr, ptr, ptr, ptr, ptr, ptr, C.c_int, ptr, ptr]),
    "fs_find_essential_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "fs_find_essential_ransac": (C.c_int, [ptr, ptr, ptr, C.c_int, C.c_int, C.c_double, C.c_double,
                                           C.c_double, C.c_double, C.c_double, C.c_int, ptr, ptr, ptr, ptr,
                                           ptr, ptr, C.c_size_t, ptr]),
    "fs_ransac_subsets": (C.c_int, [ptr, C.c_int, C.c_int, ptr, ptr]),
    "fs_score_essential": (C.c_int, [ptr, ptr, ptr, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                     C.c_double, ptr, C.c_int, ptr, ptr]),
    "fs_score_pnp": (C.c_int, [ptr, ptr, ptr, C.c_int, C.c_int, C.POINTER(FSCamera), C.c_double, ptr,
                               C.c_int, ptr, ptr]),
    "fs_recover_pose_workspace_bytes": (C.c_size_t, [C.c_int]),
    "fs_recover_pose": (C.c_int, [ptr, ptr, ptr, ptr, ptr, C.c_int, C.c_int, C.c_double, C.c_double,
                                  C.c_double, C.c_double, ptr, ptr, ptr, ptr, ptr, C.c_size_t, ptr]),
    "fs_recover_pose_static_max_pairs": (C.c_int, []),
    "fs_recover_pose_static": (C.c_int, [ptr, ptr, ptr, ptr, ptr, C.c_int, C.c_int, C.c_double, C.c_double,
                                         C.c_double, C.c_double, ptr, ptr, ptr, ptr, ptr]),
    "fs_solve_pnp_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "fs_solve_pnp_ransac": (C.c_int, [ptr, ptr, ptr, C.c_int, C.c_int, C.POINTER(FSCamera), C.c_double,
                                      C.c_double, C.c_int, C.c_int, ptr, ptr, ptr, ptr, ptr, ptr, ptr, ptr,
                                      C.c_size_t, ptr]),
    "fs_orb_search_for_initialization": (C.c_int, [ptr, ptr, ptr, ptr, C.c_int, C.c_int, ptr, ptr, ptr, C.c_int, ptr,
                                                   C.c_int, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float,
                                                   C.c_float, ptr, ptr, ptr]),
    "fs_orb_stereo_matches": (C.c_int, [ptr, ptr, ptr, ptr, ptr, ptr, ptr, ptr, C.c_int, C.c_int, ptr, ptr, C.c_int,
                                        C.c_int, C.c_float, C.c_float, ptr, ptr, ptr, ptr]),
    "fs_render_forest": (C.c_int, [ptr, C.POINTER(FSCamera), C.c_double, C.c_double, ptr, C.c_int, C.c_int, C.c_int,
                                   ptr, ptr, ptr]),
    "fs_rodrigues": (C.c_int, [ptr, C.c_int, ptr, ptr]),
    "fs_compose_trajectory": (C.c_int, [ptr, ptr, ptr, C.c_int, ptr, ptr, ptr]),
    "fs_transform_points": (C.c_int, [ptr, ptr, C.c_int, C.c_int, ptr, ptr, ptr]),
}

_lib = None


class ForestSlamError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the CUDA extension; raise (never fall back) if it is absent or incomplete."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ForestSlamError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise ForestSlamError(f"{LIB_PATH} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != 0:
        lib = load()
        msg = lib.fs_last_error_string().decode(errors="replace")
        kind = lib.fs_status_string(status).decode()
        raise ForestSlamError(f"{what} failed: {kind} ({status}) {msg}")
