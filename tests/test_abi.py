"""The C-ABI library loads and exports every symbol include/forestslam_b200.h declares (no compute calls)."""
import ctypes
import os
import re

from tests.util import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "forestslam_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fs_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from forest_slam_b200 import _lib
    names = _declared()
    assert len(names) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_table_matches_header():
    from forest_slam_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()
    lib = _lib.load()
    assert lib.fs_version() >= 100
    assert lib.fs_status_string(0) == b"FS_OK"
    assert lib.fs_status_string(-2) == b"FS_ERR_WORKSPACE (workspace too small)"


def test_no_oracle_in_product():
    """The product package never imports the oracle or cv2 (the oracle is test infrastructure)."""
    pkg = os.path.join(ROOT, "forest_slam_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py") and fn != "selfcheck.py":
            txt = open(os.path.join(pkg, fn)).read()
            assert "import cv2" not in txt and "from oracle" not in txt and "import oracle" not in txt, fn


def test_missing_library_fails_loudly(monkeypatch):
    from forest_slam_b200 import _lib
    import pytest
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libforestslam_b200.so")
    with pytest.raises(_lib.ForestSlamError):
        _lib.load()


def test_survey_8b_names_exist():
    """the drop-in module exports the named functions with the named keyword arguments."""
    import inspect
    from forest_slam_b200 import api
    want = {"match_mutual_nn": {"threshold", "ratio"}, "bf_match_hamming": {"cross_check", "ratio"},
            "find_essential_mat": {"focal", "pp", "prob", "threshold", "max_iters"},
            "recover_pose": {"focal", "pp", "distance_thresh"},
            "solve_pnp_ransac": {"reprojection_error", "confidence", "iterations_count"},
            "backproject_stereo": {"baseline", "zmin", "zmax"}}
    for name, kws in want.items():
        fn = getattr(api, name)
        assert callable(fn), name
        assert kws <= set(inspect.signature(fn).parameters), (name, kws - set(inspect.signature(fn).parameters))
