"""forest_slam_b200 — B200-native (sm_100a) matching + pose hot path of Forest-SLAM behind the call shapes
of ros_ws/src/mono_slam.py and ros_ws/src/stereo_slam.py.  See DESIGN.md / INTEGRATION.md."""
from . import _lib  # noqa: F401
from ._lib import ForestSlamError, LIB_PATH  # noqa: F401

__all__ = ["_lib", "ForestSlamError", "LIB_PATH"]


def __getattr__(name):
    # the operator layer needs torch; import it lazily so that `import forest_slam_b200` stays cheap
    if name == "api":
        import importlib
        return importlib.import_module(".api", __name__)
    raise AttributeError(name)
