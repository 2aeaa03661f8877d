"""B200-native volume-rendering device for Open Volume Renderer (the hot path only: ray marching / delta tracking of a
structured-regular scalar volume through a 1-D transfer function, behind the ovr::MainRenderer plugin API).

The directory name carries a hyphen (the reference repo's name); import it as `ovr_b200` (alias module at the repo
root) or with importlib.import_module("open-volume-renderer_b200").
"""
from . import _capi as capi  # noqa: F401
from . import scenes  # noqa: F401
from .renderer import FrameBufferData, MainRenderer, create_renderer  # noqa: F401
from .scenes import Camera, Scene, TransferFunction, Volume  # noqa: F401

__all__ = ["capi", "scenes", "MainRenderer", "FrameBufferData", "create_renderer", "Camera", "Scene", "TransferFunction",
           "Volume"]
from . import scene_io  # noqa: F401,E402
from . import distributed  # noqa: F401,E402
