"""moss-tts_b200: B200-native (sm_100a) MOSS-TTS synthesis hot path.

Python host mirroring the reference's method surface; all arithmetic in libmosstts_sm100
(include/mosstts.h).  Importing the package does not need a GPU; constructing a model does.
"""
from .config import LMConfig, CodecConfig, CodecStage          # noqa: F401
from . import _lib                                               # noqa: F401

__all__ = ["LMConfig", "CodecConfig", "CodecStage"]
