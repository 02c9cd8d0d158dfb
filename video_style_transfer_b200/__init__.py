"""B200-native (sm_100a) hot path of tanmud/video_style_transfer: the AnimateDiff-XL motion
module and the UnZipLoRA content/style attention, behind the reference's own plugin surface.

Importing the package loads libvst_b200.so; there is no CPU or eager-PyTorch fallback.
"""
from . import _lib  # noqa: F401  (fails loudly when the CUDA library is missing)
from . import ops  # noqa: F401

__version__ = "0.1.0"
