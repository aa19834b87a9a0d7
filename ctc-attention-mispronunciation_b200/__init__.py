"""B200-native CTC hot path for dyustc/CTC-Attention-Mispronunciation (drop-in host side).

Import as ``mddctc`` (see mddctc.py at the repository root) or with
``importlib.import_module("ctc-attention-mispronunciation_b200")``.
"""
from . import _lib  # noqa: F401
