"""Import alias: ``import mddctc`` == the package directory ``ctc-attention-mispronunciation_b200`` (whose name is
not a Python identifier)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("ctc-attention-mispronunciation_b200")
for _sub in ("_lib", "_rt", "loss", "lm", "decoder", "mdd", "corpus", "prefix", "dist"):
    try:
        _m = importlib.import_module("ctc-attention-mispronunciation_b200." + _sub)
        sys.modules[__name__ + "." + _sub] = _m
        setattr(_pkg, _sub, _m)
    except ModuleNotFoundError as e:        # optional submodules not written yet
        if _sub not in str(e):
            raise
sys.modules[__name__] = _pkg
