"""Plugin file selected by `Model.name: TransMIL` in the DeepGraft/*.yaml configs
(code/models/model_interface.py:1257-1274 imports `models.TransMIL` and takes the
class of the same name).  The implementation is the B200-native package."""
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from transmil_b200 import PPEG, TransLayer, TransMIL  # noqa: E402,F401

__all__ = ["TransMIL", "TransLayer", "PPEG"]
