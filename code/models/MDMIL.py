"""Plugin file selected by `Model.name: MDMIL` (code/models/model_interface.py:1257-1274 imports
`models.MDMIL` and takes the class of the same name; reference file code/models/MDMIL.py:60-113).
Same B200-native path as TransMIL with the fixed 1024-d stem and the `_fc2` classifier."""
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from transmil_b200 import MDMIL, PPEG, TransLayer  # noqa: E402,F401

__all__ = ["MDMIL", "TransLayer", "PPEG"]
