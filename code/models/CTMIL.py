"""Plugin file selected by `Model.name: CTMIL` (code/models/model_interface.py:1257-1274 imports `models.CTMIL` and takes the
class of the same name; reference file code/models/CTMIL.py).  The Nystrom path behind it is the B200-native one of TransMIL
(tokens in, TM_FC1_TOKENS); the conv stem in front of it is stock torch modules."""
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from transmil_b200 import CTMIL, PPEG, TransLayer  # noqa: E402,F401

__all__ = ["CTMIL", "TransLayer", "PPEG"]
