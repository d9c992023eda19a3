"""finufft_type1 / finufft_type2 with the pytorch_finufft.functional signature (reference call sites
efgpnd.py:1496-1499, 1533-1549, 1679), served by oracle/fast_nufft.py."""
import os
import sys

_ORACLE = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ORACLE not in sys.path:
    sys.path.append(_ORACLE)
from fast_nufft import finufft_type1, finufft_type2  # noqa: E402,F401
