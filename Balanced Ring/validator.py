"""Balanced Ring: validation pass."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.compat import validate  # noqa: F401,E402
