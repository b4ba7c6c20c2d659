"""Disbalanced All-Reduce: rank-0 test evaluation with macro/weighted/micro P/R/F1."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.compat import evaluate  # noqa: F401,E402
