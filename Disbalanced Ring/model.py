"""Disbalanced Ring: network definitions."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.models import EnhancedCNNModel, ResBlock, ResNet50  # noqa: F401,E402
