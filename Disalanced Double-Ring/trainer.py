"""Disalanced Double-Ring: training loop API (reference-signature wrappers over ldnn_b200.trainer)."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.compat import make_train_global, train_local_epoch  # noqa: F401,E402

TOPOLOGY, BALANCE = "double_ring", "disbalanced"
train_global = make_train_global(TOPOLOGY, BALANCE)
