"""Balanced All-Reduce: exchange API.  Host-level functions run over torch.distributed;
the training entrypoint uses the fused P2P kernels (ldnn_b200/parallel/fused.py) on GPU."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.compat import average_gradients_weighted, average_weights_weighted, average_gradients_equal, average_weights_equal  # noqa: F401,E402
