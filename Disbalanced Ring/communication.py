"""Disbalanced Ring: exchange API.  Host-level functions run over torch.distributed;
the training entrypoint uses the fused P2P kernels (ldnn_b200/parallel/fused.py) on GPU."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.compat import ring_all_reduce_equal, ring_all_reduce_weighted  # noqa: F401,E402
