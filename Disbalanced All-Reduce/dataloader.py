"""Disbalanced All-Reduce: data API (disbalanced sharding pinned)."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from functools import partial as _partial

from ldnn_b200 import compat as _c  # noqa: E402

get_loaders = _partial(_c.get_loaders, balance="disbalanced")
get_dataset_partition = _c.get_dataset_partition
get_next_dataset_partition = _c.get_next_dataset_partition
get_subset_loaders = _c.get_subset_loaders
estimate_epoch_duration = _c.estimate_epoch_duration
