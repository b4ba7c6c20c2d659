"""Balanced Ring: graphs (six upstream file names + time_throughput.png)."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.utils.viz import (ensure_directory_exists, plot_accuracy_distribution_per_epoch_global,  # noqa: F401,E402
                                 plot_loss_distribution_by_worker, plot_loss_distribution_per_epoch,
                                 plot_loss_distribution_per_epoch_global, plot_metrics_global, plot_metrics_total,
                                 plot_time_throughput)
