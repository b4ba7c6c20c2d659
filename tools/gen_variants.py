"""Writes the six variant directories (thin, variant-pinned entrypoints).

The upstream project *is* six copy-pasted directories; here they are generated shims over
one shared package, so `cd "Balanced Ring" && python main.py …` keeps working.
Run:  python tools/gen_variants.py
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VARIANTS = {
    "Balanced All-Reduce": ("allreduce", "balanced"),
    "Balanced Ring": ("ring", "balanced"),
    "Balanced Double-Ring": ("double_ring", "balanced"),
    "Disbalanced All-Reduce": ("allreduce", "disbalanced"),
    "Disbalanced Ring": ("ring", "disbalanced"),
    "Disalanced Double-Ring": ("double_ring", "disbalanced"),   # upstream spelling kept
}
BOOT = '''import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
'''
COMM = {
    "allreduce": "average_gradients_weighted, average_weights_weighted, average_gradients_equal, average_weights_equal",
    "ring": "ring_all_reduce_equal, ring_all_reduce_weighted",
    "double_ring": "double_ring_all_reduce, double_ring_all_reduce_weighted",
}


def files(name, topo, bal):
    tag = f'TOPOLOGY, BALANCE = "{topo}", "{bal}"\n'
    yield "main.py", f'''"""{name}: entrypoint.  Same flags as upstream's main.py plus the new ones
(see ldnn_b200/config.py); run under torchrun / torch.distributed.launch, one process per GPU."""
{BOOT}from ldnn_b200.cli import main as _main

{tag}

def main(argv=None):
    return _main(argv, topology=TOPOLOGY, balance=BALANCE)


if __name__ == "__main__":
    main()
'''
    yield "trainer.py", f'''"""{name}: training loop API (reference-signature wrappers over ldnn_b200.trainer)."""
{BOOT}from ldnn_b200.compat import make_train_global, train_local_epoch  # noqa: F401,E402

{tag}train_global = make_train_global(TOPOLOGY, BALANCE)
'''
    yield "communication.py", f'''"""{name}: exchange API.  Host-level functions run over torch.distributed;
the training entrypoint uses the fused P2P kernels (ldnn_b200/parallel/fused.py) on GPU."""
{BOOT}from ldnn_b200.compat import {COMM[topo]}  # noqa: F401,E402
'''
    extra = "" if bal == "balanced" else ", fixed_ratio=0.5"
    yield "dataloader.py", f'''"""{name}: data API ({bal} sharding pinned)."""
{BOOT}from functools import partial as _partial

from ldnn_b200 import compat as _c  # noqa: E402

get_loaders = _partial(_c.get_loaders, balance="{bal}")
get_dataset_partition = _c.get_dataset_partition
get_next_dataset_partition = _c.get_next_dataset_partition
get_subset_loaders = _c.get_subset_loaders
estimate_epoch_duration = _c.estimate_epoch_duration
'''
    yield "model.py", f'''"""{name}: network definitions."""
{BOOT}from ldnn_b200.models import EnhancedCNNModel, ResBlock, ResNet50  # noqa: F401,E402
'''
    yield "validator.py", f'''"""{name}: validation pass."""
{BOOT}from ldnn_b200.compat import validate  # noqa: F401,E402
'''
    yield "evaluator.py", f'''"""{name}: rank-0 test evaluation with macro/weighted/micro P/R/F1."""
{BOOT}from ldnn_b200.compat import evaluate  # noqa: F401,E402
'''
    yield "vizualizator.py", f'''"""{name}: graphs (six upstream file names + time_throughput.png)."""
{BOOT}from ldnn_b200.utils.viz import (ensure_directory_exists, plot_accuracy_distribution_per_epoch_global,  # noqa: F401,E402
                                 plot_loss_distribution_by_worker, plot_loss_distribution_per_epoch,
                                 plot_loss_distribution_per_epoch_global, plot_metrics_global, plot_metrics_total,
                                 plot_time_throughput)
'''


for name, (topo, bal) in VARIANTS.items():
    d = os.path.join(ROOT, name)
    os.makedirs(d, exist_ok=True)
    for fn, body in files(name, topo, bal):
        with open(os.path.join(d, fn), "w") as f:
            f.write(body)
print("wrote", len(VARIANTS), "variant directories")
