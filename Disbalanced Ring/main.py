"""Disbalanced Ring: entrypoint.  Same flags as upstream's main.py plus the new ones
(see ldnn_b200/config.py); run under torchrun / torch.distributed.launch, one process per GPU."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from ldnn_b200.cli import main as _main

TOPOLOGY, BALANCE = "ring", "disbalanced"


def main(argv=None):
    return _main(argv, topology=TOPOLOGY, balance=BALANCE)


if __name__ == "__main__":
    main()
