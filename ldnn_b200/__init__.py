"""ldnn_b200 — B200-native data-parallel training with fused peer-to-peer gradient exchange.

Capabilities of Sanasar1/Learning-Deep-Neural-Network-In-Distributed-Computing-Environment
(six variants: All-Reduce / Ring / Double-Ring × Balanced / Disbalanced), re-designed for
Blackwell: PyTorch as the framework layer, hand-written sm_100a CUDA (tcgen05/TMEM/TMA
implicit-GEMM convolutions, fused BN/ReLU/residual, fused head, P2P exchange kernels with the
optimizer in their tail) as the compute path, NCCL only as bootstrap and measured baseline.
"""
__version__ = "0.1.0"

from .config import Config, parse_args  # noqa: F401
