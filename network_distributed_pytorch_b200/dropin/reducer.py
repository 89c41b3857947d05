"""Top-level ``reducer`` module for unmodified upstream scripts (``from reducer import
PowerSGDReducer``, ddp_powersgd_guide_cifar10/ddp_init.py:20).  Put this directory first on sys.path."""
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
if _root not in _sys.path:
    _sys.path.insert(0, _root)

from network_distributed_pytorch_b200.reducer import *  # noqa: F401,F403,E402
from network_distributed_pytorch_b200.reducer import PowerSGDReducer, Reducer, all_reduce, n_bits, orthogonalize  # noqa: F401,E402
