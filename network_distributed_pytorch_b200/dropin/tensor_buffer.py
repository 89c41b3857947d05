"""Top-level ``tensor_buffer`` module for unmodified upstream scripts (``import tensor_buffer as tb``,
ddp_powersgd_guide_cifar10/reducer.py:4).  Put this directory first on sys.path."""
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
if _root not in _sys.path:
    _sys.path.insert(0, _root)

from network_distributed_pytorch_b200.tensor_buffer import TensorBuffer, all_gather, all_reduce  # noqa: F401,E402
