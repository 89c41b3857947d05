"""B200-native PowerSGD data-parallel reducer: drop-in for the ``reducer.py`` /
``tensor_buffer.py`` pair of jaeyong-song/network_distributed_pytorch.

    from network_distributed_pytorch_b200.reducer import PowerSGDReducer
    from network_distributed_pytorch_b200.tensor_buffer import TensorBuffer

or, to leave the upstream training scripts untouched, put
``network_distributed_pytorch_b200/dropin`` first on ``sys.path`` so that ``from reducer import
PowerSGDReducer`` (ddp_init.py:20) resolves here.
"""
from .reducer import PowerSGDReducer, Reducer, all_reduce, n_bits, orthogonalize  # noqa: F401
from .tensor_buffer import TensorBuffer  # noqa: F401

__all__ = ["PowerSGDReducer", "Reducer", "TensorBuffer", "orthogonalize", "all_reduce", "n_bits"]
