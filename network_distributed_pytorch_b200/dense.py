"""Uncompressed comparison point: the dense gradient averaging of the upstream plain-DDP script
(``average_gradients``, ddp_guide_cifar10/ddp_init.py:57-62: one ``all_reduce`` + divide per parameter), done as ONE
flat all-reduce: gather all gradients into a single buffer (TensorBuffer's gather kernel), one SUM all-reduce, divide,
scatter.  Same ``reduce(grad_in, grad_out, memory_out)`` signature as PowerSGDReducer so the two can be swapped in a
training loop; ``memory_out`` is left untouched (no compression error)."""
import torch

from .tensor_buffer import TensorBuffer


class DenseReducer:
    def __init__(self, random_seed=None, device=None):
        self.device = device
        if torch.distributed.is_available():
            self.n_workers = torch.distributed.get_world_size()
            self.rank = torch.distributed.get_rank()
        else:
            self.n_workers, self.rank = 1, 0
        self._tb = None
        self._key = None

    def reduce(self, grad_in, grad_out, memory_out=None):
        """grad_out[i] = mean over workers of grad_in[i]; returns the communicated bits."""
        key = tuple(tuple(t.shape) for t in grad_in)
        if self._tb is None or self._key != key:
            self._tb = TensorBuffer(list(grad_in))  # packs on construction
            self._key = key
        else:
            self._tb.pack(list(grad_in))
        if self.n_workers > 1:
            torch.distributed.all_reduce(self._tb.buffer)
            self._tb.buffer /= self.n_workers
        self._tb.unpack(list(grad_out))
        return self._tb.bits()
