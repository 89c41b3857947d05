"""Host mirror of the upstream ``tensor_buffer.py`` (ddp_powersgd_guide_cifar10/tensor_buffer.py):
same class, methods and module functions; the flatten / unflatten copies are a single
gather / scatter kernel over a device pointer table instead of ``torch.cat`` and a Python loop of
``tensor[:] = entry`` (tensor_buffer.py:19, :27-36).  The flat layout is bit-identical to
``torch.cat([t.view(-1) for t in tensors])``: running offsets, no padding, dtype of the inputs.
"""
import ctypes

import torch


class TensorBuffer():
    """
    Packs multiple tensors into one flat buffer for efficient
    intra-worker communication.
    """

    def __init__(self, tensors):
        tensors = list(tensors)
        if len(tensors) == 0:
            # upstream: torch.cat([]) (tensor_buffer.py:19)
            raise RuntimeError("torch.cat(): expected a non-empty list of Tensors")
        indices = [0]
        for tensor in tensors:
            indices.append(indices[-1] + tensor.nelement())
        self._start_idx = indices[:-1]
        self._end_idx = indices[1:]
        self._tensors = tensors

        first = tensors[0]
        if first.device.type != "cuda":
            raise RuntimeError("TensorBuffer (B200 build) packs CUDA tensors only; there is no CPU fallback")
        for t in tensors:
            if t.dtype != first.dtype or t.device != first.device:
                raise RuntimeError("TensorBuffer: all tensors must share dtype and device (torch.cat would raise too)")
        from . import _lib

        self._libmod = _lib
        self._lib = _lib.load()
        self._device = first.device
        handle = ctypes.c_void_p()
        _lib.check(self._lib.psgd_tb_create(len(tensors), _lib.i64_array([t.nelement() for t in tensors]),
                                            first.element_size(), first.device.index, ctypes.byref(handle)),
                   self._lib)
        self._h = handle
        self.buffer = torch.empty(indices[-1], dtype=first.dtype, device=first.device)
        self.pack()  # copies

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.psgd_tb_destroy(h)
            except Exception:
                pass

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self._device).cuda_stream)

    def _same_layout(self, tensors):
        if len(tensors) != len(self._tensors):
            return False
        return all(t.is_contiguous() and t.dtype == self.buffer.dtype and t.device == self.buffer.device
                   and t.nelement() == e - s for t, s, e in zip(tensors, self._start_idx, self._end_idx))

    def __getitem__(self, index):
        return self.buffer[self._start_idx[index] : self._end_idx[index]].view(*self._tensors[index].shape)

    def __len__(self):
        return len(self._tensors)

    def pack(self, tensors=None):
        # Optional. init already does this.
        if tensors is None:
            tensors = self._tensors
        tensors = list(tensors)
        if self._same_layout(tensors):
            ptrs = self._libmod.ptr_array([t.data_ptr() for t in tensors])
            self._libmod.check(self._lib.psgd_tb_pack(self._h, ptrs, self.buffer.data_ptr(), self._stream()), self._lib)
        else:  # broadcasting / strided / converting assignments keep upstream's per-entry semantics
            for tensor, entry in zip(tensors, self):
                entry[:] = tensor

    def unpack(self, tensors):
        tensors = list(tensors)
        if self._same_layout(tensors):
            ptrs = self._libmod.ptr_array([t.data_ptr() for t in tensors])
            self._libmod.check(self._lib.psgd_tb_unpack(self._h, self.buffer.data_ptr(), ptrs, self._stream()), self._lib)
        else:
            for tensor, entry in zip(tensors, self):
                tensor[:] = entry

    def nelement(self):
        return self.buffer.nelement()

    def element_size(self):
        return self.buffer.element_size()

    def bits(self):
        return 8 * self.nelement() * self.element_size()

    def all_reduce(self, async_op=False):
        return torch.distributed.all_reduce(self.buffer, async_op=async_op)

    def all_gather(self, async_op=False):
        n_workers = torch.distributed.get_world_size() if torch.distributed.is_available() else 1
        buffers = [torch.empty_like(self.buffer) for i in range(n_workers)]
        handle = all_gather(buffers, self.buffer, async_op=async_op)
        if async_op:
            return buffers, handle
        else:
            return buffers


def all_reduce(*args, **kwargs):
    if torch.distributed.is_available() and torch.distributed.get_world_size() > 1:
        return torch.distributed.all_reduce(*args, **kwargs)


def all_gather(out_list, in_tensor, **kwargs):
    if torch.distributed.is_available() and torch.distributed.get_world_size() > 1:
        return torch.distributed.all_gather(out_list, in_tensor, **kwargs)
    else:
        assert len(out_list) == 1
        out_list[0].data = in_tensor
