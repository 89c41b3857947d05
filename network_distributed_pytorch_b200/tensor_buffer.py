"""Flat-buffer packer with the interface of the upstream ``tensor_buffer.py``
(ddp_powersgd_guide_cifar10/tensor_buffer.py:4-69): ``TensorBuffer(tensors)`` with ``buffer``, indexing, ``len``,
``pack``, ``unpack``, ``nelement``, ``element_size``, ``bits``, ``all_reduce``, ``all_gather`` and the two guarded
module-level collectives.  The semantics are upstream's -- the flat layout is bit-identical to
``torch.cat([t.view(-1) for t in tensors])`` (running offsets, no padding, dtype of the inputs) -- the mechanics are
not: flattening and un-flattening are ONE gather / scatter kernel over a device pointer table
(``psgd_tb_pack`` / ``psgd_tb_unpack``) instead of ``torch.cat`` plus a Python loop of per-tensor copies.
CUDA tensors only; there is no CPU fallback.
"""
import ctypes
from itertools import accumulate

import torch
import torch.distributed as dist


def _world_size():
    return dist.get_world_size() if dist.is_available() else 1


class TensorBuffer:
    """One contiguous buffer holding the elements of a list of tensors, in list order."""

    def __init__(self, tensors):
        members = list(tensors)
        if not members:
            # what torch.cat([]) says upstream (tensor_buffer.py:19)
            raise RuntimeError("torch.cat(): expected a non-empty list of Tensors")
        head = members[0]
        if head.device.type != "cuda":
            raise RuntimeError("TensorBuffer (B200 build) packs CUDA tensors only; there is no CPU fallback")
        if any(t.dtype != head.dtype or t.device != head.device for t in members):
            raise RuntimeError("TensorBuffer: all tensors must share dtype and device (torch.cat would raise too)")

        counts = [t.nelement() for t in members]
        self._bounds = [0, *accumulate(counts)]  # prefix sums: entry i starts at _bounds[i], ends at _bounds[i + 1]
        self._tensors = members
        self._device = head.device

        from . import _lib

        self._libmod = _lib
        self._lib = _lib.load()
        handle = ctypes.c_void_p()
        _lib.check(self._lib.psgd_tb_create(len(members), _lib.i64_array(counts), head.element_size(),
                                            head.device.index, ctypes.byref(handle)), self._lib)
        self._h = handle
        self.buffer = torch.empty(self._bounds[-1], dtype=head.dtype, device=head.device)
        self.pack()

    def __del__(self):
        handle, self._h = getattr(self, "_h", None), None
        if handle:
            try:
                self._lib.psgd_tb_destroy(handle)
            except Exception:
                pass

    # upstream exposes these two lists; keep them readable
    @property
    def _start_idx(self):
        return self._bounds[:-1]

    @property
    def _end_idx(self):
        return self._bounds[1:]

    def __len__(self):
        return len(self._tensors)

    def __getitem__(self, index):
        lo, hi = self._bounds[index], self._bounds[index + 1]
        return self.buffer.narrow(0, lo, hi - lo).view(self._tensors[index].shape)

    # ------------------------------------------------------------------------------------------------
    def _kernel_ok(self, tensors):
        """True when the list can go through the pointer-table kernel: same count, sizes, dtype, device, packed."""
        if len(tensors) != len(self._tensors):
            return False
        spans = zip(tensors, self._bounds, self._bounds[1:])
        return all(t.is_contiguous() and t.dtype == self.buffer.dtype and t.device == self.buffer.device
                   and t.nelement() == hi - lo for t, lo, hi in spans)

    def _launch(self, entry, tensors, *, into_buffer):
        ptrs = self._libmod.ptr_array([t.data_ptr() for t in tensors])
        stream = ctypes.c_void_p(torch.cuda.current_stream(self._device).cuda_stream)
        flat = self.buffer.data_ptr()
        rc = entry(self._h, ptrs, flat, stream) if into_buffer else entry(self._h, flat, ptrs, stream)
        self._libmod.check(rc, self._lib)

    def pack(self, tensors=None):
        """buffer <- tensors (the constructor has already done this for the original list)."""
        source = self._tensors if tensors is None else list(tensors)
        if self._kernel_ok(source):
            self._launch(self._lib.psgd_tb_pack, source, into_buffer=True)
            return
        # broadcasting, strided or converting assignments: upstream's per-entry semantics
        for i, t in enumerate(source[: len(self)]):
            self[i][...] = t

    def unpack(self, tensors):
        """tensors <- buffer."""
        targets = list(tensors)
        if self._kernel_ok(targets):
            self._launch(self._lib.psgd_tb_unpack, targets, into_buffer=False)
            return
        for i, t in enumerate(targets[: len(self)]):
            t[...] = self[i]

    # ------------------------------------------------------------------------------------------------
    def nelement(self):
        return self.buffer.nelement()

    def element_size(self):
        return self.buffer.element_size()

    def bits(self):
        return self.buffer.nelement() * self.buffer.element_size() * 8

    def all_reduce(self, async_op=False):
        # unguarded on purpose, like upstream: needs an initialised process group even for one worker
        return dist.all_reduce(self.buffer, async_op=async_op)

    def all_gather(self, async_op=False):
        copies = [torch.empty_like(self.buffer) for _ in range(_world_size())]
        work = all_gather(copies, self.buffer, async_op=async_op)
        return (copies, work) if async_op else copies


def all_reduce(*args, **kwargs):
    """dist.all_reduce, skipped (returns None) when there is a single worker."""
    if dist.is_available() and dist.get_world_size() > 1:
        return dist.all_reduce(*args, **kwargs)
    return None


def all_gather(out_list, in_tensor, **kwargs):
    """dist.all_gather; with a single worker the only output simply aliases the input."""
    if dist.is_available() and dist.get_world_size() > 1:
        return dist.all_gather(out_list, in_tensor, **kwargs)
    assert len(out_list) == 1
    out_list[0].data = in_tensor
    return None
