"""Fused form of the caller-side update that follows ``reducer.reduce`` in the upstream training loops
(ddp_powersgd_guide_cifar10/ddp_init.py:166-178, same lines in the DistilBERT script):

    for g, momentum in zip(grads, momenta):
        momentum.data = g.clone()            # first step          | momentum.mul_(lambda).add_(g)   # later steps
        g.data[:] += momentum
    for param, g in zip(model.parameters(), grads):
        param.data.add_(g, alpha=-lr)

Upstream that is four to five kernel launches per parameter; here it is one launch over a tile table of all
parameters.  CUDA fp32 only (no CPU fallback)."""
import ctypes

import torch

from . import _lib


class FusedMomentumSGD:
    def __init__(self, params, lr, momentum):
        params = list(params)
        if not params:
            raise ValueError("no parameters")
        self.device = params[0].device
        if self.device.type != "cuda":
            raise RuntimeError("FusedMomentumSGD (B200 build) needs CUDA tensors; there is no CPU fallback")
        self.lr, self.momentum = float(lr), float(momentum)
        self._n = len(params)
        self._lib = _lib.load()
        h = ctypes.c_void_p()
        _lib.check(self._lib.psgd_sgd_create(self._n, _lib.i64_array([p.nelement() for p in params]),
                                             self.device.index, ctypes.byref(h)), self._lib)
        self._h = h
        self._first = True

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.psgd_sgd_destroy(h)
            except Exception:
                pass

    def step(self, params, grads, momenta):
        """m = g (first call) or lambda*m + g;  g += m;  p -= lr*g  -- in place on all three lists."""
        params, grads, momenta = list(params), list(grads), list(momenta)
        if not (len(params) == len(grads) == len(momenta) == self._n):
            raise ValueError("params, grads and momenta must have the length given at construction")
        for lst in (params, grads, momenta):
            for t in lst:
                if t.dtype != torch.float32 or not t.is_contiguous() or t.device != self.device:
                    raise RuntimeError("FusedMomentumSGD needs contiguous float32 tensors on its device")
        arrs = [_lib.ptr_array([t.data_ptr() for t in lst]) for lst in (params, grads, momenta)]
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self._lib.psgd_sgd_step(self._h, *arrs, self.lr, self.momentum, int(self._first), stream), self._lib)
        self._first = False
