"""Peer-memory exchange for the packed all-reduces of a step (reducer.py:126,132,145) inside one NVLink domain.

The buffers the P = M Q and Q = M^T P kernels write live in torch symmetric memory (every rank maps every other rank's
buffer and signal pad).  Two families of implementation, all this package's own kernels:

* fused (default): no all-reduce launch at all.  The Gram-Schmidt kernel takes a flag barrier and loads its rows
  summed over the ranks; the streamed finalize kernel (and the P = M Q kernel of a later power iteration) does the same
  for Q in its prologue (psgd_orthogonalize_peer / psgd_finalize_peer / psgd_p_from_m_peer).  Three ways of forming the
  sum: "push" -- the producing kernels store their results with ``multimem.st`` into slot [rank] of every rank's receive
  buffer while they run, and after the barrier the consumers add W local slots in slot order (nothing crosses NVLink
  on the critical path; needs the multicast mapping); "multimem" -- one ``multimem.ld_reduce`` per element, the
  NVSwitch adds the copies; "pull" -- one load per rank added in rank order (needs no multicast mapping).  Push and
  pull are bit-identical on all ranks by construction, multimem was measured to be.
* kernel: the round-1 stand-alone kernels between the phases (two-shot through the multicast mapping where the fabric
  has one, else one-shot pull) -- kept as a comparison point and as the parity cross-check.

Set-up is a collective; if it fails on any rank every rank falls back to torch.distributed.all_reduce (NCCL).

PSGD_PEER_ALLREDUCE: 0 = NCCL, 1 = this module, unset = this module from 2 ranks on (inside one node).
PSGD_PEER_MODE: push | multimem | pull | kernel; unset: push where the fabric gives the buffers a multicast mapping,
else pull.  Measured DistilBERT r=4 step on B200s (profiles/r02_allreduce_modes.md): one GPU 0.237 ms; two GPUs push 0.267,
multimem 0.270, pull 0.272, NCCL 0.274, stand-alone kernels 0.290; eight GPUs push 0.280, multimem 0.291, stand-alone
kernels 0.294, NCCL 0.308, pull 0.328 (reads (W-1) x the message).  multimem sums came out bit-identical on all eight ranks.
"""
import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib

_SLOT0 = 256          # first signal-pad slot used by this package (torch's own barrier uses the low slots)
MAX_WORLD = 8
_MODE_IDS = {"multimem": 1, "pull": 2, "push": 3}


class PeerExchange:
    def __init__(self, device, n_pf, n_q, mode):
        import torch.distributed._symmetric_memory as symm

        self.device = device
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self._lib = _lib.load()
        group = dist.group.WORLD
        pad4 = lambda n: (max(n, 4) + 3) // 4 * 4
        self.stride = [pad4(n_pf), pad4(n_q)]
        slots = self.world if mode in ("push", "auto") else 1  # push: one receive slot per source rank
        self.buf = [symm.empty(slots * self.stride[0], dtype=torch.float32, device=device),
                    symm.empty(slots * self.stride[1], dtype=torch.float32, device=device)]
        self.hdl = [symm.rendezvous(b, group) for b in self.buf]
        for h in self.hdl:
            if len(h.buffer_ptrs) != self.world or h.signal_pad_size < 4 * (_SLOT0 + 2 * MAX_WORLD):
                raise RuntimeError("unexpected symmetric memory geometry")
        self._in = [_lib.ptr_array(list(h.buffer_ptrs)) for h in self.hdl]
        self._sig = [_lib.ptr_array(list(h.signal_pad_ptrs)) for h in self.hdl]
        # NVSwitch multicast mapping of the two buffers (0 when the fabric has none)
        self._mc = [int(getattr(h, "multicast_ptr", 0) or 0) for h in self.hdl]
        have_mc = all(self._mc) and os.environ.get("PSGD_PEER_MULTICAST", "1") == "1"
        if mode == "auto":
            mode = "push" if have_mc else "pull"
        if mode in ("multimem", "push") and not have_mc:
            mode = "pull"
        self.fused = mode in _MODE_IDS
        if self.fused:
            self.kind = "fused-" + mode
            self.multicast = False
        else:
            if not have_mc:
                self._mc = [0, 0]
            self.multicast = bool(self._mc[0])
            self.kind = "kernel-multicast" if self.multicast else "kernel-pull"
        self.epoch = [0, 0]
        self._done = torch.zeros(2, dtype=torch.int32, device=device)  # per-buffer CTA completion word of the multicast kernel
        for b in self.buf:
            b.zero_()
        # The flag slots must start below the first epoch.  Symmetric allocations are recycled, so a pad may still hold
        # the epochs of an earlier exchange (a second reducer in the same process): every rank clears its own slots
        # before anybody may signal.
        for h in self.hdl:
            pad = h.get_signal_pad(self.rank, (h.signal_pad_size // 4,), torch.int32)
            pad[_SLOT0 : _SLOT0 + 2 * MAX_WORLD].zero_()
        self._h = None
        self.push = mode == "push"
        if self.push:
            handle = ctypes.c_void_p()
            _lib.check(self._lib.psgd_peer_create_push(
                self.world, self.rank, device.index, ctypes.c_void_p(self.buf[0].data_ptr()), ctypes.c_void_p(self._mc[0]),
                self.stride[0], self._sig[0], ctypes.c_void_p(self.buf[1].data_ptr()), ctypes.c_void_p(self._mc[1]),
                self.stride[1], self._sig[1], _SLOT0, ctypes.byref(handle)), self._lib)
            self._h = handle
        elif self.fused:
            handle = ctypes.c_void_p()
            _lib.check(self._lib.psgd_peer_create(
                self.world, self.rank, device.index, self._in[0], self._sig[0],
                ctypes.c_void_p(self._mc[0] if mode == "multimem" else 0), self._in[1], self._sig[1],
                ctypes.c_void_p(self._mc[1] if mode == "multimem" else 0), _SLOT0, _MODE_IDS[mode], ctypes.byref(handle)),
                self._lib)
            self._h = handle
        torch.cuda.synchronize(device)
        dist.barrier()

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.psgd_peer_destroy(h)
            except Exception:
                pass

    def describe(self):
        return {"fused-push": "none as a launch: pushed by the producing kernels (multimem.st), summed from local slots inside the Gram-Schmidt and finalize kernels",
                "fused-pull": "none as a launch: summed on load inside the Gram-Schmidt and finalize kernels (one load per rank, rank order)",
                "fused-multimem": "none as a launch: summed on load inside the Gram-Schmidt and finalize kernels (multimem.ld_reduce in the NVSwitch)",
                "kernel-multicast": "own kernel, NVSwitch multicast (multimem), two-shot",
                "kernel-pull": "own kernel, peer pull"}[self.kind]

    @property
    def handle(self):
        return self._h

    def allreduce(self, which, out, n, div=1.0, div_from=0):
        """Stand-alone kernels ("kernel" mode): out[:n] = sum over ranks of buf[which][:n]  (which: 0 = P || flat,
        1 = Q).  With the multicast kernel the elements from div_from on leave as sum / div (callers check .multicast
        before asking for that)."""
        if self.fused:
            raise RuntimeError("a fused exchange has no stand-alone all-reduce")
        if div != 1.0 and not self._mc[which]:
            raise RuntimeError("the pull kernel returns plain sums")
        self.epoch[which] += 1
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        if self._mc[which]:
            _lib.check(self._lib.psgd_peer_allreduce_mc(self._in[which], self._sig[which], ctypes.c_void_p(self._mc[which]),
                                                        out.data_ptr(), n, self.rank, self.world, _SLOT0,
                                                        self.epoch[which], ctypes.c_float(div), div_from,
                                                        self._done.data_ptr() + 4 * which, self.device.index, stream),
                       self._lib)
            return
        _lib.check(self._lib.psgd_peer_allreduce(self._in[which], self._sig[which], out.data_ptr(), n, self.rank,
                                                 self.world, _SLOT0, self.epoch[which], self.device.index, stream),
                   self._lib)


def enabled_for(world, setting):
    """PSGD_PEER_ALLREDUCE: '1' / '0' force the peer exchange on / off; unset (or 'auto') turns it on from two ranks."""
    if world < 2 or world > MAX_WORLD or setting == "0":
        return False
    return True


def try_setup(device, n_pf, n_q):
    """Collective.  Returns a PeerExchange on every rank, or None on every rank."""
    world = dist.get_world_size()
    if not enabled_for(world, os.environ.get("PSGD_PEER_ALLREDUCE", "auto")):
        return None
    if device.type != "cuda" or dist.get_backend() != "nccl":
        return None
    mode = os.environ.get("PSGD_PEER_MODE", "auto")
    if mode not in ("auto", "push", "pull", "multimem", "kernel"):
        raise ValueError("PSGD_PEER_MODE must be auto, push, pull, multimem or kernel")
    ex, ok = None, 1
    try:
        ex = PeerExchange(device, n_pf, n_q, mode)
    except Exception as exc:  # noqa: BLE001 -- any failure means "use NCCL", decided jointly below
        ok = 0
        if os.environ.get("PSGD_DEBUG"):
            print("peer exchange unavailable:", repr(exc))
    flag = torch.tensor([ok], device=device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return ex if int(flag.item()) == 1 else None
