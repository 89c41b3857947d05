"""Peer-memory exchange for the two packed all-reduces of a step (reducer.py:126,132,145) inside one NVLink domain.

The input buffers of the two phases live in torch symmetric memory (every rank maps every other rank's buffer and
signal pad); the sum itself is this package's own kernel.  Where the fabric gives the buffers a multicast mapping
(NVSwitch), psgd_peer_allreduce_mc sums rank k's slice inside the switch (multimem.ld_reduce) and stores it into every
rank's copy (multimem.st); otherwise psgd_peer_allreduce has each rank read all ranks' inputs and add them in rank
order.  Either way every rank gets the same bits.  Set-up is a collective; if it fails on any rank every rank falls
back to torch.distributed.all_reduce (NCCL).

PSGD_PEER_ALLREDUCE: 1 / 0 force it on / off; unset, the multicast kernel is used from 8 ranks on and NCCL below
(PSGD_PEER_MULTICAST=0 forces the pull kernel where the exchange is on).  Measured DistilBERT r=4 step on B200s:
2 GPUs NCCL 0.280 ms, pull 0.278, multicast 0.302; 8 GPUs NCCL 0.307, pull 0.318, multicast 0.294."""
import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib

_SLOT0 = 256          # first signal-pad slot used by this package (torch's own barrier uses the low slots)
MAX_WORLD = 8


class PeerExchange:
    def __init__(self, device, n_pf, n_q):
        import torch.distributed._symmetric_memory as symm

        self.device = device
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self._lib = _lib.load()
        group = dist.group.WORLD
        pad4 = lambda n: (max(n, 4) + 3) // 4 * 4
        self.buf = [symm.empty(pad4(n_pf), dtype=torch.float32, device=device),
                    symm.empty(pad4(n_q), dtype=torch.float32, device=device)]
        self.hdl = [symm.rendezvous(b, group) for b in self.buf]
        for h in self.hdl:
            if len(h.buffer_ptrs) != self.world or h.signal_pad_size < 4 * (_SLOT0 + 2 * MAX_WORLD):
                raise RuntimeError("unexpected symmetric memory geometry")
        self._in = [_lib.ptr_array(list(h.buffer_ptrs)) for h in self.hdl]
        self._sig = [_lib.ptr_array(list(h.signal_pad_ptrs)) for h in self.hdl]
        # NVSwitch multicast mapping of the two buffers (0 when the fabric has none): in-switch two-shot sum
        self._mc = [int(getattr(h, "multicast_ptr", 0) or 0) for h in self.hdl]
        if os.environ.get("PSGD_PEER_MULTICAST", "1") != "1" or not all(self._mc):
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
        torch.cuda.synchronize(device)
        dist.barrier()

    def describe(self):
        return {"kernel-multicast": "own kernel, NVSwitch multicast (multimem), two-shot",
                "kernel-pull": "own kernel, peer pull"}[self.kind]

    def allreduce(self, which, out, n, div=1.0, div_from=0):
        """out[:n] = sum over ranks of buf[which][:n]  (which: 0 = P || flat, 1 = Q).  With the multicast kernel the
        elements from div_from on leave as sum / div (callers check .multicast before asking for that)."""
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


AUTO_MIN_WORLD = 8   # from this many ranks on the multicast kernel beat NCCL in the measurements


def mode_for(world, setting):
    """What PSGD_PEER_ALLREDUCE asks for: 'off', 'on' (either kernel) or 'multicast-only'.
    '1' / '0' force it on / off; unset (or 'auto') turns the multicast kernel on from AUTO_MIN_WORLD ranks."""
    if world < 2 or world > MAX_WORLD or setting == "0":
        return "off"
    if setting == "1":
        return "on"
    return "multicast-only" if world >= AUTO_MIN_WORLD else "off"


def try_setup(device, n_pf, n_q):
    """Collective.  Returns a PeerExchange on every rank, or None on every rank."""
    world = dist.get_world_size()
    mode = mode_for(world, os.environ.get("PSGD_PEER_ALLREDUCE", "auto"))
    if mode == "off" or device.type != "cuda" or dist.get_backend() != "nccl":
        return None
    ex, ok = None, 1
    try:
        ex = PeerExchange(device, n_pf, n_q)
        if mode == "multicast-only" and not ex.multicast:
            ok = 0  # the pull kernel loses to NCCL at this size
    except Exception as exc:  # noqa: BLE001 -- any failure means "use NCCL", decided jointly below
        ok = 0
        if os.environ.get("PSGD_DEBUG"):
            print("peer exchange unavailable:", repr(exc))
    flag = torch.tensor([ok], device=device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return ex if int(flag.item()) == 1 else None
