"""The upstream reducer's op stream, op for op, on any torch device.  TEST / BASELINE INFRASTRUCTURE ONLY.

``powersgd_oracle.OracleReducer`` emulates a world inside one process and is shaped for checking.  This
class is shaped for *timing*: it issues the very sequence of library calls the upstream
``PowerSGDReducer.reduce`` issues (ddp_powersgd_guide_cifar10/reducer.py:43-170) -- one
``torch.matmul(..., out=)`` per matrix and phase written in place, the real ``torch.distributed``
collectives, the eager column-by-column Gram-Schmidt, the ``sub`` + ``copy_`` pair for the local error,
``torch.cat`` for the flat buffer and one ``copy_`` per <=1-D tensor on the way back -- so that

* on the GPU box it is "the reference algorithm as stock PyTorch ops on the same B200"
  (bench.py ``torch_gpu_baseline``: ~1,530 kernels per DistilBERT step), and
* on the host it is the gloo world-size-2 CPU baseline of BASELINE.md section 4
  (bench.py ``cpu_baseline_gloo2``).

The upstream tree itself is Python that cannot travel to the GPU box; ``tests/test_oracle_golden.py``
pins this restatement bit for bit to vectors produced by the unmodified upstream file.

Only ``tests/`` and bench.py's baseline legs import this module; the product never does.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .powersgd_oracle import bit_count, burned_rng, gram_schmidt_inplace


def _world():
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def _sum_all_reduce(t, async_op=False):
    """reducer.py:193-195 (guarded) -- returns a work handle when async."""
    if _world() > 1:
        return dist.all_reduce(t, async_op=async_op)
    return None


class StreamReducer:
    """reducer.py:26-170 with the upstream constructor signature."""

    def __init__(self, random_seed, device, n_power_iterations=0, reuse_query=True, rank=1):
        assert n_power_iterations == 0  # reducer.py:30
        self.rng = burned_rng(random_seed)  # reducer.py:8-12
        self.device = torch.device(device)
        self.n_workers = _world()
        self.rank = rank
        self.reuse_query = reuse_query
        self.p_memory = None
        self.q_memory = None

    def _draw(self, q):  # reducer.py:36-38
        torch.manual_seed(int(self.rng.randint(1_000_000_000)))
        q.data[:] = torch.randn(*q.shape, device=self.device)

    def reduce(self, grad_in, grad_out, memory_out):
        W = self.n_workers
        bits = 0
        # reducer.py:53-62
        lo = [(a, b, c) for a, b, c in zip(grad_in, grad_out, memory_out) if a.ndimension() <= 1]
        hi = [(a, b, c) for a, b, c in zip(grad_in, grad_out, memory_out) if a.ndimension() > 1]

        # reducer.py:72-98
        views = [a.view(a.shape[0], -1) for a, _, _ in hi]
        ranks = [min(v.shape[0], v.shape[1], self.rank) for v in views]
        first = self.p_memory is None
        if first:
            self.p_memory = torch.empty(sum(v.shape[0] * r for v, r in zip(views, ranks)), device=self.device)
            self.q_memory = torch.empty(sum(v.shape[1] * r for v, r in zip(views, ranks)), device=self.device)
        ps, qs, po, qo = [], [], 0, 0
        for v, r in zip(views, ranks):
            n, m = v.shape
            ps.append(self.p_memory[po : po + n * r].view(n, r))
            qs.append(self.q_memory[qo : qo + m * r].view(m, r))
            po += n * r
            qo += m * r

        # reducer.py:101-111
        if first or not self.reuse_query:
            for q in qs:
                self._draw(q)

        # reducer.py:121-128
        for v, q, p in zip(views, qs, ps):
            torch.matmul(v, q, out=p)
        _sum_all_reduce(self.p_memory)
        bits += bit_count(self.p_memory)
        self.p_memory.data[:] /= W

        # reducer.py:131-133 (a fresh cat every step, asynchronous collective)
        flat = torch.cat([a.reshape(-1) for a, _, _ in lo])
        handle = _sum_all_reduce(flat, async_op=True)
        bits += bit_count(flat)

        # reducer.py:136-137
        for p in ps:
            gram_schmidt_inplace(p)

        # reducer.py:140-147
        for v, q, p in zip(views, qs, ps):
            torch.matmul(v.t(), p, out=q)
        _sum_all_reduce(self.q_memory)
        bits += bit_count(self.q_memory)
        self.q_memory.data[:] /= W

        # reducer.py:158-163: the product lands in grad_out's storage, then sub + copy_ for the local error
        for p, q, v, (a, out, mem) in zip(ps, qs, views, hi):
            torch.matmul(p, q.t(), out=out.data.view(v.shape))
            mem.data[:] = a - out

        # reducer.py:166-168
        if handle is not None:
            handle.wait()
        flat /= W
        off = 0
        for _, out, _ in lo:
            k = out.nelement()
            out.data[:] = flat[off : off + k].view(out.shape)
            off += k
        return bits
