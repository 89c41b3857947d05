"""CPU restatement of the reference PowerSGD reducer.  TEST INFRASTRUCTURE ONLY.

This file is the checker, not the product: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline / ``--impl reference`` legs may import it.  Nothing in
``network_distributed_pytorch_b200/`` imports it, and the product path raises when the
CUDA library is missing instead of falling back to anything here.

What it restates (all citations relative to the upstream tree, the two copies of
each file are byte-identical):

* ``ddp_powersgd_guide_cifar10/reducer.py:7-19``    -> :func:`burned_rng`, :class:`OracleReducer.__init__`
* ``ddp_powersgd_guide_cifar10/reducer.py:36-41``   -> :meth:`OracleReducer._draw_query`
* ``ddp_powersgd_guide_cifar10/reducer.py:43-170``  -> :meth:`OracleReducer.reduce`
* ``ddp_powersgd_guide_cifar10/reducer.py:180-191`` -> :func:`gram_schmidt_inplace`
* ``ddp_powersgd_guide_cifar10/reducer.py:197-198`` -> :func:`bit_count`
* ``ddp_powersgd_guide_cifar10/tensor_buffer.py:4-57`` -> :class:`FlatPack`

Parity pin: the reference ships no tests or golden vectors, so
the pin is the reference itself, executed in the build container by
``tests/golden/make_golden.py`` (which imports the upstream ``reducer.py`` by path).
``tests/test_oracle_golden.py`` checks this restatement against those fixtures
bit-for-bit on CPU fp32.

Differences from the upstream code that are deliberate:

* works without ``torch.distributed``: a "world" is emulated by handing ``reduce`` one
  gradient list per worker and summing in worker order (exact for world 2, where the
  sum is commutative; for larger worlds the collective's summation order is
  implementation-defined upstream as well);
* ``dtype=torch.float64`` gives the high-precision arbiter used for tolerance arguments;
* ``n_power_iterations >= 1`` repeats steps 3-7 (upstream asserts 0, reducer.py:30);
* the Q initialisation can be injected (``q_init=``) so a CUDA run and a CPU run can be
  compared on the same query although ``torch.randn`` differs per device.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch

BURN_DRAWS = 128 * 1024 * 1024  # reducer.py:9-11
_BURN_CHUNK = 8 * 1024 * 1024
_burn_cache = {}


def burned_rng(seed: int) -> np.random.RandomState:
    """RandomState(seed) advanced past the 128 Mi normal draws of reducer.py:8-12.

    The legacy polar method is rejection based, so the draws have to be replayed; they
    are replayed in chunks (the cached second gaussian travels in the state, so chunking
    does not change the stream) and the resulting state is cached per seed.
    """
    if seed not in _burn_cache:
        rng = np.random.RandomState(seed)
        left = BURN_DRAWS
        while left:
            k = min(left, _BURN_CHUNK)
            rng.randn(k)
            left -= k
        _burn_cache[seed] = rng.get_state()
    out = np.random.RandomState()
    out.set_state(_burn_cache[seed])
    return out


def bit_count(t: torch.Tensor) -> int:
    """reducer.py:197-198."""
    return 8 * t.nelement() * t.element_size()


def gram_schmidt_inplace(mat: torch.Tensor, eps: float = 1e-8) -> None:
    """Modified Gram-Schmidt, column by column, in place (reducer.py:180-191).

    Same op sequence as the TorchScript function: normalise column i by
    ``sqrt(sum(col**2)) + eps`` (eps outside the root), then remove its component from
    every later column with ``rest -= sum(col*rest, dim=0) * col``.
    """
    eps_t = torch.tensor(eps, dtype=mat.dtype, device=mat.device)
    n_cols = mat.shape[1]
    for i in range(n_cols):
        col = mat[:, i : i + 1]
        col /= torch.sqrt(torch.sum(col ** 2)) + eps_t
        if i + 1 < n_cols:
            rest = mat[:, i + 1 :]
            rest -= torch.sum(col * rest, dim=0) * col


class FlatPack:
    """tensor_buffer.py:4-57 restated: concatenate, view back, copy out."""

    def __init__(self, tensors: Sequence[torch.Tensor]):
        ends = np.cumsum([0] + [t.nelement() for t in tensors]).tolist()
        self.starts, self.ends = ends[:-1], ends[1:]
        self.shapes = [tuple(t.shape) for t in tensors]
        self.buffer = torch.cat([t.reshape(-1) for t in tensors])  # raises on empty list, as upstream

    def __len__(self):
        return len(self.shapes)

    def __getitem__(self, i):
        return self.buffer[self.starts[i] : self.ends[i]].view(*self.shapes[i])

    def pack(self, tensors):
        for i, t in enumerate(tensors):
            self[i][:] = t

    def unpack(self, tensors):
        for i, t in enumerate(tensors):
            t[:] = self[i]

    def bits(self):
        return bit_count(self.buffer)


class OracleReducer:
    """reducer.py:26-170 for an emulated world of ``n_workers`` replicas in one process.

    ``reduce`` takes ``grad_in``/``grad_out``/``memory_out`` either as plain tensor lists
    (world 1) or as lists of per-worker lists.  P/Q live in ``p_memory``/``q_memory`` with
    the upstream packed layout (running offsets, no padding, reducer.py:86-98); after the
    emulated all-reduce every worker holds the same P and Q, so one copy is kept.
    """

    def __init__(self, random_seed: int, device="cpu", n_power_iterations: int = 0,
                 reuse_query: bool = True, rank: int = 1, n_workers: int = 1,
                 dtype: torch.dtype = torch.float32):
        self.rng = burned_rng(random_seed)
        self.device = torch.device(device)
        self.n_workers = n_workers
        self.n_power_iterations = n_power_iterations
        self.rank = rank
        self.reuse_query = reuse_query
        self.dtype = dtype
        self.p_memory: Optional[torch.Tensor] = None
        self.q_memory: Optional[torch.Tensor] = None
        self.last_query_init: Optional[torch.Tensor] = None  # flat copy of Q as drawn

    # reducer.py:36-41 -- note the global torch generator is reseeded on purpose
    def _draw_query(self, q: torch.Tensor) -> None:
        torch.manual_seed(int(self.rng.randint(1_000_000_000)))
        q[:] = torch.randn(*q.shape, device=self.device).to(self.dtype)

    @staticmethod
    def _as_world(lists, n_workers):
        if len(lists) and isinstance(lists[0], torch.Tensor):
            assert n_workers == 1
            return [list(lists)]
        assert len(lists) == n_workers
        return [list(l) for l in lists]

    def reduce(self, grad_in, grad_out, memory_out, q_init: Optional[torch.Tensor] = None) -> int:
        W = self.n_workers
        gin, gout, gmem = (self._as_world(x, W) for x in (grad_in, grad_out, memory_out))
        bits = 0

        # reducer.py:53-62: split by dimensionality, order preserved
        vec_idx = [i for i, t in enumerate(gin[0]) if t.ndimension() <= 1]
        mat_idx = [i for i, t in enumerate(gin[0]) if t.ndimension() > 1]

        # reducer.py:72-98: shapes, ranks, packed offsets
        dims = []
        for i in mat_idx:
            n = gin[0][i].shape[0]
            m = gin[0][i].nelement() // n
            dims.append((n, m, min(n, m, self.rank)))
        first = self.p_memory is None
        if first:
            self.p_memory = torch.empty(sum(n * r for n, _, r in dims), device=self.device, dtype=self.dtype)
            self.q_memory = torch.empty(sum(m * r for _, m, r in dims), device=self.device, dtype=self.dtype)
        ps, qs, po, qo = [], [], 0, 0
        for n, m, r in dims:
            ps.append(self.p_memory[po : po + n * r].view(n, r))
            qs.append(self.q_memory[qo : qo + m * r].view(m, r))
            po += n * r
            qo += m * r

        # reducer.py:101-111
        if not (self.reuse_query and not first):
            for q in qs:
                self._draw_query(q)
            if q_init is not None:
                self.q_memory[:] = q_init.to(self.dtype).to(self.device)
            self.last_query_init = self.q_memory.clone()

        mats = [[gin[w][i].reshape(d[0], d[1]).to(self.dtype) for i, d in zip(mat_idx, dims)] for w in range(W)]

        for it in range(self.n_power_iterations + 1):
            # reducer.py:121-128: P = M Q per worker, SUM all-reduce, divide
            for k in range(len(dims)):
                acc = torch.matmul(mats[0][k], qs[k])
                for w in range(1, W):
                    acc = acc + torch.matmul(mats[w][k], qs[k])
                ps[k][:] = acc
            bits += bit_count(self.p_memory)
            self.p_memory[:] /= W

            if it == 0:
                # reducer.py:131-133: flat buffer of the <=1-D tensors, SUM all-reduce started
                packs = [FlatPack([gin[w][i].to(self.dtype) for i in vec_idx]) for w in range(W)]
                flat = packs[0].buffer.clone()
                for w in range(1, W):
                    flat = flat + packs[w].buffer
                bits += packs[0].bits()

            # reducer.py:136-137
            for p in ps:
                gram_schmidt_inplace(p)

            # reducer.py:140-147: Q = M^T P_hat per worker, SUM all-reduce, divide
            for k in range(len(dims)):
                acc = torch.matmul(mats[0][k].t(), ps[k])
                for w in range(1, W):
                    acc = acc + torch.matmul(mats[w][k].t(), ps[k])
                qs[k][:] = acc
            bits += bit_count(self.q_memory)
            self.q_memory[:] /= W

        # reducer.py:158-163: decompress, local error (only for the >=2-D tensors)
        for k, i in enumerate(mat_idx):
            if W == 1 and gout[0][i].dtype == self.dtype and gout[0][i].is_contiguous():
                # one worker: as upstream, the product is written straight into grad_out's storage (reducer.py:161), so
                # the timed CPU baseline does not pay a pass the reference does not make
                out = gout[0][i].view(dims[k][0], dims[k][1])
                torch.matmul(ps[k], qs[k].t(), out=out)
                gmem[0][i].reshape(-1)[:] = (mats[0][k] - out).reshape(-1).to(gmem[0][i].dtype)
                continue
            approx = torch.matmul(ps[k], qs[k].t())
            for w in range(W):
                gout[w][i].reshape(-1)[:] = approx.reshape(-1).to(gout[w][i].dtype)
                gmem[w][i].reshape(-1)[:] = (mats[w][k] - approx).reshape(-1).to(gmem[w][i].dtype)

        # reducer.py:166-168: mean of the flat buffer scattered to the outputs only
        flat /= W
        packs[0].buffer = flat
        for w in range(W):
            packs[0].unpack([gout[w][i] for i in vec_idx])
        # communicated bits use the wire dtype of the reference (fp32) whatever self.dtype is
        return bits * 4 // self.p_memory.element_size() if self.p_memory.element_size() != 4 else bits
