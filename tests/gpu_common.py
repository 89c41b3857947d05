"""Helpers shared by the -m gpu tests (imported lazily; nothing here runs without CUDA)."""
import os
import tempfile

import numpy as np
import torch
import torch.distributed as dist

from oracle.powersgd_oracle import OracleReducer
from tests.helpers import flatten, per_tensor_mem_err, per_tensor_rel_err

TOL = 1e-5  # BASELINE.json north_star: fp32 relative tolerance per tensor (Frobenius)


def ensure_group():
    if not dist.is_initialized():
        f = tempfile.NamedTemporaryFile(delete=False)
        f.close()
        os.unlink(f.name)
        dist.init_process_group("gloo", init_method="file://" + f.name, world_size=1, rank=0)


def within_tolerance(res, tol=TOL):
    """The parity gate of one step: every tensor within `tol` of the fp32 oracle, or -- for tensors whose
    rank-r factor is ill-conditioned (square P when r = min(n, m): modified Gram-Schmidt then loses
    ~cond(P) * eps in ANY fp32 implementation, SURVEY.md section 7 'Tolerance') -- at least as close to the
    fp64 oracle as the fp32 oracle itself is (factor 2 slack)."""
    bad = []
    for key in ("out", "mem"):
        for i, e in enumerate(res[key]):
            if e < tol:
                continue
            mine64, ref64 = res[key + "_vs64"][i], res[key + "_ref_vs64"][i]
            if not (mine64 <= 2.0 * ref64 + 1e-7 and e < 20 * tol):
                bad.append((key, i, e, mine64, ref64))
    return bad


def run_pair(shapes, rank, steps=1, reuse=True, seed=714, data_seed=0, extra=0, offset_elems=0, scale_cols=False,
             arbiter=True, sync_q=False):
    """Run the CUDA reducer and the CPU oracle on the same seeded inputs and the same query.

    Returns a list (per step) of dicts with per-tensor relative errors of grad_out / memory_out and the
    relative errors of the packed P / Q buffers.  ``offset_elems`` places every CUDA tensor that many
    elements into a larger allocation (exercises the unaligned code paths).  ``sync_q`` also hands the oracle's Q to the
    CUDA reducer after every step, so that with ``reuse_query`` every step is a single-step comparison."""
    from network_distributed_pytorch_b200.reducer import PowerSGDReducer

    ensure_group()
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(data_seed)

    def on_gpu(t):
        if offset_elems == 0:
            return t.to(dev)
        big = torch.zeros(t.nelement() + offset_elems + 4, device=dev)
        v = big[offset_elems : offset_elems + t.nelement()].view(t.shape)
        v.copy_(t)
        return v

    red = PowerSGDReducer(seed, dev, 0, reuse, rank, extra_power_iterations=extra)
    orc = OracleReducer(seed, "cpu", n_power_iterations=extra, reuse_query=reuse, rank=rank)
    orc64 = (OracleReducer(seed, "cpu", n_power_iterations=extra, reuse_query=reuse, rank=rank, dtype=torch.float64)
             if arbiter else None)
    mem_c = [torch.zeros(s) for s in shapes]
    mem_g = [on_gpu(t) for t in mem_c]
    results = []
    for step in range(steps):
        g = [torch.randn(s, generator=gen) for s in shapes]
        if scale_cols:  # steep spectrum: columns scaled geometrically
            g = [t * torch.logspace(0, -3, t.shape[-1]) if t.ndimension() > 1 else t for t in g]
        send_c = [a + b for a, b in zip(g, mem_c)]
        send_g = [on_gpu(a) + b for a, b in zip(g, mem_g)] if offset_elems == 0 else [on_gpu(a + b.cpu()) for a, b in zip(g, mem_g)]
        out_c = [torch.zeros(s) for s in shapes]
        out_g = [on_gpu(torch.zeros(s)) for s in shapes]
        drew = red.p_memory is None or not reuse
        bits_g = red.reduce(send_g, out_g, mem_g)
        torch.cuda.synchronize()
        # same query for the oracle: take what the CUDA run drew (torch.randn differs per device)
        q_init = None
        if drew:
            q_init = _last_q_init(red, shapes, rank, step, seed)
        mem_prev = [t.double() for t in mem_c] if arbiter else None
        q_prev = orc.q_memory.clone() if orc.q_memory is not None else None
        bits_c = orc.reduce(send_c, out_c, mem_c, q_init=q_init)
        assert bits_g == bits_c
        if not arbiter:  # large sets: fp32 oracle only
            results.append(dict(
                out=per_tensor_rel_err(flatten(out_g), flatten(out_c), shapes),
                mem=per_tensor_mem_err(flatten(mem_g), flatten(mem_c), flatten(send_c), shapes),
                p=per_tensor_rel_err(red.p_memory.cpu().numpy(), orc.p_memory.numpy(), [(red.p_memory.nelement(),)])[0],
                q=per_tensor_rel_err(red.q_memory.cpu().numpy(), orc.q_memory.numpy(), [(red.q_memory.nelement(),)])[0],
                send_equal=all(torch.equal(a.cpu(), b) for a, b in zip(send_g, send_c))))
            for a, b in zip(mem_g, mem_c):
                a.copy_(b)
            if sync_q:
                red.q_memory.copy_(orc.q_memory)
            continue
        # fp64 arbiter, single step from the same state as the fp32 runs
        out_d = [torch.zeros(s, dtype=torch.float64) for s in shapes]
        if q_init is None and q_prev is not None:
            orc64.q_memory = q_prev.double()
            orc64.p_memory = orc.p_memory.double()
        bits_d = orc64.reduce([t.double() for t in send_c], out_d, mem_prev, q_init=q_init)
        assert bits_d == bits_c
        results.append(dict(
            out=per_tensor_rel_err(flatten(out_g), flatten(out_c), shapes),
            mem=per_tensor_mem_err(flatten(mem_g), flatten(mem_c), flatten(send_c), shapes),
            p=per_tensor_rel_err(red.p_memory.cpu().numpy(), orc.p_memory.numpy(), [(red.p_memory.nelement(),)])[0],
            q=per_tensor_rel_err(red.q_memory.cpu().numpy(), orc.q_memory.numpy(), [(red.q_memory.nelement(),)])[0],
            send_equal=all(torch.equal(a.cpu(), b) for a, b in zip(send_g, send_c)),
            out_vs64=per_tensor_rel_err(flatten(out_g), flatten(out_d), shapes),
            out_ref_vs64=per_tensor_rel_err(flatten(out_c), flatten(out_d), shapes),
            mem_vs64=per_tensor_mem_err(flatten(mem_g), flatten(mem_prev), flatten(send_c), shapes),
            mem_ref_vs64=per_tensor_mem_err(flatten(mem_c), flatten(mem_prev), flatten(send_c), shapes),
        ))
        # keep the two trajectories on identical inputs: feed the oracle's memory to both
        for a, b in zip(mem_g, mem_c):
            a.copy_(b)
        if sync_q:
            red.q_memory.copy_(orc.q_memory)
    return results


def _last_q_init(red, shapes, rank, step, seed):
    """Re-draw on the GPU exactly what PowerSGDReducer._set_random drew at this step."""
    from oracle.powersgd_oracle import burned_rng

    rng = burned_rng(seed)
    n_mat = sum(1 for s in shapes if len(s) > 1)
    for _ in range(step * n_mat):  # earlier draw rounds (reuse_query=False draws every step)
        rng.randint(1_000_000_000)
    keep = torch.random.get_rng_state(), torch.cuda.get_rng_state()
    parts = []
    for s in shapes:
        if len(s) <= 1:
            continue
        n = s[0]
        m = int(np.prod(s[1:]))
        r = min(n, m, rank)
        torch.manual_seed(int(rng.randint(1_000_000_000)))
        parts.append(torch.randn(m, r, device=red.device).reshape(-1).cpu())
    torch.random.set_rng_state(keep[0])
    torch.cuda.set_rng_state(keep[1])
    return torch.cat(parts)
