"""CPU test double for the engine interface of network_distributed_pytorch_b200.engine.

TEST INFRASTRUCTURE.  It implements the same phase calls as CudaPlan on CPU tensors with the
oracle's arithmetic, so the reducer's *host* logic (list splitting, plan caching, Q seeding,
collective ordering, folded divides, bit accounting) can be exercised without a GPU, including
world_size-2 runs over gloo.  The product never constructs it: PowerSGDReducer builds a
CudaEngine unless a test passes ``_engine=``.
"""
import torch

from oracle.powersgd_oracle import gram_schmidt_inplace


class OraclePlan:
    def __init__(self, mat_dims, rank, vec_lens):
        self.mats = []
        po = qo = 0
        for n, m in mat_dims:
            r = min(n, m, rank)
            self.mats.append(dict(n=n, m=m, r=r, p_off=po, q_off=qo))
            po += n * r
            qo += m * r
        self.p_elems, self.q_elems = po, qo
        self.vec_lens = list(vec_lens)
        self.vec_elems = sum(vec_lens)
        self.calls = []

    def _views(self, p_mem, q_mem):
        for i, d in enumerate(self.mats):
            p = p_mem[d["p_off"] : d["p_off"] + d["n"] * d["r"]].view(d["n"], d["r"]) if p_mem is not None else None
            q = q_mem[d["q_off"] : d["q_off"] + d["m"] * d["r"]].view(d["m"], d["r"]) if q_mem is not None else None
            yield i, d, p, q

    def set_ptrs(self, mat_in, mat_out, mat_mem, vec_in, vec_out, grad=None, err=None, vec_grad=None, vec_err=None):
        self.calls.append("set_ptrs")
        self.mat_in, self.mat_out, self.mat_mem = mat_in, mat_out, mat_mem
        self.vec_in, self.vec_out, self.grad, self.err = vec_in, vec_out, grad, err
        self.vec_grad, self.vec_err = vec_grad, vec_err

    def pack_vec_fused(self, flat):
        self.calls.append("pack_vec_fused")
        for s, g, e in zip(self.vec_in, self.vec_grad, self.vec_err):
            s[:] = g + e
        flat[:] = torch.cat([t.reshape(-1) for t in self.vec_in])

    def pack_vec(self, flat):
        self.calls.append("pack_vec")
        flat[:] = torch.cat([t.reshape(-1) for t in self.vec_in])

    def p_from_m(self, p_mem, q_mem, fuse_add=False, q_div=1.0):
        self.calls.append("p_from_m")
        for i, d, p, q in self._views(p_mem, q_mem):
            if fuse_add:
                self.mat_in[i][:] = self.grad[i] + self.err[i]
            torch.matmul(self.mat_in[i].view(d["n"], d["m"]), q / q_div if q_div != 1.0 else q, out=p)

    def orthogonalize(self, p_mem, n_workers):
        self.calls.append("orthogonalize")
        p_mem /= n_workers
        for _, _, p, _ in self._views(p_mem, None):
            gram_schmidt_inplace(p)

    def q_from_mt(self, p_mem, q_mem):
        self.calls.append("q_from_mt")
        for i, d, p, q in self._views(p_mem, q_mem):
            torch.matmul(self.mat_in[i].view(d["n"], d["m"]).t(), p, out=q)

    def finalize(self, p_mem, q_mem, n_workers):
        self.calls.append("finalize")
        for i, d, p, q in self._views(p_mem, q_mem):
            approx = torch.matmul(p, (q / n_workers).t())
            m = self.mat_in[i].view(d["n"], d["m"])
            self.mat_out[i].view(d["n"], d["m"])[:] = approx
            self.mat_mem[i].view(d["n"], d["m"])[:] = m - approx

    def finish(self, flat, q_mem, n_workers):
        self.calls.append("finish")
        off = 0
        for t, k in zip(self.vec_out, self.vec_lens):
            t.reshape(-1)[:] = flat[off : off + k] / n_workers
            off += k
        if q_mem is not None:
            q_mem /= n_workers


class OracleEngine:
    name = "oracle-test-double"

    def __init__(self, device="cpu"):
        self.device = torch.device(device)
        self.plans = []

    def make_plan(self, mat_dims, rank, vec_lens):
        plan = OraclePlan(mat_dims, rank, vec_lens)
        self.plans.append(plan)
        return plan

    def empty(self, n):
        return torch.empty(n, dtype=torch.float32, device=self.device)
