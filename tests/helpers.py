"""Shared helpers for the test-suite: golden fixture access and flat <-> list plumbing."""
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["w1_r2_reuse", "w1_r4_fresh", "w2_r2_reuse", "w1_r8_clip", "w2_r3_fresh"]


class Golden:
    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.z = z
        self.shapes = [tuple(int(x) for x in s.split(",")) for s in z["shapes"]]
        self.rank = int(z["rank"])
        self.reuse = bool(z["reuse"])
        self.world = int(z["world"])
        self.steps = int(z["steps"])
        self.seed = int(z["seed"])

    def get(self, worker, step, key):
        return self.z[f"w{worker}_s{step}_{key}"]


def unflatten(flat, shapes, device="cpu", dtype=torch.float32):
    """Split a flat numpy array into freshly allocated tensors of the given shapes."""
    out, off = [], 0
    for s in shapes:
        k = int(np.prod(s)) if len(s) else 1
        out.append(torch.from_numpy(np.ascontiguousarray(flat[off : off + k])).to(dtype).reshape(s).to(device).contiguous())
        off += k
    assert off == flat.size
    return out


def flatten(tensors):
    return torch.cat([t.detach().reshape(-1).cpu() for t in tensors]).numpy()


def rel_err(a, b):
    """Relative Frobenius error |a-b| / |b| (0/0 -> 0)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b)
    num = np.linalg.norm(a - b)
    return 0.0 if num == 0 else num / max(den, 1e-300)


def per_tensor_rel_err(flat_a, flat_b, shapes):
    out, off = [], 0
    for s in shapes:
        k = int(np.prod(s)) if len(s) else 1
        out.append(rel_err(flat_a[off : off + k], flat_b[off : off + k]))
        off += k
    return out


def per_tensor_mem_err(flat_a, flat_b, flat_in, shapes):
    """Error of memory_out per tensor.  memory = M - P Q^T, so its absolute error is that of the
    decompressed gradient; when the rank-r approximation is exact (r = min(n, m), e.g. a (2, 768)
    classifier at rank 4) the reference's memory is itself pure rounding noise of size ~1e-7 |M| and a
    ratio of two noises says nothing.  Those tensors (|mem_ref| < 1e-3 |M|) are measured against |M|,
    all others against |mem_ref| as usual."""
    out, off = [], 0
    for s in shapes:
        k = int(np.prod(s)) if len(s) else 1
        a = np.asarray(flat_a[off : off + k], np.float64)
        b = np.asarray(flat_b[off : off + k], np.float64)
        m = np.asarray(flat_in[off : off + k], np.float64)
        nb, nm = np.linalg.norm(b), np.linalg.norm(m)
        den = nm if nb < 1e-3 * nm else nb
        num = np.linalg.norm(a - b)
        out.append(0.0 if num == 0 else num / max(den, 1e-300))
        off += k
    return out


class MediumGolden:
    """The reference-generated case at realistic sizes (tests/golden/w2_r4_medium.npz): world 2, rank 4, one step.
    Inputs are regenerated from their seeds by the very function the generating script used and checked against the
    stored checksums; P / Q are stored in full, grad_out / memory_out as sampled rows + per-tensor norms."""

    def __init__(self):
        import importlib.util

        spec = importlib.util.spec_from_file_location("_make_golden", os.path.join(GOLDEN_DIR, "make_golden.py"))
        self.gen = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(self.gen)
        self.z = np.load(os.path.join(GOLDEN_DIR, self.gen.MEDIUM_NAME + ".npz"))
        self.shapes = [tuple(int(x) for x in s.split(",")) for s in self.z["shapes"]]
        assert self.shapes == [tuple(s) for s in self.gen.MEDIUM_SHAPES]
        self.rank, self.world, self.seed = int(self.z["rank"]), int(self.z["world"]), int(self.z["seed"])

    def inputs(self, worker):
        """Regenerated send buffers; None when this torch build's CPU generator does not reproduce them."""
        send = self.gen.medium_inputs(worker)
        sumsq = np.array([float((t.double() ** 2).sum()) for t in send])
        head = np.stack([np.resize(t.reshape(-1)[:4].numpy(), 4) for t in send])
        if not (np.allclose(sumsq, self.z[f"w{worker}_in_sumsq"], rtol=1e-12) and np.array_equal(head, self.z[f"w{worker}_in_head"])):
            return None
        return send

    def get(self, worker, key):
        return self.z[f"w{worker}_{key}"]

    def check_outputs(self, worker, gout, gmem, tol):
        """Worst relative error of grad_out / memory_out of `worker` over the stored rows, vectors and norms."""
        worst = 0.0
        for i, s in enumerate(self.shapes):
            o = gout[i].detach().cpu()
            if len(s) > 1:
                rows = self.gen.medium_rows(i, s[0])
                m = gmem[i].detach().cpu()
                worst = max(worst, rel_err(o.reshape(s[0], -1)[rows].numpy(), self.get(worker, f"out_rows{i}")))
                ref_m = self.get(worker, f"mem_rows{i}")
                got_m = m.reshape(s[0], -1)[rows].numpy()
                if self.get(worker, "mem_norm")[i] < 1e-3 * np.sqrt(self.get(worker, "in_sumsq")[i]):
                    # exact-rank tensor: the reference's memory is rounding noise (see per_tensor_mem_err)
                    worst = max(worst, np.linalg.norm(got_m.astype(np.float64) - ref_m) /
                                np.sqrt(self.get(worker, "in_sumsq")[i] * len(rows) / s[0]))
                else:
                    worst = max(worst, rel_err(got_m, ref_m))
                    worst = max(worst, abs(float(m.double().norm()) - self.get(worker, "mem_norm")[i]) / self.get(worker, "mem_norm")[i])
                worst = max(worst, abs(float(o.double().norm()) - self.get(worker, "out_norm")[i]) / self.get(worker, "out_norm")[i])
            else:
                worst = max(worst, rel_err(o.numpy(), self.get(worker, f"out_vec{i}")))
        assert worst < tol, worst
        return worst
