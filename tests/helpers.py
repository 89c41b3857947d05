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
