"""Generate tests/golden/*.npz by running the UNMODIFIED upstream reducer on CPU/gloo.

Run in the build container only (it needs /root/reference, which does not exist on
the GPU box):

    python tests/golden/make_golden.py

The upstream modules are imported by path (``ddp_powersgd_guide_cifar10/reducer.py``
with its own ``tensor_buffer.py`` first on ``sys.path``); nothing is copied.  The only
instrumentation is a wrapper around ``PowerSGDReducer._set_random`` that records the
query exactly as the reference drew it.  Each case stores, per step and per worker, the
flat concatenation of ``grad_in``, ``grad_out`` and ``memory_out`` plus ``p_memory``,
``q_memory``, the drawn query and the returned bit count.  The caller-side error
feedback (``send = g + e``, ddp_init.py:155-157) is applied between steps so the later
steps exercise a non-zero memory.
"""
import importlib.util
import os
import sys
import tempfile
import warnings

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REF_DIR = "/root/reference/ddp_powersgd_guide_cifar10"
HERE = os.path.dirname(os.path.abspath(__file__))

SHAPES_A = [(8, 3, 3, 3), (6,), (16, 8), (16,), (1, 5), (5, 1), (12, 20), (3,)]
SHAPES_B = [(2, 24), (2,), (24, 36), (36,), (10, 7, 2), (7,), (33, 5), (1,)]

CASES = {
    # name: (shapes, compression rank, reuse_query, world, steps)
    "w1_r2_reuse": (SHAPES_A, 2, True, 1, 3),
    "w1_r4_fresh": (SHAPES_A, 4, False, 1, 3),
    "w2_r2_reuse": (SHAPES_A, 2, True, 2, 3),
    "w1_r8_clip": (SHAPES_B, 8, True, 1, 2),
    "w2_r3_fresh": (SHAPES_B, 3, False, 2, 2),
}
SEED = 714  # ddp_init.py:23


def load_reference():
    sys.path.insert(0, REF_DIR)
    try:
        spec = importlib.util.spec_from_file_location("_upstream_reducer", os.path.join(REF_DIR, "reducer.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(REF_DIR)
    return mod


def flat(ts):
    return torch.cat([t.reshape(-1) for t in ts]).numpy().copy()


def worker(rank, world, name, init_file, out_dir):
    warnings.filterwarnings("ignore")
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method="file://" + init_file, world_size=world, rank=rank)
    ref = load_reference()
    shapes, crank, reuse, _, steps = CASES[name]

    drawn = []
    orig = ref.PowerSGDReducer._set_random

    def recording(self, vector):
        orig(self, vector)
        drawn.append(vector.detach().reshape(-1).clone())

    ref.PowerSGDReducer._set_random = recording
    red = ref.PowerSGDReducer(random_seed=SEED, device=torch.device("cpu"), n_power_iterations=0,
                              reuse_query=reuse, rank=crank)
    gen = torch.Generator().manual_seed(1000 + rank)
    mem = [torch.zeros(s) for s in shapes]
    out = {}
    for step in range(steps):
        g = [torch.randn(s, generator=gen) for s in shapes]
        send = [a + b for a, b in zip(g, mem)]
        gout = [torch.zeros(s) for s in shapes]
        drawn.clear()
        bits = red.reduce(send, gout, mem)
        out[f"s{step}_in"] = flat(send)
        out[f"s{step}_out"] = flat(gout)
        out[f"s{step}_mem"] = flat(mem)
        out[f"s{step}_p"] = red.p_memory.numpy().copy()
        out[f"s{step}_q"] = red.q_memory.numpy().copy()
        out[f"s{step}_qinit"] = torch.cat(drawn).numpy().copy() if drawn else np.zeros(0, np.float32)
        out[f"s{step}_bits"] = np.int64(bits)
    np.savez(os.path.join(out_dir, f"{name}.rank{rank}.npz"), **out)
    dist.destroy_process_group()


# ---- one case at realistic sizes (world 2, rank 4, one step with a non-zero error memory folded into the input) ----
# Full-size inputs and outputs would be ~160 MB, so the fixture keeps what pins every kernel path and nothing more:
# the inputs are regenerated from their seeds (a checksum guards the regeneration), P and Q are stored in full
# (K1, K2, K3, both all-reduces), grad_out / memory_out as SAMPLED_ROWS seeded rows per matrix plus per-tensor
# Frobenius norms (K4 + K5), the <=1-D outputs in full.
MEDIUM_NAME = "w2_r4_medium"
MEDIUM_SHAPES = [(3072, 768), (3072,), (768, 3072), (768,), (30522, 64), (64,), (64, 3, 7, 7), (2, 768), (2,)]
MEDIUM_RANK, MEDIUM_WORLD, SAMPLED_ROWS = 4, 2, 16


def medium_inputs(rank):
    """The seeded send buffers (g + e) of worker `rank`; used by this script and by the tests that replay the case."""
    gen = torch.Generator().manual_seed(5000 + rank)
    g = [torch.randn(s, generator=gen) for s in MEDIUM_SHAPES]
    e = [0.1 * torch.randn(s, generator=gen) if len(s) > 1 else torch.zeros(s) for s in MEDIUM_SHAPES]
    return [a + b for a, b in zip(g, e)]


def medium_rows(i, n):
    """Row indices sampled from matrix i (seeded, sorted, first and last row always in)."""
    rs = np.random.RandomState(100 + i)
    k = min(n, SAMPLED_ROWS)
    rows = set(rs.choice(n, size=k, replace=False).tolist()) | {0, n - 1}
    return np.array(sorted(rows), np.int64)


def medium_worker(rank, world, init_file, out_dir):
    warnings.filterwarnings("ignore")
    torch.set_num_threads(4)
    dist.init_process_group("gloo", init_method="file://" + init_file, world_size=world, rank=rank)
    ref = load_reference()
    drawn = []
    orig = ref.PowerSGDReducer._set_random

    def recording(self, vector):
        orig(self, vector)
        drawn.append(vector.detach().reshape(-1).clone())

    ref.PowerSGDReducer._set_random = recording
    red = ref.PowerSGDReducer(random_seed=SEED, device=torch.device("cpu"), n_power_iterations=0, reuse_query=True,
                              rank=MEDIUM_RANK)
    send = medium_inputs(rank)
    gout = [torch.zeros(s) for s in MEDIUM_SHAPES]
    mem = [torch.zeros(s) for s in MEDIUM_SHAPES]
    bits = red.reduce(send, gout, mem)
    out = {"bits": np.int64(bits), "p": red.p_memory.numpy().copy(), "q": red.q_memory.numpy().copy(),
           "qinit": torch.cat(drawn).numpy().copy(),
           "in_sumsq": np.array([float((t.double() ** 2).sum()) for t in send]),
           "in_head": np.stack([np.resize(t.reshape(-1)[:4].numpy(), 4) for t in send]),
           "out_norm": np.array([float(t.double().norm()) for t in gout]),
           "mem_norm": np.array([float(t.double().norm()) for t in mem])}
    for i, s in enumerate(MEDIUM_SHAPES):
        if len(s) > 1:
            rows = medium_rows(i, s[0])
            out[f"out_rows{i}"] = gout[i].reshape(s[0], -1)[rows].numpy().copy()
            out[f"mem_rows{i}"] = mem[i].reshape(s[0], -1)[rows].numpy().copy()
        else:
            out[f"out_vec{i}"] = gout[i].numpy().copy()
    np.savez(os.path.join(out_dir, f"medium.rank{rank}.npz"), **out)
    dist.destroy_process_group()


def make_medium():
    with tempfile.TemporaryDirectory() as tmp:
        mp.spawn(medium_worker, args=(MEDIUM_WORLD, os.path.join(tmp, "rdzv"), tmp), nprocs=MEDIUM_WORLD, join=True)
        merged = {"shapes": np.array([",".join(map(str, s)) for s in MEDIUM_SHAPES]), "rank": np.int64(MEDIUM_RANK),
                  "world": np.int64(MEDIUM_WORLD), "seed": np.int64(SEED), "sampled_rows": np.int64(SAMPLED_ROWS)}
        for r in range(MEDIUM_WORLD):
            z = np.load(os.path.join(tmp, f"medium.rank{r}.npz"))
            for k in z.files:
                merged[f"w{r}_{k}"] = z[k]
    np.savez_compressed(os.path.join(HERE, f"{MEDIUM_NAME}.npz"), **merged)
    print("wrote", MEDIUM_NAME)


def main():
    assert os.path.isdir(REF_DIR), "run this where /root/reference is mounted"
    if len(sys.argv) > 1 and sys.argv[1] == "medium":
        make_medium()
        return
    for name, (shapes, crank, reuse, world, steps) in CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            init_file = os.path.join(tmp, "rdzv")
            mp.spawn(worker, args=(world, name, init_file, tmp), nprocs=world, join=True)
            merged = {"shapes": np.array([",".join(map(str, s)) for s in shapes]),
                      "rank": np.int64(crank), "reuse": np.int64(reuse), "world": np.int64(world),
                      "steps": np.int64(steps), "seed": np.int64(SEED)}
            for r in range(world):
                z = np.load(os.path.join(tmp, f"{name}.rank{r}.npz"))
                for k in z.files:
                    merged[f"w{r}_{k}"] = z[k]
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **merged)
        print("wrote", name)

    # known answers for the seed stream (reducer.py:8-12 then :37)
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    rng = np.random.RandomState(SEED)
    rng.randn(128 * 1024 * 1024)
    seeds = [int(rng.randint(1_000_000_000)) for _ in range(8)]
    torch.manual_seed(seeds[0])
    first = torch.randn(147, 2).numpy()
    np.savez_compressed(os.path.join(HERE, "seed_stream.npz"), seed=np.int64(SEED), seeds=np.array(seeds, np.int64),
                        first_randn_147x2=first)
    print("seeds", seeds)
    make_medium()


if __name__ == "__main__":
    main()
