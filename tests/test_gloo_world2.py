"""The N>1 host path on CPU: two gloo ranks run the reducer mirror (oracle-backed test double as
the engine) and must reproduce what two ranks of the real upstream reducer produced."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.helpers import Golden, flatten, rel_err, unflatten


def _worker(rank, world, name, init_file, out_dir):
    import warnings

    warnings.filterwarnings("ignore")
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method="file://" + init_file, world_size=world, rank=rank)
    from network_distributed_pytorch_b200.reducer import PowerSGDReducer
    from tests.fake_engine import OracleEngine

    g = Golden(name)
    red = PowerSGDReducer(g.seed, torch.device("cpu"), 0, g.reuse, g.rank, _engine=OracleEngine())
    assert red.n_workers == world
    mem = [torch.zeros(s) for s in g.shapes]
    res = {}
    for step in range(g.steps):
        gin = unflatten(g.get(rank, step, "in"), g.shapes)
        gout = [torch.zeros(s) for s in g.shapes]
        res[f"bits{step}"] = np.int64(red.reduce(gin, gout, mem))
        res[f"out{step}"] = flatten(gout)
        res[f"mem{step}"] = flatten(mem)
        res[f"p{step}"] = red.p_memory.numpy().copy()
        res[f"q{step}"] = red.q_memory.numpy().copy()
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), **res)
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["w2_r2_reuse", "w2_r3_fresh"])
def test_two_ranks_match_reference_vectors(name):
    g = Golden(name)
    with tempfile.TemporaryDirectory() as tmp:
        mp.spawn(_worker, args=(2, name, os.path.join(tmp, "rdzv"), tmp), nprocs=2, join=True)
        r = [np.load(os.path.join(tmp, f"r{k}.npz")) for k in range(2)]
    for step in range(g.steps):
        assert int(r[0][f"bits{step}"]) == int(r[1][f"bits{step}"]) == int(g.get(0, step, "bits"))
        # after the all-reduces every rank holds the same P, Q and grad_out
        for key in ("p", "q", "out"):
            assert np.array_equal(r[0][f"{key}{step}"], r[1][f"{key}{step}"])
        assert rel_err(r[0][f"p{step}"], g.get(0, step, "p")) < 1e-6
        assert rel_err(r[0][f"q{step}"], g.get(0, step, "q")) < 1e-6
        for k in range(2):
            assert rel_err(r[k][f"out{step}"], g.get(k, step, "out")) < 1e-6
            assert rel_err(r[k][f"mem{step}"], g.get(k, step, "mem")) < 1e-6
