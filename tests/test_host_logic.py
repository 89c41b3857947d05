"""Host-side logic of the reducer mirror, exercised on CPU through the oracle-backed test double:
list splitting, packed layout, seed stream, folded divides, bit accounting, error behaviour."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist

from network_distributed_pytorch_b200 import reducer as R
from tests.fake_engine import OracleEngine
from tests.helpers import GOLDEN_CASES, Golden, flatten, rel_err, unflatten


@pytest.fixture(scope="module", autouse=True)
def single_process_group():
    if not dist.is_initialized():
        f = tempfile.NamedTemporaryFile(delete=False)
        f.close()
        os.unlink(f.name)
        dist.init_process_group("gloo", init_method="file://" + f.name, world_size=1, rank=0)
    yield
    if dist.is_initialized():
        dist.destroy_process_group()


def make(rank=2, reuse=True, **kw):
    return R.PowerSGDReducer(random_seed=714, device=torch.device("cpu"), n_power_iterations=0, reuse_query=reuse,
                             rank=rank, _engine=OracleEngine(), **kw)


@pytest.mark.parametrize("name", [c for c in GOLDEN_CASES if c.startswith("w1")])
def test_world1_matches_reference_vectors(name):
    g = Golden(name)
    torch.set_num_threads(1)
    red = make(rank=g.rank, reuse=g.reuse)
    assert red.p_memory is None and red.q_memory is None
    mem = [torch.zeros(s) for s in g.shapes]
    for step in range(g.steps):
        gin = unflatten(g.get(0, step, "in"), g.shapes)
        gout = [torch.zeros(s) for s in g.shapes]
        bits = red.reduce(gin, gout, mem)
        assert bits == int(g.get(0, step, "bits"))
        # the packed buffers must have the reference layout bit for bit
        assert np.array_equal(red.p_memory.numpy(), g.get(0, step, "p"))
        assert np.array_equal(red.q_memory.numpy(), g.get(0, step, "q"))
        assert rel_err(flatten(gout), g.get(0, step, "out")) < 1e-6
        assert rel_err(flatten(mem), g.get(0, step, "mem")) < 1e-6


def test_attributes_and_seed_side_effect():
    red = make(rank=3)
    assert red.rank == 3 and red.n_workers == 1 and red.reuse_query is True
    assert red.device == torch.device("cpu")
    shapes = [(6, 5), (4,)]
    gin = [torch.ones(s) for s in shapes]
    red.reduce(gin, [torch.zeros(s) for s in shapes], [torch.zeros(s) for s in shapes])
    # reducer.py:37 leaves the global generator seeded with the last Q seed
    a = torch.randn(3)
    torch.manual_seed(663468714)
    torch.randn(5, 3)
    b = torch.randn(3)
    assert torch.equal(a, b)


def test_phase_order_and_plan_cache():
    eng = OracleEngine()
    red = R.PowerSGDReducer(714, torch.device("cpu"), rank=2, _engine=eng)
    shapes = [(6, 5), (4,), (3, 2, 2)]
    for _ in range(2):
        gin = [torch.randn(s) for s in shapes]
        red.reduce(gin, [torch.zeros(s) for s in shapes], [torch.zeros(s) for s in shapes])
    assert len(eng.plans) == 1  # same shapes -> same plan
    one = ["set_ptrs", "pack_vec", "p_from_m", "orthogonalize", "q_from_mt", "finalize", "finish"]
    assert eng.plans[0].calls == one + one


def test_memory_of_vectors_is_never_written():
    red = make()
    shapes = [(6, 5), (4,)]
    gin = [torch.randn(s) for s in shapes]
    mem = [torch.full(s, 7.0) for s in shapes]
    out = [torch.zeros(s) for s in shapes]
    red.reduce(gin, out, mem)
    assert torch.equal(mem[1], torch.full((4,), 7.0))  # reducer.py:166-168 only touches `out`
    assert torch.equal(out[1], gin[1])
    assert torch.allclose(out[0] + mem[0], gin[0], atol=1e-6)


def test_error_behaviour():
    with pytest.raises(AssertionError):  # reducer.py:30
        R.PowerSGDReducer(714, torch.device("cpu"), n_power_iterations=1, _engine=OracleEngine())
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):  # product engine refuses CPU
        R.PowerSGDReducer(714, torch.device("cpu"))
    red = make()
    with pytest.raises(RuntimeError, match="contiguous"):
        t = torch.randn(5, 6).t()
        red.reduce([t, torch.zeros(3)], [torch.zeros(6, 5), torch.zeros(3)], [torch.zeros(6, 5), torch.zeros(3)])
    with pytest.raises(RuntimeError, match="float32"):
        red.reduce([torch.zeros(2, 2, dtype=torch.float64), torch.zeros(3)], [torch.zeros(2, 2), torch.zeros(3)],
                   [torch.zeros(2, 2), torch.zeros(3)])
    with pytest.raises(RuntimeError, match="non-empty"):  # tensor_buffer.py:19 via reducer.py:131
        red.reduce([torch.zeros(2, 2)], [torch.zeros(2, 2)], [torch.zeros(2, 2)])
    with pytest.raises(NotImplementedError):
        R.Reducer(714, torch.device("cpu"), _engine=OracleEngine()).reduce([], [], [])


def test_extra_power_iterations_match_oracle():
    from oracle.powersgd_oracle import OracleReducer

    shapes = [(12, 20), (5,)]
    torch.manual_seed(3)
    gin = [torch.randn(s) for s in shapes]
    red = make(rank=2, extra_power_iterations=1)
    o1, m1 = [torch.zeros(s) for s in shapes], [torch.zeros(s) for s in shapes]
    b1 = red.reduce(gin, o1, m1)
    orc = OracleReducer(714, rank=2, n_power_iterations=1)
    o2, m2 = [torch.zeros(s) for s in shapes], [torch.zeros(s) for s in shapes]
    b2 = orc.reduce(gin, o2, m2)
    assert b1 == b2
    assert rel_err(flatten(o1), flatten(o2)) < 1e-6 and rel_err(flatten(m1), flatten(m2)) < 1e-6


def test_reduce_fused_equals_add_then_reduce():
    torch.manual_seed(11)
    shapes = [(12, 20), (5,), (6, 2, 3), (7,)]
    g = [torch.randn(s) for s in shapes]
    e = [torch.randn(s) for s in shapes]
    # upstream order: send = g + e ; reduce(send, grads, memories)
    a = make(rank=2)
    send_a = [x + y for x, y in zip(g, e)]
    ga, ea = [t.clone() for t in g], [t.clone() for t in e]
    bits_a = a.reduce(send_a, ga, ea)
    b = make(rank=2)
    gb, eb, send_b = [t.clone() for t in g], [t.clone() for t in e], [torch.zeros(s) for s in shapes]
    bits_b = b.reduce_fused(gb, eb, send_b)
    assert bits_a == bits_b
    for x, y in zip(ga + ea + send_a, gb + eb + send_b):
        assert torch.equal(x, y)


def test_module_helpers():
    assert R.n_bits(torch.zeros(10)) == 320
    assert R.all_reduce(torch.zeros(1)) is None  # world 1: skipped (reducer.py:193-195)
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        R.orthogonalize(torch.randn(4, 2))


def test_peer_allreduce_selection():
    """PSGD_PEER_ALLREDUCE: forced on / off, otherwise on from two ranks (inside one node, up to MAX_WORLD) (peer.py)."""
    from network_distributed_pytorch_b200 import peer

    assert not peer.enabled_for(1, "auto") and peer.enabled_for(2, "auto") and peer.enabled_for(8, "auto")
    assert not peer.enabled_for(8, "0") and peer.enabled_for(2, "1")
    assert not peer.enabled_for(peer.MAX_WORLD + 1, "1")
