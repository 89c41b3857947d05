"""Two real GPUs over NCCL (skipped on a single-GPU box): the CUDA reducer on two ranks must reproduce what two
ranks of the upstream reducer produced (golden world-2 fixtures), and both ranks must end with identical P, Q
and grad_out."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["PSGD_ROOT"])
from tests.helpers import Golden, flatten, per_tensor_mem_err, per_tensor_rel_err, rel_err, unflatten
from network_distributed_pytorch_b200.reducer import PowerSGDReducer
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
TOL = 1e-5
red = None
for name in ("w2_r2_reuse", "w2_r3_fresh", "w2_r2_reuse"):
    if red is not None:  # let go of the previous reducer, so that its symmetric buffers and signal pads get recycled
        import gc
        peer_seen = red._peer is not None
        red._set_random = None
        red = None
        gc.collect()
        torch.cuda.synchronize()
    g = Golden(name)
    red = PowerSGDReducer(g.seed, dev, 0, g.reuse, g.rank)
    assert red.n_workers == 2
    state = {"buf": None, "off": 0}
    def fake(vector, red=red, state=state):
        red.rng.randint(1_000_000_000)
        k = vector.nelement()
        vector.data[:] = state["buf"][state["off"]:state["off"] + k].view(vector.shape)
        state["off"] += k
    red._set_random = fake
    mem = [torch.zeros(s, device=dev) for s in g.shapes]
    for step in range(g.steps):
        state["buf"] = torch.from_numpy(g.get(0, step, "qinit")).to(dev); state["off"] = 0
        gin = unflatten(g.get(rank, step, "in"), g.shapes, device=dev)
        gout = [torch.zeros(s, device=dev) for s in g.shapes]
        bits = red.reduce(gin, gout, mem)
        torch.cuda.synchronize()
        assert bits == int(g.get(0, step, "bits"))
        assert rel_err(red.p_memory.cpu().numpy(), g.get(0, step, "p")) < TOL
        assert rel_err(red.q_memory.cpu().numpy(), g.get(0, step, "q")) < TOL
        assert max(per_tensor_rel_err(flatten(gout), g.get(rank, step, "out"), g.shapes)) < TOL
        assert max(per_tensor_mem_err(flatten(mem), g.get(rank, step, "mem"), g.get(rank, step, "in"), g.shapes)) < TOL
        # identical across ranks after the all-reduces
        for t in (red.p_memory, red.q_memory, torch.cat([x.reshape(-1) for x in gout])):
            both = [torch.empty_like(t) for _ in range(2)]
            dist.all_gather(both, t.contiguous())
            assert torch.equal(both[0], both[1])
        for t, ref in zip(mem, unflatten(g.get(rank, step, "mem"), g.shapes, device=dev)):
            t.copy_(ref)
# one extra power iteration (upstream asserts 0): the all-reduces run twice per step, the flat buffer travels once
from oracle.powersgd_oracle import OracleReducer
from tests.gpu_common import _last_q_init
shapes = [(96, 64), (64,), (40, 200), (7,), (300, 3072), (5,)]
gen = torch.Generator().manual_seed(5)
ins = [[torch.randn(s, generator=gen) for s in shapes] for _ in range(2)]
red._set_random = None
red = PowerSGDReducer(714, dev, 0, True, 4, extra_power_iterations=1)
gin = [t.to(dev) for t in ins[rank]]
gout = [torch.zeros(s, device=dev) for s in shapes]
mem = [torch.zeros(s, device=dev) for s in shapes]
red.reduce(gin, gout, mem)
torch.cuda.synchronize()
orc = OracleReducer(714, "cpu", n_power_iterations=1, reuse_query=True, rank=4, n_workers=2)
o_out = [[torch.zeros(s) for s in shapes] for _ in range(2)]
o_mem = [[torch.zeros(s) for s in shapes] for _ in range(2)]
orc.reduce(ins, o_out, o_mem, q_init=_last_q_init(red, shapes, 4, 0, 714))
assert rel_err(red.p_memory.cpu().numpy(), orc.p_memory.numpy()) < TOL
assert rel_err(red.q_memory.cpu().numpy(), orc.q_memory.numpy()) < TOL
assert max(per_tensor_rel_err(flatten(gout), flatten(o_out[rank]), shapes)) < TOL
assert max(per_tensor_mem_err(flatten(mem), flatten(o_mem[rank]), flatten(ins[rank]), shapes)) < TOL
# the reference-generated case at realistic sizes (tests/golden/w2_r4_medium.npz): streamed kernels, cluster Gram-Schmidt
from tests.helpers import MediumGolden
mg = MediumGolden()
mine = mg.inputs(rank)
if mine is not None:
    red._set_random = None
    red = PowerSGDReducer(mg.seed, dev, 0, True, mg.rank)
    qbuf = {"buf": torch.from_numpy(mg.get(0, "qinit")).to(dev), "off": 0}
    def fake_q(vector, red=red, st=qbuf):
        red.rng.randint(1_000_000_000)
        k = vector.nelement()
        vector.data[:] = st["buf"][st["off"]:st["off"] + k].view(vector.shape)
        st["off"] += k
    red._set_random = fake_q
    gin = [t.to(dev) for t in mine]
    gout = [torch.zeros(s, device=dev) for s in mg.shapes]
    mem = [torch.zeros(s, device=dev) for s in mg.shapes]
    bits = red.reduce(gin, gout, mem)
    torch.cuda.synchronize()
    assert bits == int(mg.get(0, "bits"))
    assert rel_err(red.p_memory.cpu().numpy(), mg.get(0, "p")) < TOL
    assert rel_err(red.q_memory.cpu().numpy(), mg.get(0, "q")) < TOL
    mg.check_outputs(rank, gout, mem, TOL)
    for t in (red.p_memory, red.q_memory, torch.cat([x.reshape(-1) for x in gout])):
        both = [torch.empty_like(t) for _ in range(2)]
        dist.all_gather(both, t.contiguous())
        assert torch.equal(both[0], both[1])
    red._set_random = None
dist.destroy_process_group()
print("rank", rank, "ok", "peer=%s" % (red._peer is not None), "kind=%s" % (red._peer.kind if red._peer is not None else "nccl"))
'''


MODES = {  # name -> environment
    "fused-push": {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "push"},
    "fused-pull": {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "pull"},
    "fused-multimem": {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "multimem"},
    "kernel-multicast": {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "kernel"},
    "kernel-pull": {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "kernel", "PSGD_PEER_MULTICAST": "0"},
    "nccl": {"PSGD_PEER_ALLREDUCE": "0"},
    "default": {},
}


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
# the all-reduces taken into the kernels (two ways of summing), the stand-alone kernels, torch.distributed (NCCL), and
# whatever the package picks when nothing is set
@pytest.mark.parametrize("peer", list(MODES))
def test_two_gpus_match_two_rank_reference(tmp_path, peer):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, PSGD_ROOT=ROOT, PSGD_DEBUG="1")
    for k in ("PSGD_PEER_ALLREDUCE", "PSGD_PEER_MODE", "PSGD_PEER_MULTICAST"):
        env.pop(k, None)
    env.update(MODES[peer])
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                         env=env, capture_output=True, text=True, timeout=420)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert out.stdout.count(" ok ") == 2, out.stdout[-2000:]
    print(out.stdout[-400:])
    if peer == "nccl":
        assert "kind=nccl" in out.stdout
    elif peer != "default":
        assert "peer=True" in out.stdout, out.stdout[-2000:]
        if ("kind=" + peer) not in out.stdout:
            pytest.skip("no multicast mapping on this fabric: the pull form ran instead")
