"""The drop-in boundary, demonstrated with the upstream files themselves: the unmodified training scripts
(ddp_powersgd_guide_cifar10/ddp_init.py and ddp_powersgd_distillBERT_IMDb/ddp_init.py: `from reducer import
PowerSGDReducer` at :20, the constructor at :121 / :162, the call at :161 / :205) run `setup(); run_task(); cleanup()`
once with this package's `dropin/` modules first on sys.path and once with the reference's own reducer; the trained
parameters and the communicated-bit sequences must agree.  Needs the reference tree (absent on the GPU box: skipped)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_REF = os.path.isdir("/root/reference/ddp_powersgd_guide_cifar10")


def _start(which, kind, out, device="cpu"):
    return subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "upstream_loop.py"), which, kind, out, device],
                            stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, cwd=ROOT)


def _finish(proc, out):
    stdout, stderr = proc.communicate(timeout=900)
    assert proc.returncode == 0, stdout[-2000:] + stderr[-3000:]
    return np.load(out), stdout


@pytest.mark.skipif(not HAVE_REF, reason="needs /root/reference (build container only)")
@pytest.mark.parametrize("which", ["cifar", "imdb"])
def test_unmodified_upstream_loop_runs_on_the_dropin(which, tmp_path):
    p_ours = _start(which, "dropin", str(tmp_path / "ours.npz"))  # the two runs are independent processes
    p_ref = _start(which, "reference", str(tmp_path / "ref.npz"))
    ours, log = _finish(p_ours, str(tmp_path / "ours.npz"))
    ref, _ = _finish(p_ref, str(tmp_path / "ref.npz"))
    assert "dropin" in str(ours["reducer_file"]) and "/root/reference/" in str(ref["reducer_file"])
    assert "All Task Finished" in log
    assert int(ours["calls"]) == int(ref["calls"]) >= 2
    assert ours["bits"].tolist() == ref["bits"].tolist()
    # same seeds, same Q draws, same side effects on the global generator (DataLoader shuffling follows it): the two
    # runs see the same batches and end with the same parameters up to fp32 rounding of the reducer arithmetic
    a, b = ours["params"].astype(np.float64), ref["params"].astype(np.float64)
    assert np.isfinite(a).all()
    assert np.linalg.norm(a - b) / np.linalg.norm(b) < 1e-5


@pytest.mark.gpu
def test_flat_module_names_resolve_to_the_cuda_build(tmp_path):
    """What `from reducer import PowerSGDReducer` / `import tensor_buffer` give a script whose sys.path starts with
    dropin/: the CUDA-backed classes, working on a GPU."""
    code = r'''
import os, sys, tempfile
sys.path.insert(0, os.path.join(sys.argv[1], "network_distributed_pytorch_b200", "dropin"))
sys.path.insert(1, sys.argv[1])
import torch, torch.distributed as dist
from reducer import PowerSGDReducer, Reducer, orthogonalize, all_reduce, n_bits
import tensor_buffer as tb
import network_distributed_pytorch_b200.reducer as ours
assert PowerSGDReducer is ours.PowerSGDReducer and tb.TensorBuffer.__module__.startswith("network_distributed_pytorch_b200")
dist.init_process_group("gloo", init_method="file://" + os.path.join(tempfile.mkdtemp(), "r"), world_size=1, rank=0)
dev = torch.device("cuda", 0)
shapes = [(96, 64), (64,), (40, 200), (7,)]
g = [torch.randn(s, device=dev) for s in shapes]
out = [torch.zeros(s, device=dev) for s in shapes]
mem = [torch.zeros(s, device=dev) for s in shapes]
red = PowerSGDReducer(random_seed=714, device=dev, n_power_iterations=0, reuse_query=True, rank=2)
bits = red.reduce(g, out, mem)
torch.cuda.synchronize()
assert bits == 32 * ((96 + 40) * 2 + (64 + 200) * 2 + 71)
for a, o, e in zip(g, out, mem):
    if a.ndimension() > 1:
        assert torch.allclose(o + e, a, atol=1e-5)
    else:
        assert torch.equal(o, a)
buf = tb.TensorBuffer([g[1], g[3]])
assert torch.equal(buf.buffer, torch.cat([g[1], g[3]])) and buf.bits() == 32 * 71
print("flat names ok")
'''
    script = tmp_path / "flat.py"
    script.write_text(code)
    proc = subprocess.run([sys.executable, str(script), ROOT], capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0 and "flat names ok" in proc.stdout, proc.stdout[-2000:] + proc.stderr[-3000:]
