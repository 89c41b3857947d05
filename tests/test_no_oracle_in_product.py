"""The oracle is test infrastructure: nothing under the package may import it, and bench.py may only reach it from
its CPU legs (cpu_baseline / --impl reference).  The product path has no CPU fallback either."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "network_distributed_pytorch_b200")


def _py_files(top):
    for base, _, files in os.walk(top):
        for f in files:
            if f.endswith(".py"):
                yield os.path.join(base, f)


def test_package_never_imports_the_oracle():
    pat = re.compile(r"^\s*(from\s+oracle\b|import\s+oracle\b)", re.M)
    offenders = [p for p in _py_files(PKG) if pat.search(open(p).read())]
    assert offenders == []


def test_bench_reaches_the_oracle_only_from_its_cpu_legs():
    src = open(os.path.join(ROOT, "bench.py")).read()
    uses = [m.start() for m in re.finditer(r"\boracle\b\s*(\.|import)|from\s+oracle", src)]
    assert uses, "bench.py is expected to time the oracle port as its CPU baseline"
    # every use sits inside cpu_reduce_time(), the one function both CPU legs call
    start = src.index("def cpu_reduce_time(")
    end = src.index("\ndef ", start + 1)
    assert all(start < u < end for u in uses)


def test_engine_refuses_cpu_devices():
    import pytest
    import torch

    from network_distributed_pytorch_b200.engine import CudaEngine

    with pytest.raises(RuntimeError):
        CudaEngine(torch.device("cpu"))
