"""The oracle is test infrastructure: nothing under the package may import it, and bench.py may only reach it from
its baseline legs and its multi-GPU self-check.  The product path has no CPU fallback either."""
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


# bench.py may execute oracle/ only as a baseline (CPU legs, the reference op stream on the GPU) or as the checker
# of the multi-GPU self-test -- each a function of its own, none of them inside the timed region of the B200 arm
BENCH_ORACLE_SITES = {"cpu_reduce_time", "_gloo_worker", "parity_check", "torch_gpu_baseline", "stream_reducer_factory"}


def test_bench_reaches_the_oracle_only_from_its_baseline_and_checker_legs():
    import ast

    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    sites = set()

    def visit(node, fn):
        for child in ast.iter_child_nodes(node):
            name = child.name if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef)) else fn
            if isinstance(child, ast.ImportFrom) and (child.module or "").split(".")[0] == "oracle":
                sites.add(fn)
            if isinstance(child, ast.Import) and any(a.name.split(".")[0] == "oracle" for a in child.names):
                sites.add(fn)
            # nested helpers count as their top-level function
            visit(child, name if fn is None else fn)

    visit(tree, None)
    assert sites, "bench.py is expected to time the oracle port as its CPU baseline"
    assert sites <= BENCH_ORACLE_SITES, sites
    # the timed region of the B200 arm lies between the two event records of run_b200; none of the sites is called there
    src = open(os.path.join(ROOT, "bench.py")).read()
    body = src[src.index("def run_b200("):]
    timed = body[body.index("ev0.record()"):body.index("ev1.record()")]
    assert not any(site + "(" in timed for site in BENCH_ORACLE_SITES)


def test_engine_refuses_cpu_devices():
    import pytest
    import torch

    from network_distributed_pytorch_b200.engine import CudaEngine

    with pytest.raises(RuntimeError):
        CudaEngine(torch.device("cpu"))
