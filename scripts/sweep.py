"""Tuning sweep (GPU box): per-phase kernel times of reduce() for several tile-size settings.
Usage: python scripts/sweep.py [workload] [rank] > gpurun_out/sweep.txt"""
import itertools
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bench import padded_total, padded_views  # noqa: E402
from network_distributed_pytorch_b200 import workloads  # noqa: E402
from network_distributed_pytorch_b200.reducer import PowerSGDReducer  # noqa: E402
from tests.gpu_common import ensure_group  # noqa: E402


def measure(shapes, rank, env, steps=20):
    for k, v in env.items():
        os.environ[k] = str(v)
    dev = torch.device("cuda", 0)
    total = padded_total(shapes)
    gen = torch.Generator(device=dev).manual_seed(1000)
    gin = padded_views(torch.randn(total, device=dev, generator=gen), shapes)
    gout = padded_views(torch.zeros(total, device=dev), shapes)
    mem = padded_views(torch.zeros(total, device=dev), shapes)
    red = PowerSGDReducer(714, dev, 0, True, rank)
    for _ in range(3):
        red.reduce(gin, gout, mem)
    plan = next(iter(red._plans.values()))
    plan.timing_enable(True)
    plan.timing_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        red.reduce(gin, gout, mem)
    e1.record()
    torch.cuda.synchronize()
    ph = plan.timing_read()
    for k in env:
        os.environ.pop(k, None)
    return e0.elapsed_time(e1) / steps, {k: v[0] / max(1, v[1]) for k, v in ph.items()}


def main():
    ensure_group()
    wl = sys.argv[1] if len(sys.argv) > 1 else "distilbert"
    rank = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    shapes = workloads.WORKLOADS[wl]()
    print("workload", wl, "rank", rank)
    base = measure(shapes, rank, {})
    print("default", round(base[0], 4), {k: round(v * 1e3, 1) for k, v in base[1].items()})
    sweeps = (("PSGD_STREAM", [0]),
              ("PSGD_SLOT_BYTES", [12288, 24576, 36864, 49152, 61440]),
              ("PSGD_GSS_DIV", [2, 4]),
              ("PSGD_T3_BYTES", [131072, 262144, 524288]),
              ("PSGD_T3_MAXBLK", [24, 32, 48]))
    if len(sys.argv) > 3:
        sweeps = tuple((kv.split("=")[0], [int(x) for x in kv.split("=")[1].split(",")]) for kv in sys.argv[3:])
    for name, vals in sweeps:
        for v in vals:
            ms, ph = measure(shapes, rank, {name: v})
            print(f"{name}={v}", round(ms, 4), {k: round(x * 1e3, 1) for k, x in ph.items()})


if __name__ == "__main__":
    main()
