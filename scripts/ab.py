"""A/B of plan-time switches (GPU box): unbracketed step time of reduce() per workload and setting.
Usage: python scripts/ab.py "PSGD_PDL=0" "PSGD_PDL=1 PSGD_ORTH_MERGE=0" ... [--wl distilbert:4,resnet50:4]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bench import padded_total, padded_views  # noqa: E402
from network_distributed_pytorch_b200 import workloads  # noqa: E402
from network_distributed_pytorch_b200.reducer import PowerSGDReducer  # noqa: E402
from tests.gpu_common import ensure_group  # noqa: E402


FRESH = False  # --fresh: rewrite the gradients between steps (as backward does), time each reduce() on its own


def measure(shapes, rank, env, steps=50, reps=3):
    for k, v in env.items():
        os.environ[k] = str(v)
    dev = torch.device("cuda", 0)
    total = padded_total(shapes)
    gen = torch.Generator(device=dev).manual_seed(1000)
    flat = torch.randn(total, device=dev, generator=gen)
    gin = padded_views(flat, shapes)
    master = flat.clone() if FRESH else None
    gout = padded_views(torch.zeros(total, device=dev), shapes)
    mem = padded_views(torch.zeros(total, device=dev), shapes)
    red = PowerSGDReducer(714, dev, 0, True, rank)
    for _ in range(5):
        red.reduce(gin, gout, mem)
    best = 1e9
    for _ in range(reps if FRESH else 0):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        torch.cuda.synchronize()
        for a, b in evs:
            flat.copy_(master)  # what the last step left in L2 is of no use; the tail of this write is what backward leaves
            a.record()
            red.reduce(gin, gout, mem)
            b.record()
        torch.cuda.synchronize()
        best = min(best, sum(a.elapsed_time(b) for a, b in evs) / steps)
    for _ in range(0 if FRESH else reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            red.reduce(gin, gout, mem)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / steps)
    chk = float(sum(o.double().sum() for o in gout[:8]))
    for k in env:
        os.environ.pop(k, None)
    return best, chk


def main():
    ensure_group()
    args = sys.argv[1:]
    global FRESH
    if "--fresh" in args:
        FRESH = True
        args.remove("--fresh")
    wls = "distilbert:4,distilbert:8,distilbert:16,resnet50:4,bert-large:8"
    if "--wl" in args:
        i = args.index("--wl")
        wls = args[i + 1]
        del args[i:i + 2]
    settings = [dict(kv.split("=") for kv in a.split()) for a in args] or [{}]
    for item in wls.split(","):
        wl, rank = item.split(":")
        shapes = workloads.WORKLOADS[wl]()
        for env in settings:
            ms, chk = measure(shapes, int(rank), env)
            print(f"{wl:12s} r={rank:>2s} {str(env):60s} {ms * 1e3:8.1f} us  checksum {chk:.9e}", flush=True)


if __name__ == "__main__":
    main()
