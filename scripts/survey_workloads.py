"""dev tool (GPU box): reduce() step time and per-phase kernel times for the BASELINE.json configurations."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from network_distributed_pytorch_b200 import workloads
from scripts.sweep import measure
from tests.gpu_common import ensure_group

ensure_group()
peak = 6555.2
for wl, rank in [("resnet18", 2), ("distilbert", 1), ("distilbert", 2), ("distilbert", 4), ("distilbert", 8),
                 ("distilbert", 16), ("resnet50", 4), ("bert-large", 8), ("resnet152", 4)]:
    shapes = workloads.WORKLOADS[wl]()
    st = workloads.stats(shapes, rank)
    ms, ph = measure(shapes, rank, {}, steps=10)
    torch.cuda.empty_cache()
    print(f"{wl:11s} r={rank:2d}  step {ms*1e3:8.1f} us  algo {st['algo_bytes']/1e6:8.1f} MB  "
          f"{st['algo_bytes']/ms/1e6:7.0f} GB/s ({st['algo_bytes']/ms/1e6/peak:4.2f} of peak, all phases bracketed)  "
          + " ".join(f"{k}={v*1e3:.0f}" for k, v in ph.items()), flush=True)
