"""Dev tool (library built with -DPSGD_ORTH_TS): where K2's time goes -- globaltimer stamps of every CTA."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from bench import padded_total, padded_views  # noqa: E402
from network_distributed_pytorch_b200 import _lib, workloads  # noqa: E402
from network_distributed_pytorch_b200.reducer import PowerSGDReducer  # noqa: E402
from tests.gpu_common import ensure_group  # noqa: E402

ensure_group()
wl, rank = (sys.argv[1] if len(sys.argv) > 1 else "distilbert:4").split(":")
shapes = workloads.WORKLOADS[wl]()
dev = torch.device("cuda", 0)
total = padded_total(shapes)
gin = padded_views(torch.randn(total, device=dev), shapes)
gout = padded_views(torch.zeros(total, device=dev), shapes)
mem = padded_views(torch.zeros(total, device=dev), shapes)
red = PowerSGDReducer(714, dev, 0, True, int(rank))
lib = ctypes.CDLL(_lib.load()._name)
for _ in range(20):
    red.reduce(gin, gout, mem)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (64 * 12))()
assert lib.psgd_debug_orth_ts(buf) == 0
ts = np.array(buf, dtype=np.uint64).reshape(64, 12).astype(np.int64)
used = ts[:, 0] > 0
t0 = ts[used, 0].min()
np.set_printoptions(linewidth=200)
print("columns: entry, after dependency wait, rows loaded, steps 0..R-1 ..., stored (ns after the first CTA's entry)")
for b in np.nonzero(used)[0]:
    row = ts[b]
    print(b, [int(x - t0) if x > 0 else -1 for x in row])
