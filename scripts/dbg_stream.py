"""dev tool (GPU box): per-CTA cycle breakdown of the streamed P = M Q kernel."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["PSGD_STREAM_DEBUG"] = "1"
import numpy as np, torch
from bench import padded_total, padded_views
from network_distributed_pytorch_b200 import workloads
from network_distributed_pytorch_b200.reducer import PowerSGDReducer
from tests.gpu_common import ensure_group
ensure_group()
wl = sys.argv[1] if len(sys.argv) > 1 else "distilbert"
rank = int(sys.argv[2]) if len(sys.argv) > 2 else 4
shapes = workloads.WORKLOADS[wl]()
dev = torch.device("cuda", 0)
total = padded_total(shapes)
gin = padded_views(torch.randn(total, device=dev), shapes)
gout = padded_views(torch.zeros(total, device=dev), shapes)
mem = padded_views(torch.zeros(total, device=dev), shapes)
red = PowerSGDReducer(714, dev, 0, True, rank)
for _ in range(5):
    red.reduce(gin, gout, mem)
torch.cuda.synchronize()
plan = next(iter(red._plans.values()))
buf = (ctypes.c_uint64 * (8 * 1024))()
n = plan._lib.psgd_debug_read(plan._h, buf, 8 * 1024)
a = np.array(buf[:n], dtype=np.uint64).reshape(-1, 8).astype(np.float64)
a = a[a[:, 0] > 0]
t0 = a[:, 0].min()
print("CTAs", len(a))
print("start spread ns: max-min", a[:, 0].max() - t0)
print("end   ns rel: min %.0f mean %.0f max %.0f" % ((a[:, 1] - t0).min(), (a[:, 1] - t0).mean(), (a[:, 1] - t0).max()))
for k, name in ((5, "total cycles"), (2, "desc wait"), (3, "data wait"), (4, "Q stage"), (6, "units"), (7, "work cycles")):
    print("%-12s min %9.0f mean %9.0f max %9.0f" % (name, a[:, k].min(), a[:, k].mean(), a[:, k].max()))
