import os, sys, time
sys.path.insert(0, os.getcwd())
import torch
from tests.gpu_common import ensure_group
from network_distributed_pytorch_b200 import workloads
from network_distributed_pytorch_b200.reducer import PowerSGDReducer
from bench import padded_total, padded_views
ensure_group()
dev = torch.device("cuda", 0)
for wl, rank in (("distilbert", 4), ("resnet18", 2), ("resnet50", 4)):
    shapes = workloads.WORKLOADS[wl]()
    total = padded_total(shapes)
    gin = padded_views(torch.randn(total, device=dev), shapes)
    gout = padded_views(torch.zeros(total, device=dev), shapes)
    mem = padded_views(torch.zeros(total, device=dev), shapes)
    red = PowerSGDReducer(714, dev, 0, True, rank)
    for _ in range(3):
        red.reduce(gin, gout, mem)
    torch.cuda.synchronize()
    ref_out = [t.clone() for t in gout]; ref_mem = [t.clone() for t in mem]; ref_q = red.q_memory.clone()
    # eager timing
    def timeit(fn, n=200):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); h0 = time.perf_counter(); e0.record()
        for _ in range(n): fn()
        host = (time.perf_counter() - h0) / n * 1e6
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3, host
    t_eager = timeit(lambda: red.reduce(gin, gout, mem))
    try:
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            red.reduce(gin, gout, mem)
        torch.cuda.current_stream().wait_stream(s)
        with torch.cuda.graph(g):
            red.reduce(gin, gout, mem)
        t_graph = timeit(lambda: g.replay())
        print(wl, rank, "eager us/step (gpu, host):", [round(x, 1) for x in t_eager], "graph:", [round(x, 1) for x in t_graph])
    except Exception as exc:
        print(wl, "capture failed:", repr(exc)[:500])
