#!/usr/bin/env python
"""Benchmark of the PowerSGD reduce hot path (BASELINE.json metric: reduce ms/step and effective
gradient GB/s) on the DistilBERT gradient set, compression rank 4 (BASELINE configs[1]).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N ...            # the reference algorithm on the host CPUs

A "step" is one ``PowerSGDReducer.reduce(grad_in, grad_out, memory_out)`` over the full gradient set
(all ranks: same shapes, rank-specific seeded data; weak scaling).  ``value`` is the whole-job
effective gradient bandwidth = N * 4*(E_hi+E_1) bytes / step time, with the gradients resident in HBM;
``e2e`` is the same quantity with the gradients starting in pinned host memory and grad_out read back,
both copies inside the timed region.  One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "powersgd_reduce_effective_grad_bandwidth"
UNIT = "GB/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="distilbert")
    ap.add_argument("--rank", type=int, default=4, help="PowerSGD compression rank")
    ap.add_argument("--no-reuse-query", action="store_true")
    ap.add_argument("--extra-power-iterations", type=int, default=0)
    ap.add_argument("--e2e-steps", type=int, default=40)  # long enough that pipeline fill and drain (one copy each) are ~2 %
    ap.add_argument("--cpu-reps", type=int, default=5)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--no-phase-timing", action="store_true")
    ap.add_argument("--no-clock-probe", action="store_true", help="profiling runs: skip the 1.5 s clock probe")
    ap.add_argument("--skip-train", action="store_true", help="skip the training samples/sec leg")
    ap.add_argument("--train-steps", type=int, default=8)
    ap.add_argument("--skip-parity", action="store_true", help="N > 1: skip the self-check against the N-worker oracle")
    ap.add_argument("--skip-configs", action="store_true", help="skip the BASELINE configs 1, 3, 4, 5 block")
    ap.add_argument("--skip-baselines", action="store_true",
                    help="skip the torch-on-GPU, gloo world-2 CPU and dense comparison legs")
    return ap.parse_args()


def workload_config(args, st):
    return {
        "workload": f"{args.workload} gradient set ({st['n_mat']} matrices + {st['n_vec']} vectors, "
                    f"{st['grad_bytes'] / 1e6:.1f} MB fp32), PowerSGD rank {args.rank}, "
                    f"reuse_query={int(not args.no_reuse_query)}, power_iterations={args.extra_power_iterations} "
                    "(BASELINE.json configs[1] gradient shapes)",
        "compression_rank": args.rank,
        "grad_bytes": st["grad_bytes"],
        "algorithmic_bytes_per_step": st["algo_bytes"],
        "communicated_bits_per_step": st["bits"],
        "l2": (f"inputs ({st['grad_bytes'] / 1e6:.0f} MB) exceed the 126 MB L2 and every step also writes 2x that, so no flush is needed"
               if st["grad_bytes"] > 126e6 else
               f"the whole set ({st['grad_bytes'] / 1e6:.0f} MB) fits the 126 MB L2 and is NOT flushed between steps: "
               "a roofline fraction above 1 is possible and says so"),
    }


# ------------------------------------------------------------------------------------------------
# Baseline and checker legs.  These are the only places where bench.py touches oracle/ (the CPU restatement of
# the reference, test infrastructure): never inside the timed region of the B200 arm, never as the thing shipped.
# ------------------------------------------------------------------------------------------------
def cpu_reduce_time(args, shapes, reps):
    """One process, all host cores, world 1: the reference algorithm (oracle port) on the full gradient set."""
    import torch

    from oracle.powersgd_oracle import OracleReducer

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    gen = torch.Generator().manual_seed(1000)
    gin = [torch.randn(s, generator=gen) for s in shapes]
    gout = [torch.zeros(s) for s in shapes]
    mem = [torch.zeros(s) for s in shapes]
    orc = OracleReducer(714, "cpu", n_power_iterations=args.extra_power_iterations,
                        reuse_query=not args.no_reuse_query, rank=args.rank)
    orc.reduce(gin, gout, mem)  # warm-up: also draws Q
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        orc.reduce(gin, gout, mem)
        times.append(time.perf_counter() - t0)
    times.sort()
    return times[len(times) // 2], cores


def _gloo_worker(rank, world, workload, crank, reuse, reps, threads, init_file, out_dir):
    """One gloo rank of the CPU baseline: the upstream op stream with real collectives."""
    import warnings

    import torch
    import torch.distributed as dist

    from network_distributed_pytorch_b200 import workloads
    from oracle.reference_stream import StreamReducer

    warnings.filterwarnings("ignore")
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", init_method="file://" + init_file, world_size=world, rank=rank)
    shapes = workloads.WORKLOADS[workload]()
    gen = torch.Generator().manual_seed(1000 + rank)
    gin = [torch.randn(s, generator=gen) for s in shapes]
    gout = [torch.zeros(s) for s in shapes]
    mem = [torch.zeros(s) for s in shapes]
    red = StreamReducer(714, torch.device("cpu"), 0, reuse, crank)
    red.reduce(gin, gout, mem)  # warm-up: first call also draws Q
    times = []
    for _ in range(reps):
        dist.barrier()
        t0 = time.perf_counter()
        red.reduce(gin, gout, mem)
        times.append(time.perf_counter() - t0)
    times.sort()
    with open(os.path.join(out_dir, f"t{rank}.json"), "w") as f:
        json.dump({"median_s": times[len(times) // 2]}, f)
    dist.destroy_process_group()


def gloo_cpu_time(args, world, reps):
    """`world` gloo ranks on this host (mp.spawn + file:// rendezvous), cores // world threads each, every rank its own
    seeded gradient set; returns (max over ranks of the median seconds per reduce(), threads per rank, cores)."""
    import torch.multiprocessing as mp

    cores = os.cpu_count() or 1
    threads = max(1, cores // world)
    with tempfile.TemporaryDirectory() as tmp:
        mp.spawn(_gloo_worker, args=(world, args.workload, args.rank, not args.no_reuse_query, reps, threads,
                                     os.path.join(tmp, "rdzv"), tmp), nprocs=world, join=True)
        med = [json.load(open(os.path.join(tmp, f"t{r}.json")))["median_s"] for r in range(world)]
    return max(med), threads, cores


def run_reference(args):
    """The reference arm: the reference algorithm on the host cores, as many gloo workers as the B200 arm has GPUs
    (each with its own gradient set: the same weak-scaling definition of `value`), launched from rank 0 alone."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from network_distributed_pytorch_b200 import workloads

    shapes = workloads.WORKLOADS[args.workload]()
    st = workloads.stats(shapes, args.rank)
    n = max(1, args.gpus)
    reps = min(max(1, args.steps), 60 if n == 1 else 20)  # bounded: a DistilBERT reduce is 0.05-0.5 s on the host
    if n == 1:
        sec, cores = cpu_reduce_time(args, shapes, reps)
        threads = cores
        sample = (f"full {args.workload} gradient set, median of {reps} reduce() calls of the oracle port (same torch CPU "
                  "ops as upstream reducer.py), 1 process, world 1")
    else:
        if args.extra_power_iterations:
            raise SystemExit("the reference asserts n_power_iterations == 0; the gloo arm has no such variant")
        sec, threads, cores = gloo_cpu_time(args, n, reps)
        sample = (f"full {args.workload} gradient set per worker, {n} gloo workers on this host (mp.spawn, file:// "
                  f"rendezvous), {threads} threads each, max over workers of the median of {reps} reduce() calls of the "
                  "upstream op stream (oracle/reference_stream.py)")
    value = n * st["grad_bytes"] / sec / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n,
        "steps": reps, "warmup": max(1, args.warmup), "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, st),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "workers": n,
                         "threads_per_worker": threads, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


PARITY_SHAPES = [  # the transformer_slice and odd_columns sets of tests/test_gpu_parity.py
    (512, 768), (768,), (768, 768), (768,), (3072, 768), (3072,), (768, 3072), (2, 768), (2,),
    (64, 3, 7, 7), (64,), (33, 27), (5,), (130, 1), (1, 130), (7, 13),
]


def parity_check(dev, world, rank, modes):
    """N > 1, before the timed region: every rank reduces a seeded, rank-specific medium gradient set (two steps with
    error feedback, then one step with an extra power iteration) through each all-reduce mode in `modes`; rank 0
    gathers the inputs and results, runs the N-worker oracle on the CPU with the same query and checks grad_out,
    memory_out, P and Q (<= 1e-5 relative Frobenius per tensor) plus bit identity of P, Q and grad_out across ranks.
    Returns the report dict (rank 0) -- the caller exits non-zero when report["ok"] is false."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from network_distributed_pytorch_b200.reducer import PowerSGDReducer

    TOL = 1e-5
    shapes = PARITY_SHAPES
    numel = [int(np.prod(s)) for s in shapes]

    def flat(ts):
        return torch.cat([t.reshape(-1) for t in ts])

    def gather(t):
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t.contiguous())
        return parts

    def rel(a, b, den=None):
        a, b = a.double(), b.double()
        d = float((a - b).norm())
        return 0.0 if d == 0 else d / max(float((b if den is None else den).norm()), 1e-300)

    report = {"world": world, "tolerance": TOL, "shapes": "transformer_slice + odd_columns, rank 4", "modes": [], "ok": True}
    worst = 0.0
    identical = True
    for mode_name, env in modes:
        saved = {k: os.environ.get(k) for k in env}
        os.environ.update(env)
        try:
            for extra in (0, 1):
                red = PowerSGDReducer(714, dev, 0, True, 4, extra_power_iterations=extra)
                drawn = []
                orig = red._set_random

                def rec(v, orig=orig, drawn=drawn):
                    orig(v)
                    drawn.append(v.detach().reshape(-1).clone())

                red._set_random = rec
                gen = torch.Generator().manual_seed(4000 + 17 * rank + extra)
                mem = [torch.zeros(s, device=dev) for s in shapes]
                orc = None
                o_mem = None
                for step in range(2 if extra == 0 else 1):
                    g = [torch.randn(s, generator=gen).to(dev) for s in shapes]
                    send = [a + b for a, b in zip(g, mem)]
                    out = [torch.zeros(s, device=dev) for s in shapes]
                    drawn.clear()
                    red.reduce(send, out, mem)
                    torch.cuda.synchronize()
                    all_in, all_out, all_mem = gather(flat(send)), gather(flat(out)), gather(flat(mem))
                    all_p, all_q = gather(red.p_memory), gather(red.q_memory)
                    for parts in (all_p, all_q, all_out):
                        identical = identical and all(torch.equal(parts[0], x) for x in parts[1:])
                    err = 0.0
                    if rank == 0:
                        from oracle.powersgd_oracle import OracleReducer

                        if orc is None:
                            orc = OracleReducer(714, "cpu", n_power_iterations=extra, reuse_query=True, rank=4, n_workers=world)
                            o_mem = [[torch.zeros(s) for s in shapes] for _ in range(world)]

                        def split(f):
                            return [x.view(s) for x, s in zip(torch.split(f.cpu(), numel), shapes)]

                        o_in = [split(f) for f in all_in]
                        o_out = [[torch.zeros(s) for s in shapes] for _ in range(world)]
                        orc.reduce(o_in, o_out, o_mem, q_init=torch.cat(drawn).cpu() if drawn else None)
                        err = max(rel(all_p[0].cpu(), orc.p_memory), rel(all_q[0].cpu(), orc.q_memory))
                        for w in range(world):
                            got_o, got_m = split(all_out[w]), split(all_mem[w])
                            for i, s in enumerate(shapes):
                                err = max(err, rel(got_o[i], o_out[w][i]))
                                if len(s) > 1:  # memory of <=1-D tensors is never written (reducer.py:166-168)
                                    small = float(o_mem[w][i].double().norm()) < 1e-3 * float(o_in[w][i].double().norm())
                                    err = max(err, rel(got_m[i], o_mem[w][i], den=o_in[w][i] if small else None))
                        # both trajectories continue from the oracle's memory: every step is a single-step comparison
                        for w in range(world):
                            all_mem[w] = flat(o_mem[w]).to(dev)
                    back = torch.stack(all_mem) if rank == 0 else torch.empty(world, sum(numel), device=dev)
                    dist.broadcast(back, 0)
                    for t, x in zip(mem, torch.split(back[rank], numel)):
                        t.copy_(x.view(t.shape))
                    worst = max(worst, err)
                px = getattr(red, "_peer", None)
                kind = "nccl" if px is None else px.kind
                red._set_random = None
                del red
            report["modes"].append({"mode": mode_name, "allreduce": kind})
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
    flag = torch.tensor([0 if identical else 1], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    report["cross_rank_identical"] = int(flag.item()) == 0
    report["worst_rel_err"] = worst
    report["ok"] = bool(report["cross_rank_identical"] and worst <= TOL)
    okt = torch.tensor([1 if report["ok"] else 0], device=dev)
    dist.broadcast(okt, 0)
    report["ok"] = bool(int(okt.item()))
    return report


def torch_gpu_baseline(args, shapes, dev, world, barrier, st, steps=10):
    """The like-for-like GPU baseline: the reference op stream (one library kernel per layer and op,
    reducer.py:121-168) through stock PyTorch on this same B200 / NCCL.  Device-timed, max over ranks."""
    import torch
    import torch.distributed as dist

    from oracle.reference_stream import StreamReducer

    gen = torch.Generator(device=dev).manual_seed(1000 + (dist.get_rank() if world > 1 else 0))
    gin = [torch.randn(s, device=dev, generator=gen) for s in shapes]
    gout = [torch.zeros(s, device=dev) for s in shapes]
    mem = [torch.zeros(s, device=dev) for s in shapes]
    red = StreamReducer(714, dev, 0, not args.no_reuse_query, args.rank)
    for _ in range(3):
        red.reduce(gin, gout, mem)
    launches = None
    try:  # kernels per step, counted by the profiler on one extra step
        from torch.profiler import ProfilerActivity, profile

        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            red.reduce(gin, gout, mem)
            torch.cuda.synchronize()
        launches = sum(1 for e in prof.events() if str(getattr(e, "device_type", "")).endswith("CUDA")
                       and "memcpy" not in e.name.lower() and "memset" not in e.name.lower())
    except Exception:  # noqa: BLE001 -- a count is a nicety, the timing below is the baseline
        launches = None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        red.reduce(gin, gout, mem)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    del gin, gout, mem, red
    torch.cuda.empty_cache()
    return {"ms_per_step": ms, "value": world * st["grad_bytes"] / (ms * 1e-3) / 1e9, "unit": UNIT, "steps": steps,
            "kernel_launches_per_step": launches,
            "what": "the reference op stream of reducer.py:121-168 (per-layer torch.matmul(out=), eager Gram-Schmidt, "
                    "sub + copy_, torch.cat, torch.distributed all_reduce) as stock PyTorch fp32 ops on the same GPU(s)"}


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t0 or ts > t1 + 0.15:
                continue
            parts = [p.strip() for p in line.split(",")]
            try:
                sm.append(float(parts[0]))
                mx = max(mx, float(parts[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def padded_views(flat, shapes):
    """Views of `flat` with every tensor starting on a 16-byte boundary (identical host/device layout)."""
    views, off = [], 0
    for s in shapes:
        k = 1
        for d in s:
            k *= d
        views.append(flat[off : off + k].view(s))
        off += (k + 3) // 4 * 4
    return views


def padded_total(shapes):
    off = 0
    for s in shapes:
        k = 1
        for d in s:
            k *= d
        off += (k + 3) // 4 * 4
    return off


def peak_gbs():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy)"
    return 6650.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


def stream_reducer_factory(dev, crank):
    from oracle.reference_stream import StreamReducer

    return StreamReducer(714, dev, 0, True, crank)


# BASELINE.json configs other than the headline: (label, workload, rank, reuse_query, extra power iterations)
OTHER_CONFIGS = [
    ("config 1: ResNet-18 r=2", "resnet18", 2, True, 0),
    ("config 3: DistilBERT r=1", "distilbert", 1, True, 0),
    ("config 3: DistilBERT r=2", "distilbert", 2, True, 0),
    ("config 3: DistilBERT r=8", "distilbert", 8, True, 0),
    ("config 3: DistilBERT r=4, reuse_query off", "distilbert", 4, False, 0),
    ("config 3: DistilBERT r=4, 2 power iterations", "distilbert", 4, True, 1),
    ("config 4: ResNet-50 r=4", "resnet50", 4, True, 0),
    ("config 5: BERT-large r=8", "bert-large", 8, True, 0),
]


def run_configs(dev, world, rank, barrier, peak, sampler, min_seconds=0.6):
    """Un-bracketed reduce() ms/step of the other BASELINE.json configurations, each timed with CUDA events over at
    least `min_seconds` of back-to-back steps (long enough for clock samples), max over ranks."""
    import torch
    import torch.distributed as dist

    from network_distributed_pytorch_b200 import workloads
    from network_distributed_pytorch_b200.reducer import PowerSGDReducer

    out = []
    for label, wl, crank, reuse, extra in OTHER_CONFIGS:
        shapes = workloads.WORKLOADS[wl]()
        st = workloads.stats(shapes, crank)
        total = padded_total(shapes)
        gen = torch.Generator(device=dev).manual_seed(1000 + rank)
        in_flat = torch.randn(total, device=dev, generator=gen)
        out_flat, mem_flat = torch.zeros(total, device=dev), torch.zeros(total, device=dev)
        gin, gout, mem = padded_views(in_flat, shapes), padded_views(out_flat, shapes), padded_views(mem_flat, shapes)
        red = PowerSGDReducer(714, dev, 0, reuse, crank, extra_power_iterations=extra)
        for _ in range(3):
            red.reduce(gin, gout, mem)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(5):
            red.reduce(gin, gout, mem)
        e1.record()
        barrier()
        est = max(e0.elapsed_time(e1) / 5, 1e-3)
        steps = int(min(4000, max(30, min_seconds * 1e3 / est)))
        tsteps = torch.tensor([steps], device=dev)
        if world > 1:
            dist.all_reduce(tsteps, op=dist.ReduceOp.MAX)
        steps = int(tsteps.item())
        barrier()
        w0 = time.time()
        e0.record()
        for _ in range(steps):
            red.reduce(gin, gout, mem)
        e1.record()
        barrier()
        w1 = time.time()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms.item())
        # algorithmic bytes: the P / Q phases (K1, K2, K3 and their buffers) repeat per power iteration
        algo = st["algo_bytes"] + extra * (st["k1_bytes"] + st["k2_bytes"] + st["k3_bytes"])
        clocks = sampler.summary(w0, w1) if sampler else None
        out.append({"config": label, "workload": wl, "compression_rank": crank, "reuse_query": reuse,
                    "power_iterations": 1 + extra, "ms_per_step": ms, "steps": steps,
                    "value": world * st["grad_bytes"] / (ms * 1e-3) / 1e9, "unit": UNIT,
                    "algorithmic_bytes_per_step": algo, "hbm_roofline_frac_whole_step": algo / (ms * 1e-3) / 1e9 / peak,
                    "grad_bytes": st["grad_bytes"], "l2_resident": st["grad_bytes"] < 100e6, "clocks": clocks})
        del red, gin, gout, mem, in_flat, out_flat, mem_flat
        torch.cuda.empty_cache()
    return out


def dense_baseline(shapes, dev, world, barrier, st, steps=20):
    """Uncompressed averaging (ddp_guide_cifar10/ddp_init.py:57-62) as one gather, ONE all-reduce, one scatter."""
    import torch
    import torch.distributed as dist

    from network_distributed_pytorch_b200.dense import DenseReducer

    gen = torch.Generator(device=dev).manual_seed(1000)
    gin = [torch.randn(s, device=dev, generator=gen) for s in shapes]
    gout = [torch.zeros(s, device=dev) for s in shapes]
    red = DenseReducer(None, dev)
    for _ in range(3):
        bits = red.reduce(gin, gout)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        red.reduce(gin, gout)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    del gin, gout, red
    torch.cuda.empty_cache()
    return {"ms_per_step": ms, "value": world * st["grad_bytes"] / (ms * 1e-3) / 1e9, "unit": UNIT, "steps": steps,
            "communicated_bits_per_step": int(bits),
            "what": "DenseReducer: all gradients gathered into one flat buffer, one SUM all-reduce + divide (skipped at "
                    "world 1, as upstream), scattered back"}


def run_b200(args):
    import torch
    import torch.distributed as dist

    from network_distributed_pytorch_b200 import _lib, workloads
    from network_distributed_pytorch_b200.reducer import PowerSGDReducer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    else:
        f = tempfile.NamedTemporaryFile(delete=False)
        f.close()
        os.unlink(f.name)
        dist.init_process_group("gloo", init_method="file://" + f.name, world_size=1, rank=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- N > 1: numerical self-check of the multi-GPU path, before and outside the timed region ------------------
    parity = None
    if world > 1 and not args.skip_parity:
        modes = [("auto (the mode the timed region uses)", {})]
        if world == 2:  # cheap at two ranks: every all-reduce implementation the package has
            modes += [("nccl", {"PSGD_PEER_ALLREDUCE": "0"}),
                      ("fused-push", {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "push"}),
                      ("fused-multimem", {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "multimem"}),
                      ("fused-pull", {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "pull"}),
                      ("kernel-multicast", {"PSGD_PEER_ALLREDUCE": "1", "PSGD_PEER_MODE": "kernel"})]
        parity = parity_check(dev, world, rank, modes)
        barrier()
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "n_gpus": world, "parity": parity,
                                  "error": "multi-GPU parity check failed; nothing was timed"}), flush=True)
            dist.destroy_process_group()
            raise SystemExit(3)

    shapes = workloads.WORKLOADS[args.workload]()
    st = workloads.stats(shapes, args.rank)
    total = padded_total(shapes)
    gen = torch.Generator(device=dev).manual_seed(1000 + rank)
    in_flat = torch.randn(total, device=dev, generator=gen)
    out_flat = torch.zeros(total, device=dev)
    mem_flat = torch.zeros(total, device=dev)
    gin, gout, mem = padded_views(in_flat, shapes), padded_views(out_flat, shapes), padded_views(mem_flat, shapes)

    red = PowerSGDReducer(714, dev, 0, not args.no_reuse_query, args.rank,
                          extra_power_iterations=args.extra_power_iterations)
    lib = _lib.load()

    W = max(3, args.warmup)
    K = max(1, args.steps)
    for _ in range(W):
        red.reduce(gin, gout, mem)
    plan = next(iter(red._plans.values()))
    if not args.no_phase_timing:
        # only the dominant kernel is bracketed inside the timed region: every bracketed phase costs a
        # few microseconds of stream time (all seven together: ~13 % of the step)
        plan.timing_enable(["finalize"])
        plan.timing_read()
    bracket_every = 0 if args.no_phase_timing else 4

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.25)

    # ---- timed region: K steps, device events on the launch stream, max over ranks --------------
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = lib.psgd_launch_count()
    t_wall0 = time.time()
    ev0.record()
    h0 = time.perf_counter()
    for i in range(K):
        if bracket_every:
            # an event pair between two launches keeps the second from starting under the first's tail
            # (programmatic dependent launch), so only every fourth finalize launch is bracketed
            plan.timing_enable(["finalize"] if i % bracket_every == 0 else False)
        red.reduce(gin, gout, mem)
    host_us = (time.perf_counter() - h0) / K * 1e6  # host time to issue one step (no sync inside)
    ev1.record()
    barrier()
    t_wall1 = time.time()
    launches = lib.psgd_launch_count() - launches0
    px = getattr(red, "_peer", None)
    allreduce_kind = "none (1 GPU)" if world == 1 else "torch.distributed NCCL" if px is None else px.describe()
    ms = ev0.elapsed_time(ev1)
    phases = plan.timing_read() if not args.no_phase_timing else {}
    plan.timing_enable(False)
    all_phases = {}
    if not args.no_phase_timing:  # diagnostic breakdown, outside the timed region
        plan.timing_enable(True)
        for _ in range(min(K, 20)):
            red.reduce(gin, gout, mem)
        torch.cuda.synchronize()
        all_phases = plan.timing_read()
        plan.timing_enable(False)
    tmax = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())

    # ---- clocks: if the timed region was too short for a sample, probe under the same load -------
    clocks = None
    if sampler:
        clocks = sampler.summary(t_wall0, t_wall1)
        clocks["window"] = "timed region"
    if not args.no_clock_probe and (world > 1 or (clocks is not None and clocks["samples"] < 3)):
        p0 = time.time()
        while time.time() - p0 < 1.5:
            for _ in range(20):
                red.reduce(gin, gout, mem)
            torch.cuda.synchronize()
        p1 = time.time()
        if sampler and clocks["samples"] < 3:
            clocks = sampler.summary(p0 + 0.3, p1)
            clocks["window"] = "1.5 s probe of the same steps right after the timed region"
    barrier()

    # ---- diagnostic: the same step with the caller's error-feedback add fused (reduce_fused) ----------
    fused_ms = None
    if not args.no_phase_timing:
        send = padded_views(torch.zeros(total, device=dev), shapes)
        for _ in range(3):
            red.reduce_fused(gin, mem, send)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        f0.record()
        for _ in range(min(K, 20)):
            red.reduce_fused(gin, mem, send)
        f1.record()
        barrier()
        fused_ms = f0.elapsed_time(f1) / min(K, 20)
        # restore the state the later legs expect (gin was overwritten with decompressed gradients)
        in_flat.normal_(generator=gen)
        mem_flat.zero_()
        barrier()

    # ---- host cost of the slow path: grad_out at new addresses every step (zero_grad() re-allocates .grad upstream,
    # ddp_init.py:144), i.e. validation, list splitting, 3 x 104 data_ptr() calls and a pointer-table upload per step ----
    fresh = None
    if not args.no_phase_timing:
        out_b = torch.zeros(total, device=dev)
        gouts2 = [gout, padded_views(out_b, shapes)]
        for k in range(4):
            red.reduce(gin, gouts2[k & 1], mem)
        n_f = min(K, 20)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        f0.record()
        h0 = time.perf_counter()
        for k in range(n_f):
            red.reduce(gin, gouts2[k & 1], mem)
        fresh_host_us = (time.perf_counter() - h0) / n_f * 1e6
        f1.record()
        barrier()
        fresh = {"host_issue_us_per_step": round(fresh_host_us, 1), "ms_per_step": f0.elapsed_time(f1) / n_f,
                 "note": "grad_out alternates between two allocations, so every call takes the full host path "
                         "(validation, split, pointer table upload)"}
        del out_b, gouts2
        barrier()

    # ---- the same step replayed from a CUDA graph (PowerSGDReducer.capture): host time per step ----------
    graph_leg = None
    if world == 1 and not args.no_phase_timing and args.extra_power_iterations >= 0 and not args.no_reuse_query:
        try:
            cap = red.capture(gin, gout, mem)
            for _ in range(3):
                cap.replay()
            n_g = min(K, 50)
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            g0.record()
            h0 = time.perf_counter()
            for _ in range(n_g):
                cap.replay()
            g_host = (time.perf_counter() - h0) / n_g * 1e6
            g1.record()
            barrier()
            graph_leg = {"ms_per_step": g0.elapsed_time(g1) / n_g, "host_issue_us_per_step": round(g_host, 1), "steps": n_g,
                         "what": "PowerSGDReducer.capture(): the four kernels of a step with their programmatic dependencies "
                                 "as one CUDA graph, replayed"}
            del cap
        except Exception as exc:  # noqa: BLE001 -- a diagnostic leg
            graph_leg = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}
        in_flat.normal_(generator=gen)
        mem_flat.zero_()
        barrier()

    # ---- e2e: gradients start in pinned host memory, grad_out is read back ------------------------
    # Three streams and two device buffer sets: the H2D copy of step k+1 and the D2H copy of step k-1 run on
    # the two copy engines while step k reduces, exactly as a host-resident caller would pipeline the
    # public reduce() call.  Every step still copies all of its inputs in and its result out.
    e2e = None
    if not args.skip_e2e:
        host_in = torch.empty(total, dtype=torch.float32, pin_memory=True)
        host_out = torch.empty(total, dtype=torch.float32, pin_memory=True)
        host_in.copy_(in_flat)
        ins = [in_flat, torch.empty_like(in_flat)]
        outs = [out_flat, torch.empty_like(out_flat)]
        gins = [gin, padded_views(ins[1], shapes)]
        gouts = [gout, padded_views(outs[1], shapes)]
        s_h2d, s_d2h = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        s_main = torch.cuda.current_stream(dev)
        ev_in = [torch.cuda.Event() for _ in range(2)]     # H2D of buffer b finished
        ev_red = [torch.cuda.Event() for _ in range(2)]    # reduce() on buffer b finished
        ev_out = [torch.cuda.Event() for _ in range(2)]    # D2H of buffer b finished

        def e2e_steps(n):
            for k in range(n):
                b = k & 1
                with torch.cuda.stream(s_h2d):
                    if k >= 2:
                        s_h2d.wait_event(ev_red[b])        # reduce k-2 no longer reads in[b]
                    ins[b].copy_(host_in, non_blocking=True)   # H2D of this step's gradients
                    ev_in[b].record(s_h2d)
                s_main.wait_event(ev_in[b])
                if k >= 2:
                    s_main.wait_event(ev_out[b])           # D2H k-2 no longer reads out[b]
                red.reduce(gins[b], gouts[b], mem)         # public API call
                ev_red[b].record(s_main)
                with torch.cuda.stream(s_d2h):
                    s_d2h.wait_event(ev_red[b])
                    host_out.copy_(outs[b], non_blocking=True)  # D2H of the step's result (grad_out)
                    ev_out[b].record(s_d2h)
            s_main.wait_stream(s_d2h)
            s_main.wait_stream(s_h2d)

        E = max(2, args.e2e_steps)
        e2e_steps(2)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        e2e_steps(E)
        e1.record()
        barrier()
        ems = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
        ems = float(ems.item()) / E

        # what the link alone allows: both copies of one step running concurrently, nothing else
        def duplex_ms(reps=4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(dev)
            a.record(s_main)
            for _ in range(reps):
                s_h2d.wait_stream(s_main)
                s_d2h.wait_stream(s_main)
                with torch.cuda.stream(s_h2d):
                    ins[1].copy_(host_in, non_blocking=True)
                with torch.cuda.stream(s_d2h):
                    host_out.copy_(outs[1], non_blocking=True)
                s_main.wait_stream(s_h2d)
                s_main.wait_stream(s_d2h)
            b.record(s_main)
            torch.cuda.synchronize(dev)
            return a.elapsed_time(b) / reps

        link_ms = duplex_ms()
        e2e = {"value": world * st["grad_bytes"] / (ems * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ems,
               "steps": E, "h2d_bytes_per_step": total * 4, "d2h_bytes_per_step": total * 4,
               "path": "per step: pinned host -> HBM copy of all gradients, PowerSGDReducer.reduce(), grad_out HBM -> "
                       "pinned host; copies of neighbouring steps overlap on the two copy engines",
               "link_only_ms_per_step": link_ms,
               "link_note": "the same two copies (H2D + D2H of one step, concurrently) with no kernel running: the floor "
                            "the host link sets for ms_per_step"}
    del in_flat, out_flat, mem_flat, gin, gout, mem
    if e2e is not None:
        del ins, outs, gins, gouts, host_in, host_out
    torch.cuda.empty_cache()

    # ---- the other BASELINE.json configurations: un-bracketed step time + clocks, same rules as the headline ------
    configs = None
    if not args.skip_configs:
        configs = run_configs(dev, world, rank, barrier, peak_gbs()[0], sampler)

    # ---- comparison points on the same GPU(s): the reference op stream through stock PyTorch, and dense averaging ----
    torch_base, dense = None, None
    if not args.skip_baselines:
        if args.extra_power_iterations == 0:
            try:
                torch_base = torch_gpu_baseline(args, shapes, dev, world, barrier, st)
            except Exception as exc:  # noqa: BLE001 -- a baseline must not take the headline down
                torch_base = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}
        try:
            dense = dense_baseline(shapes, dev, world, barrier, st)
        except Exception as exc:  # noqa: BLE001
            dense = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}

    # ---- training samples/sec (BASELINE.json metric, second half): the upstream loop on synthetic tokens ----
    training = None
    if not args.skip_train and args.workload in ("distilbert", "bert-large", "resnet18", "resnet50"):
        try:
            from network_distributed_pytorch_b200 import harness

            training = {}
            seq = 512 if args.workload in ("distilbert", "bert-large") else 224
            bsz = 16 if args.workload in ("distilbert", "bert-large") else 256
            legs = [("upstream", "upstream", lambda: PowerSGDReducer(714, dev, 0, True, args.rank)),
                    ("fused", "fused", lambda: PowerSGDReducer(714, dev, 0, True, args.rank))]
            if not args.skip_baselines:
                legs.append(("torch_gpu_baseline", "upstream", lambda: stream_reducer_factory(dev, args.rank)))
            for name, loop, factory in legs:
                res = harness.measure_samples_per_sec(
                    args.workload, factory, dev, batch=bsz, seq_len=seq,
                    steps=args.train_steps, warmup=3, loop=loop, world=world, barrier=barrier)
                tms = torch.tensor([res["ms_per_step"]], device=dev)
                if world > 1:
                    dist.all_reduce(tms, op=dist.ReduceOp.MAX)
                res["ms_per_step"] = float(tms.item())
                res["samples_per_sec"] = world * bsz / (res["ms_per_step"] * 1e-3)
                training[name] = res
            training["note"] = ("random-init model of the named architecture, random tokens, bf16 autocast forward/backward, "
                                "fp32 parameters and reducer; 'upstream' = the reference loop with this reducer dropped in, "
                                "'fused' = same arithmetic through reduce_fused + FusedMomentumSGD, 'torch_gpu_baseline' = "
                                "the reference loop with the reference reducer's own op stream (stock PyTorch on the same GPU)")
        except Exception as exc:  # the headline metric must not depend on this leg
            training = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}

    if sampler:
        sampler.stop()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ms_per_step = ms / K
    value = world * st["grad_bytes"] / (ms_per_step * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (finalize: reads M, P, Q; writes out and mem) -----------
    peak, peak_src = peak_gbs()
    roofline, phase_report = None, {}
    algo = {"pack": st["e_1"] * 8, "p_from_m": st["k1_bytes"], "orthogonalize": st["k2_bytes"],
            "q_from_mt": st["k3_bytes"], "q_reduce": 8 * st["e_q"], "finalize": st["k4_bytes"], "finish": st["e_1"] * 8}
    kernel_ms = 0.0
    for name, (msum, cnt) in all_phases.items():
        if cnt:
            avg = msum / cnt
            kernel_ms += avg
            phase_report[name] = {"ms": round(avg, 5), "GBps": round(algo[name] / (avg * 1e-3) / 1e9, 1)}
    if phases.get("finalize", (0, 0))[1]:
        avg = phases["finalize"][0] / phases["finalize"][1]
        achieved = st["k4_bytes"] / (avg * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get(f"k_finalize:{args.workload}:r{args.rank}")
        roofline = {"bound": "hbm", "kernel": "k_finalize (out = P Q^T, mem = M - out)", "achieved": achieved,
                    "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                    "algorithmic_bytes_per_launch": st["k4_bytes"], "ms_per_launch": avg, "peak_source": peak_src,
                    "bracketed_launches": int(phases["finalize"][1]),
                    "bracketing": "CUDA events inside the library around every 4th finalize launch of the timed region",
                    "whole_step": {"ms_per_step": ms_per_step,
                                   "achieved": st["algo_bytes"] / (ms_per_step * 1e-3) / 1e9,
                                   "frac": st["algo_bytes"] / (ms_per_step * 1e-3) / 1e9 / peak,
                                   "note": "all kernels + launch gaps of one reduce(), algorithmic bytes of SURVEY 8(d)"}}

    cpu, cpu_gloo2 = None, None
    if world == 1 and not args.skip_cpu_baseline:
        sec, cores = cpu_reduce_time(args, shapes, max(1, args.cpu_reps))
        cpu = {"value": st["grad_bytes"] / sec / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
               "ms_per_step": sec * 1e3,
               "sample": f"full {args.workload} gradient set, median of {max(1, args.cpu_reps)} reduce() calls of the "
                         "oracle port (same torch CPU ops as upstream reducer.py)"}
        if not args.skip_baselines and args.extra_power_iterations == 0:
            try:  # the configured CPU baseline of BASELINE.md section 4: gloo, world 2, cores / 2 threads per rank
                sec2, threads, cores = gloo_cpu_time(args, 2, max(1, args.cpu_reps))
                cpu_gloo2 = {"value": 2 * st["grad_bytes"] / sec2 / 1e9, "unit": UNIT, "ms_per_step": sec2 * 1e3,
                             "workers": 2, "threads_per_worker": threads, "cores": cores, "kind": "port",
                             "sample": f"full {args.workload} gradient set per worker, 2 gloo workers (mp.spawn, file:// "
                                       f"rendezvous), max over workers of the median of {max(1, args.cpu_reps)} reduce() "
                                       "calls of the upstream op stream (oracle/reference_stream.py); value = both "
                                       "workers' gradient bytes / that time"}
            except Exception as exc:  # noqa: BLE001
                cpu_gloo2 = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, st),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
        "allreduce": allreduce_kind, "parity": parity,
        "torch_gpu_baseline": torch_base, "cpu_baseline_gloo2": cpu_gloo2, "dense": dense, "configs": configs,
        "host_issue_us_per_step_fresh_ptrs": fresh, "graph_replay": graph_leg,
        "clocks": clocks, "phases": phase_report,
        "phases_note": "per-phase event brackets measured in a separate pass (they add ~13 % to the step and keep the kernels from overlapping head to tail); "
                       "inside the timed region only finalize is bracketed, on every 4th step",
        "training": training,
        "reduce_fused_ms_per_step": fused_ms,
        "reduce_fused_note": "reduce_fused(grads, memories, send): ddp_init.py:155-157 (send = g + e, 2 launches per "
                             "parameter upstream) folded into the P = M Q kernel; 2 more HBM streams than reduce()",
        "host_issue_us_per_step": round(host_us, 1),
        "algorithmic_GBps": world * st["algo_bytes"] / (ms_per_step * 1e-3) / 1e9,
        "hbm_roofline_frac_whole_step": st["algo_bytes"] / (ms_per_step * 1e-3) / 1e9 / peak,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
