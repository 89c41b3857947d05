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
        "l2": "inputs (268 MB) exceed the 126 MB L2 and every step also writes 2x that, so no flush is needed",
    }


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm (oracle port; /root/reference is Python and cannot travel)
# ------------------------------------------------------------------------------------------------
def cpu_reduce_time(args, shapes, reps):
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


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from network_distributed_pytorch_b200 import workloads

    shapes = workloads.WORKLOADS[args.workload]()
    st = workloads.stats(shapes, args.rank)
    reps = max(1, args.steps)
    # bounded: a DistilBERT reduce is ~0.1-0.3 s on the host cores, cap the run at ~60 reps
    reps = min(reps, 60)
    # warm-up inside cpu_reduce_time is one call (it includes the Q draw); extra warm-ups as requested
    sec, cores = cpu_reduce_time(args, shapes, reps)
    value = st["grad_bytes"] / sec / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": reps, "warmup": max(1, args.warmup), "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, st),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"full {args.workload} gradient set, median of {reps} reduce() calls of the "
                                   "oracle port (same torch CPU ops as upstream reducer.py), 1 process"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    allreduce_kind = ("none (1 GPU)" if world == 1 else "torch.distributed NCCL" if px is None else
                      "own kernel, NVSwitch multicast (multimem)" if px.multicast else "own kernel, peer pull")
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
    # ---- training samples/sec (BASELINE.json metric, second half): the upstream loop on synthetic tokens ----
    training = None
    if not args.skip_train and args.workload in ("distilbert", "bert-large", "resnet18", "resnet50"):
        try:
            from network_distributed_pytorch_b200 import harness

            del in_flat, out_flat, mem_flat, gin, gout, mem
            torch.cuda.empty_cache()
            training = {}
            seq = 512 if args.workload in ("distilbert", "bert-large") else 224
            bsz = 16 if args.workload in ("distilbert", "bert-large") else 256
            for loop in ("upstream", "fused"):
                res = harness.measure_samples_per_sec(
                    args.workload, lambda: PowerSGDReducer(714, dev, 0, True, args.rank), dev, batch=bsz, seq_len=seq,
                    steps=args.train_steps, warmup=3, loop=loop, world=world, barrier=barrier)
                tms = torch.tensor([res["ms_per_step"]], device=dev)
                if world > 1:
                    dist.all_reduce(tms, op=dist.ReduceOp.MAX)
                res["ms_per_step"] = float(tms.item())
                res["samples_per_sec"] = world * bsz / (res["ms_per_step"] * 1e-3)
                training[loop] = res
            training["note"] = ("random-init model of the named architecture, random tokens, bf16 autocast forward/backward, "
                                "fp32 parameters and reducer; 'upstream' = the reference loop with this reducer dropped in, "
                                "'fused' = same arithmetic through reduce_fused + FusedMomentumSGD")
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
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy)"
    else:
        peak, peak_src = 6650.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"
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

    cpu = None
    if world == 1 and not args.skip_cpu_baseline:
        sec, cores = cpu_reduce_time(args, shapes, max(1, args.cpu_reps))
        cpu = {"value": st["grad_bytes"] / sec / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
               "ms_per_step": sec * 1e3,
               "sample": f"full {args.workload} gradient set, median of {max(1, args.cpu_reps)} reduce() calls of the "
                         "oracle port (same torch CPU ops as upstream reducer.py)"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, st),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
        "allreduce": allreduce_kind,
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
