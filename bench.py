#!/usr/bin/env python
"""bench.py -- DDIM sample-steps/s of the B200-native denoiser path (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A *step* is one DDIM step (denoiser forward + fused update, functions/denoising.py:21-50) over one batch.
Workload (config.workload): BASELINE.json configs[2]/[3] -- audio.yml sampling shape x[64, 2, 8192, 256] PER GPU
(weak scaling: config 4's 512 samples over 8 GPUs is 64 per rank), eta = 0, uniform 1000-step schedule, synthetic
N(0,1) input and seeded random-init weights (norm.2.weight randomised so the residual blocks are live).

  value         sample-steps/s with xt resident in HBM (CUDA events, barrier + synchronize both sides, max over ranks)
  e2e           the same through the public API ``generalized_steps`` with a pinned HOST x per call and both snapshot
                lists read back to the host every step (select_index=None, as the reference does)
  roofline      the kernel class with the largest share of the timed region: algorithmic bytes (or FLOPs) / CUDA-event
                time on the launching stream (library profiler), against MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (port of the reference path) on this box's host cores, bounded sample (rank 0, N=1 only)

``--impl reference`` times the reference's CPU implementation of the path instead (the oracle port: the reference is
Python/PyTorch and /root/reference does not travel to the GPU box), all host threads, bounded sample per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "ddim_sample_steps_per_sec"
UNIT = "sample-steps/s"
F_SIZE = 256


_JSON_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner, written with the C runtime
    before NCCL_DEBUG_FILE is honoured), so file descriptor 1 is re-pointed at stderr for the whole process and the real stdout
    is kept aside for ``emit``."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="samples per GPU (default: 64, BASELINE configs[2]; --mode train: 32, configs[1])")
    ap.add_argument("--t-size", type=int, default=None, help="time frames (default: audio.yml sampling.t_size 8192; --mode train: model.t_size 1024)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=0, help="sample groups of the pipelined host path (0 = the API's default)")
    ap.add_argument("--cpu-sample-steps", type=int, default=2)
    ap.add_argument("--mode", default="sample", choices=["sample", "train", "parity"],
                    help="sample: DDIM sample-steps/s (the headline, BASELINE configs[2]); train: noise_estimation_loss fwd+bwd samples/s "
                         "(configs[1]/[4]); parity: configs[0] -- 50 DDIM steps, batch 4, T=1024 on the GPU and on the CPU oracle, "
                         "per-step eps and final-sample error")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --batch samples per GPU (the default line); strong: --total-samples split over the ranks (BASELINE configs[3]: "
                         "512 samples -> 256 / 128 / 64 per rank at N = 2 / 4 / 8), shards that do not fit are walked as sample groups")
    ap.add_argument("--total-samples", type=int, default=512)
    ap.add_argument("--graph", default="auto", choices=["auto", "on", "off"], help="CUDA-graph replay of the forward in the device-timed loop (auto: small groups only)")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity / batch-position objects of the default line")
    ap.add_argument("--no-small-batch", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the bounded train leg of the default line")
    ap.add_argument("--no-train-graph", action="store_true", help="train leg: issue forward + backward eagerly instead of replaying the captured step")
    ap.add_argument("--train-batch", type=int, default=32)
    ap.add_argument("--train-t-size", type=int, default=1024)
    ap.add_argument("--train-steps", type=int, default=5)
    ap.add_argument("--train-warmup", type=int, default=3)
    ap.add_argument("--parity-steps", type=int, default=50)
    ap.add_argument("--parity-batch", type=int, default=4)
    ap.add_argument("--parity-t-size", type=int, default=1024)
    args = ap.parse_args()
    train = args.mode == "train" and args.impl != "reference"
    if args.t_size is None:
        args.t_size = 1024 if train else 8192
    if args.batch is None:
        args.batch = 32 if train else 64
    return args


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tensor=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured (MEASURED_PEAKS.json; bf16 sustained)")
    return dict(hbm=6650.0, tensor=1400.0, source="fallback (B200_PROFILING.md)")


def workload_name(batch, t_size):
    return (f"audio.yml DDIM sampling step, x[{batch},2,{t_size},{F_SIZE}] per GPU, eta=0, uniform 1000-step schedule "
            f"(BASELINE configs[2]/[3] shape)")


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms while the timed region runs (B200_PROFILING.md clocks line)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [c.strip() for c in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ synthetic model / schedule
def alphas_cumprod():
    """runners/diffusion.py:46-49,109-115: linear betas in float64, cast to fp32, cumprod in fp32."""
    import numpy as np
    import torch
    betas = torch.from_numpy(np.linspace(1e-4, 2e-2, 1000, dtype=np.float64)).float()
    return (1.0 - betas).cumprod(dim=0)


def synthetic_state_dict():
    """Reference default init under seed 0, then every norm.2.weight <- 0.5 randn (seed 1) so the residual blocks are not
    the identity.  CPU fp32 state_dict with the reference's 389 keys."""
    import torch
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    torch.manual_seed(0)
    m = Model(audio_config("torch.FloatTensor"))
    g = torch.Generator().manual_seed(1)
    sd = m.state_dict()
    for k, v in sd.items():
        if k.endswith("norm.2.weight"):
            v.copy_(0.5 * torch.randn(v.shape, generator=g))
    return sd


# ------------------------------------------------------------------------------------------------ CPU legs (oracle = checker / baseline only)
class _Recorder:
    """Wraps the oracle model: keeps every (xt, t, eps) of the run so the CUDA denoiser can be evaluated teacher-forced."""

    def __init__(self, inner, keep):
        self.inner, self.keep, self.calls = inner, keep, []

    def __call__(self, x, t):
        et = self.inner(x, t)
        if self.keep:
            self.calls.append((x.clone(), t.clone(), et))
        return et


def cpu_steps(sd, t_size, nsteps, warm=True, record=False):
    """nsteps DDIM steps of the oracle on x[1,2,t_size,256] with every host thread; returns (seconds, cores, record) where
    record = None or dict(x, seq, calls=[(xt, t, eps)], xs, x0) of the timed run (the parity object's reference side)."""
    import torch
    from oracle import ddim_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = _Recorder(O.OracleModel(sd), record)
    al = alphas_cumprod()
    g = torch.Generator().manual_seed(1234)
    if warm:
        O.generalized_steps(torch.randn(1, 2, 32, F_SIZE, generator=g), [999], O.OracleModel(sd), al, set(), eta=0.0)
    x = torch.randn(1, 2, t_size, F_SIZE, generator=g)
    seq = list(range(0, 1000, 1000 // 50))[-nsteps:] if nsteps <= 50 else list(range(1000))[-nsteps:]
    t0 = time.perf_counter()
    xs, x0 = O.generalized_steps(x, seq, model, al, None if record else set(), eta=0.0)
    dt = time.perf_counter() - t0
    rec = dict(x=x, seq=seq, calls=model.calls, xs=xs, x0=x0) if record else None
    return dt, cores, rec


def oracle_state_dict():
    """Weights for the CPU legs that must not load the product library (``--impl reference``): the oracle's own seeded synthetic
    state_dict (same 389 keys / shapes, live norm.2.weight).  CPU time per step does not depend on the weight values."""
    from oracle import ddim_oracle as O
    return O.make_state_dict(seed=0)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sd = oracle_state_dict()
    assert "ddim_audio_b200" not in sys.modules, "the reference arm must not load the product package"
    for _ in range(min(args.warmup, 1)):
        cpu_steps(sd, min(args.t_size, 1024), 1)
    per = []
    for _ in range(args.steps):
        dt, cores, _ = cpu_steps(sd, args.t_size, 1, warm=False)
        per.append(dt)
    total = sum(per)
    value = args.steps / total
    sample = f"x[1,2,{args.t_size},{F_SIZE}] (1 of the workload's {args.batch} samples per GPU), 1 DDIM step per bench step, fp32, torch CPU ops"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": workload_name(args.batch, args.t_size), "batch_per_gpu": args.batch, "t_size": args.t_size,
                                        "f_size": F_SIZE, "reference_arm": "CPU oracle port of the reference path (the Python reference does not travel to the GPU box)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def roofline_from_profile(prof, K):
    """Per-kernel-class table + the dominant class's roofline object from the library profiler's records."""
    pk = peaks()
    tot_ms = sum(p["ms"] for p in prof) or 1.0
    kern = []
    t_roof = 0.0
    for p in sorted(prof, key=lambda q: -q["ms"]):
        tb = p["bytes"] / (pk["hbm"] * 1e9) * 1e3
        tf = p["flops"] / (pk["tensor"] * 1e12) * 1e3
        bound = "hbm" if tb >= tf else "tensor"
        t_roof += max(tb, tf)
        kern.append({"name": p["name"], "share": round(p["ms"] / tot_ms, 4), "ms_per_call": round(p["ms"] / p["calls"], 5), "calls_per_step": p["calls"] / K,
                     "bound": bound, "frac": round(max(tb, tf) / p["ms"], 4) if p["ms"] > 0 else None,
                     "gbs": round(p["bytes"] / p["ms"] * 1e-6, 1), "tflops": round(p["flops"] / p["ms"] * 1e-9, 1)})
    top = sorted(prof, key=lambda q: -q["ms"])[0]
    tb = top["bytes"] / (pk["hbm"] * 1e9)
    tf = top["flops"] / (pk["tensor"] * 1e12)
    traffic = None
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        traffic = json.load(open(tr_path)).get(top["name"])
    if tb >= tf:
        ach = top["bytes"] / (top["ms"] * 1e-3) * 1e-9
        roof = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"]}
    else:
        ach = top["flops"] / (top["ms"] * 1e-3) * 1e-12
        roof = {"bound": "tensor", "achieved": ach, "peak": pk["tensor"], "unit": "TFLOP/s", "frac": ach / pk["tensor"]}
    roof.update({"traffic": traffic, "kernel": top["name"], "share_of_step": top["ms"] / tot_ms, "launch_ms": top["ms"] / max(top["launches"], 1),
                 "algorithmic_bytes_per_launch": top["bytes"] / max(top["launches"], 1), "algorithmic_flops_per_launch": top["flops"] / max(top["launches"], 1),
                 "peak_source": pk["source"], "timed": "CUDA events on the launching stream around every launch of the timed region"})
    return roof, kern, t_roof


# ------------------------------------------------------------------------------------------------ the B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 arm has no CPU fallback (use --impl reference for the CPU baseline)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        if os.environ.get("DDIMB_NO_NUMA_BIND") is None:
            from ddim_audio_b200.runners.dist import bind_to_gpu_numa
            ok = bind_to_gpu_numa(local)                           # pinned host buffers land on the GPU's own NUMA node
            print(f"[bench] rank {rank}: NUMA binding {'on' if ok else 'unavailable'} ({len(os.sched_getaffinity(0))} cores)", file=sys.stderr)
        dist.init_process_group("nccl", device_id=dev)

    from ddim_audio_b200 import _C, ops
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.functions.denoising import ddim_loop_device, generalized_steps
    from ddim_audio_b200.models.diffusion import Model

    sd = synthetic_state_dict()
    model = Model(audio_config("torch.FloatTensor")).to(dev)
    model.load_state_dict(sd, strict=True)
    model.eval()
    al = alphas_cumprod()
    alpha_list = [1.0] + al.numpy().tolist()
    B, T = args.batch, args.t_size
    if args.scaling == "strong":
        lo = rank * (args.total_samples // world) + min(rank, args.total_samples % world)
        B = args.total_samples // world + (1 if rank < args.total_samples % world else 0)
    K, W = args.steps, max(args.warmup, 0)
    g = torch.Generator().manual_seed(1234 + rank)
    x_host = torch.randn(B, 2, T, F_SIZE, generator=g).pin_memory()
    seq_all = list(range(1000))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world > 1:
            tns = torch.tensor([v], dtype=torch.float64, device=dev)
            dist.all_reduce(tns, op=dist.ReduceOp.MAX)
            return float(tns.item())
        return v

    # ---------------- value: device-resident loop
    gmode = {"auto": None, "on": True, "off": False}[args.graph]
    if args.scaling == "strong":
        return run_strong(args, model, x_host, al, dev, world, rank, local, B, T, K, W, barrier, max_over_ranks)
    torch.cuda.reset_peak_memory_stats(dev)
    xt = x_host.to(dev, non_blocking=True)
    with torch.no_grad():
        if W:
            ddim_loop_device(xt, seq_all[-W:], model, alpha_list, 0.0, graph=gmode)
        clocks = ClockSampler(local)
        barrier()
        clocks.start()
        _C.prof_enable(True)
        _C.prof_report()
        l0 = _C.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        xt = ddim_loop_device(xt, seq_all[-(W + K):-W] if W else seq_all[-K:], model, alpha_list, 0.0, graph=gmode)
        e1.record()
        barrier()
        launches = _C.launch_count() - l0
        ms = max_over_ranks(e0.elapsed_time(e1))
        prof = _C.prof_report()
        _C.prof_enable(False)
        clk = clocks.stop()
    value = world * B * K / (ms * 1e-3)
    finite = bool(torch.isfinite(xt).all().item())
    peak_gb = torch.cuda.max_memory_allocated(dev) / 1e9
    del xt

    roof, kern, t_roof = roofline_from_profile(prof, K)

    # ---------------- e2e: public API, pinned host x in, both snapshot lists out, every step
    e2e = None
    if not args.no_e2e:
        ald = al.to(dev)
        ckw = {"chunks": args.e2e_chunks} if args.e2e_chunks > 0 else {}
        cur = x_host
        sets = [None, None]          # two recycled sets of pinned snapshot tensors: a call's input is the previous call's xs[-1]
        with torch.no_grad():
            for i in range(max(W, 2)):
                res = generalized_steps(cur, [seq_all[-1 - i]], model, ald, None, eta=0.0, out=sets[i & 1], **ckw)
                sets[i & 1] = res
                cur = res[0][-1]
            barrier()
            t0 = time.perf_counter()
            for i in range(K):
                j = (max(W, 2) + i) & 1
                res = generalized_steps(cur, [seq_all[-1 - W - i]], model, ald, None, eta=0.0, out=sets[j], **ckw)
                sets[j] = res
                cur = res[0][-1]
            barrier()
            dt = max_over_ranks(time.perf_counter() - t0)
        del sets, res, cur
        nbytes = x_host.numel() * 4
        e2e = {"value": world * B * K / dt, "unit": UNIT, "h2d_bytes_per_step": nbytes + 8 * B, "d2h_bytes_per_step": 2 * nbytes, "ms_per_step": 1e3 * dt / K,
               "pcie_GBps_per_rank": 3 * nbytes / (dt / K) / 1e9,
               "api": "generalized_steps(x_pinned_host, [t], model, alphas, None, eta=0, out=<recycled pinned lists>) once per step "
                      "(4 sample groups pipelined over PCIe)"}

    # ---------------- batch-position independence at the headline shape: samples 0 and B-1 of the full batch vs the same samples
    # run alone (the only check that exercises element offsets past 2^32; samples never interact)
    batch_pos = None
    if not args.no_parity and B >= 2:
        rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
        with torch.no_grad():
            xd = x_host.to(dev)
            tt = torch.full((B,), 999, dtype=torch.int64, device=dev)
            full = model(xd, tt)
            first, last = full[:1].clone(), full[B - 1:].clone()
            del full
            alone0 = model(xd[:1].contiguous(), tt[:1])
            alone1 = model(xd[B - 1:].contiguous(), tt[:1])
            r0, r1 = rel(first, alone0), rel(last, alone1)
            del xd, first, last, alone0, alone1
        batch_pos = {"shape": f"x[{B},2,{T},{F_SIZE}]", "samples": [0, B - 1], "eps_rel_l2_vs_batch_of_1": [r0, r1], "tolerance": 2e-2,
                     "pass": bool(max(r0, r1) <= 2e-2),
                     "note": "a different batch size changes the fp32 summation order of the GroupNorm statistics; bf16 storage amplifies that "
                             "to the path's own rounding noise, so the bound is the model tolerance"}

    # ---------------- CPU baseline (oracle; rank 0 at N=1 only) + parity of the CUDA path against that very run
    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n = args.cpu_sample_steps
        dt, cores, rec = cpu_steps(sd, T, n, record=not args.no_parity)
        cpu = {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"x[1,2,{T},{F_SIZE}] (1 of the {B} samples), {n} DDIM steps, fp32 torch CPU ops, {dt:.1f} s"}
        if rec is not None:
            parity = parity_vs_record(rec, model, al, dev)

    # ---------------- small batches (launch-bound regime): BASELINE configs[0] geometry and the reference's own default (audio.yml:76-77)
    small = None
    if rank == 0 and not args.no_small_batch:
        small = small_batch_leg(model, al, dev)

    # ---------------- training (BASELINE metric "train samples/sec", configs[1]/[4]): bounded leg, every rank, NCCL all-reduce when N > 1
    train = None
    if not args.no_train:
        del x_host
        torch.cuda.empty_cache()
        train = train_leg(sd, dev, world, rank, local, args.train_batch, args.train_t_size, args.train_steps, max(args.train_warmup, 3), train_graph=not args.no_train_graph,
                          prof_kernels=True, tail=False, cpu=False)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(B, T), "batch_per_gpu": B, "t_size": T, "f_size": F_SIZE, "eta": 0.0,
                       "storage": "bf16 NHWC activations, fp32 accumulate, fp32 x/eps", "parallelism": f"batch-sharded x{world}, no collectives",
                       "l2": "inputs larger than L2 (every activation tensor of the step exceeds 126 MB)" if B * T >= 16384 else "small debug shape: L2-resident",
                       "snapshots": "none inside the device-timed steps (config 3 selects 10 of 1000 steps); every step in e2e"},
            "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
            "step_roofline": {"t_roof_ms_per_step": t_roof / K, "frac": (t_roof / K) / (ms / K), "note": "sum over kernels of max(bytes/HBM peak, flops/tensor peak) / measured step"},
            "kernels": kern, "finite": finite, "peak_device_memory_gb": peak_gb, "parity": parity, "batch_position": batch_pos, "small_batch": small, "train": train,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_strong(args, model, x_host, al, dev, world, rank, local, B, T, K, W, barrier, max_over_ranks):
    """BASELINE configs[3]: ``--total-samples`` (512) split over the ranks; this rank's shard is device-resident and goes through the
    public ``generalized_steps`` (no snapshots), which walks it as as many sample groups as device memory requires.  One JSON line
    with "scaling": "strong"; value = total samples x steps / max-over-ranks device time."""
    import torch
    import torch.distributed as dist
    from ddim_audio_b200 import _C
    from ddim_audio_b200.functions.denoising import generalized_steps, max_group_size
    ald = al.to(dev)
    seq_all = list(range(1000))
    xd = x_host.to(dev)
    groups = -(-B // max_group_size(xd.shape, dev))
    with torch.no_grad():
        if W:
            generalized_steps(xd, seq_all[-W:], model, ald, set(), eta=0.0)
        clocks = ClockSampler(local)
        barrier()
        clocks.start()
        l0 = _C.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        generalized_steps(xd, seq_all[-(W + K):-W] if W else seq_all[-K:], model, ald, set(), eta=0.0)
        e1.record()
        barrier()
        launches = _C.launch_count() - l0
        ms = max_over_ranks(e0.elapsed_time(e1))
        clk = clocks.stop()
    if rank == 0:
        emit({"metric": METRIC, "value": args.total_samples * K / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
              "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
              "config": {"workload": f"audio.yml DDIM sampling, {args.total_samples} samples x[.,2,{T},{F_SIZE}] split over {world} GPU(s) "
                                     f"(BASELINE configs[3]), eta=0, uniform 1000-step schedule", "samples_per_gpu": B, "sample_groups_per_gpu": int(groups),
                         "t_size": T, "f_size": F_SIZE, "parallelism": f"batch-sharded x{world}, no collectives",
                         "api": "generalized_steps(x_cuda_shard, seq, model, alphas, set(), eta=0): groups sized from free device memory"},
              "clocks": clk, "gpu_launches": int(launches)})
    if world > 1:
        dist.destroy_process_group()


def parity_vs_record(rec, model, al, dev):
    """The CUDA path against the oracle run that ``cpu_baseline`` just timed (same weights, same sample x[1,2,T,256], same steps):
    per-step eps teacher-forced on the oracle's own xt, and the free-running public-API trajectory against the oracle's snapshots.
    Tolerance: DESIGN.md section 2 (bf16 activation storage, fp32 accumulate)."""
    import torch
    from ddim_audio_b200.functions.denoising import generalized_steps
    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
    eps_rel, eps_abs = [], []
    with torch.no_grad():
        for xt, t, et in rec["calls"]:
            got = model(xt.to(dev), t.to(dev)).float().cpu()
            eps_rel.append(rel(got, et))
            eps_abs.append(float((got - et).abs().max() / et.std().clamp_min(1e-30)))
        xs, x0 = generalized_steps(rec["x"].to(dev), rec["seq"], model, al.to(dev), None, eta=0.0)
    final = rel(xs[-1].cpu(), rec["xs"][-1])
    x0_rel = [rel(a.cpu(), b) for a, b in zip(x0, rec["x0"])]
    ok = max(eps_rel) <= 2e-2 and max(eps_abs) <= 0.114 and final <= 5e-2
    return {"shape": f"x{list(rec['x'].shape)}", "steps": len(rec["seq"]), "seq": [int(v) for v in rec["seq"]], "reference": "CPU oracle (the cpu_baseline run itself)",
            "per_step_eps_rel_l2": eps_rel, "max_abs_over_std": eps_abs, "x0_pred_rel_l2": x0_rel, "final_rel_l2": final,
            "tolerance": 2e-2, "tolerance_max_abs_over_std": 0.114, "tolerance_final_rel_l2": 5e-2, "pass": bool(ok)}


def small_batch_leg(model, al, dev, steps=10, warm=3):
    """Launch-bound shapes through the public API on a device-resident x (no snapshots: select_index = empty set): BASELINE
    configs[0] geometry x[4,2,1024,256] and the reference's own default x[2,2,8192,256] (audio.yml:76-77)."""
    import torch
    from ddim_audio_b200.functions.denoising import generalized_steps
    out = {}
    ald = al.to(dev)
    for name, (b, t) in {"b4_t1024": (4, 1024), "b2_t8192": (2, 8192)}.items():
        x = torch.randn(b, 2, t, F_SIZE, generator=torch.Generator().manual_seed(7)).to(dev)
        seq_w, seq_k = list(range(1000))[-warm:], list(range(1000))[-(warm + steps):-warm]
        with torch.no_grad():
            generalized_steps(x, seq_w, model, ald, set(), eta=0.0)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            generalized_steps(x, seq_k, model, ald, set(), eta=0.0)
            e1.record()
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
        out[name] = {"shape": f"x[{b},2,{t},{F_SIZE}]", "steps": steps, "sample_steps_per_s": b * steps / wall, "ms_per_step_wall": 1e3 * wall / steps,
                     "ms_per_step_device": e0.elapsed_time(e1) / steps}
    out["api"] = "generalized_steps(x_cuda, seq, model, alphas, set(), eta=0), wall clock with synchronize on both sides"
    return out


# ------------------------------------------------------------------------------------------------ training arm (--mode train)
def train_leg(sd, dev, world, rank, local, B, T, K, W, prof_kernels=True, tail=False, cpu=False, e2e=True, train_graph=True):
    """BASELINE configs[1] / [4]: K full train steps of audio.yml at x[B,2,T,256] per GPU, dropout 0.1 active -- per step:
    zero_grad, noise_estimation_loss forward + backward (functions/losses.py:4-18), NCCL all-reduce of the flat gradient buffer
    when N > 1, clip_grad_norm_ per group, AdaBelief + AdamW, LambdaLR, EMA (runners/diffusion.py:130-183), and the re-pack of the
    bf16 GEMM operands the changed weights force on the next forward.  Returns the ``train`` object of the JSON line
    (``samples_per_s`` = whole-job samples / max-over-ranks device time of the FULL step; ``fwd_bwd`` is the SURVEY.md section 8d
    metric: loss + backward + all-reduce only)."""
    import argparse

    import torch
    import torch.distributed as dist
    from ddim_audio_b200 import _C
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.functions import optim as fused
    from ddim_audio_b200.functions.losses import noise_estimation_loss
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200.models.ema import EMAHelper
    from ddim_audio_b200.runners import dist as rdist
    from ddim_audio_b200.runners.diffusion import Diffusion, antithetic_timesteps

    cfg = audio_config("torch.FloatTensor")
    model = Model(cfg).to(dev)
    model.load_state_dict(sd, strict=True)
    model.train()
    runner = Diffusion(argparse.Namespace(log_path=None), cfg, device=dev)
    optimizers, schedulers, grad_group = runner.build_training(model)
    ema = EMAHelper(mu=cfg.model.ema_rate)
    ema.register(model)
    al = alphas_cumprod().to(dev)
    g = torch.Generator().manual_seed(4321 + rank)
    x_host = torch.randn(B, 2, T, F_SIZE, generator=g).pin_memory()
    e_host = torch.randn(B, 2, T, F_SIZE, generator=g).pin_memory()
    t_host = antithetic_timesteps(B, 1000, g).pin_memory()
    params = list(model.parameters())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world > 1:
            tns = torch.tensor([v], dtype=torch.float64, device=dev)
            dist.all_reduce(tns, op=dist.ReduceOp.MAX)
            return float(tns.item())
        return v

    graphed = [None]

    def step(x0, e, t, ev=None):
        if ev:
            ev[0].record()
        if graphed[0] is not None:
            loss = graphed[0](x0, t, e)                      # captured re-pack + forward + backward (models/training.py)
        else:
            for o in optimizers.values():
                o.zero_grad(set_to_none=True)
            loss = noise_estimation_loss(model, x0, t, e, al)
            (loss / world if world > 1 else loss).backward()
        if ev:
            ev[1].record()
        rdist.allreduce_gradients(params, model.flat_grad)
        if ev:
            ev[2].record()
        for gg in grad_group.values():
            if gg.config.grad_clip is not None:
                fused.clip_grad_norm_(gg.params, gg.config.grad_clip)
        for o in optimizers.values():
            o.step()
        for sc in schedulers.values():
            sc.step()
        ema.update(model)
        if ev:
            ev[3].record()
        return loss

    x0, e, t = x_host.to(dev), e_host.to(dev), t_host.to(dev)
    step(x0, e, t)
    if train_graph:
        from ddim_audio_b200.models.training import GraphedLossBackward
        graphed[0] = GraphedLossBackward(model, al, x0, t, e, world)
    for _ in range(W):
        step(x0, e, t)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    l0 = _C.launch_count()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        loss = step(x0, e, t, evs[k])
    e1.record()
    barrier()
    launches = _C.launch_count() - l0
    ms = max_over_ranks(e0.elapsed_time(e1))
    # the per-kernel table comes from a SECOND pass of K steps with the library profiler on (an event pair around each of the ~1100
    # launches of a step costs the step ~3 ms: 57.1 against 54.0 ms); the throughput figures above are from the unprofiled pass
    prof = []
    if prof_kernels:
        held, graphed[0] = graphed[0], None     # the library profiler records at launch time: the profiled pass issues the step eagerly
        _C.prof_enable(True)
        _C.prof_report()
        for k in range(K):
            step(x0, e, t)
        torch.cuda.synchronize()
        prof = _C.prof_report()
        _C.prof_enable(False)
        graphed[0] = held
    clk = clocks.stop()
    fb_ms = max_over_ranks(sum(v[0].elapsed_time(v[2]) for v in evs) / K)
    ar_ms = max_over_ranks(sum(v[1].elapsed_time(v[2]) for v in evs) / K)
    tail_ms = max_over_ranks(sum(v[2].elapsed_time(v[3]) for v in evs) / K)
    finite = bool(torch.isfinite(loss).item()) and all(bool(torch.isfinite(p).all().item()) for p in params[:8])
    roof, kern, t_roof = roofline_from_profile(prof, K) if prof else (None, [], 0.0)
    n_el = sum(p.numel() for p in params)

    out = {"metric": "train_samples_per_sec", "samples_per_s": world * B * K / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / K,
           "steps": K, "warmup": W, "n_gpus": world, "batch_per_gpu": B, "t_size": T,
           "workload": f"audio.yml train step, x[{B},2,{T},{F_SIZE}] per GPU, dropout 0.1: loss + backward" + (" + NCCL all-reduce" if world > 1 else "") +
                       " + clip + AdaBelief/AdamW + EMA + operand re-pack (BASELINE configs[1]/[4] shape)",
           "fwd_bwd": {"samples_per_s": world * B / (fb_ms * 1e-3), "ms_per_step": fb_ms, "note": "zero_grad + loss + backward + all-reduce only (SURVEY.md section 8d train metric)"},
           "allreduce_ms": ar_ms, "allreduce_bytes": int(model.flat_grad.numel() * 4) if world > 1 else 0,
           "optimizer_tail": {"ms": tail_ms, "elements": n_el, "GBps": (12.0 + 28.0 + 12.0) * n_el / max(tail_ms, 1e-9) / 1e6,
                              "what": "clip x groups + AdaBelief + AdamW + LambdaLR + EMA, device time between events"},
           "gpu_launches": int(launches), "roofline": roof,
           "step_roofline": {"t_roof_ms_per_step": t_roof / K, "frac": (t_roof / K) / (ms / K)} if prof else None,
           "kernels": kern[:int(os.environ.get("DDIMB_BENCH_TRAIN_KERNELS", "12"))], "clocks": clk, "finite": finite, "last_loss": float(loss.item()), "dtype": "bf16",
           "timing": "CUDA events around K full steps, barrier + synchronize both sides, max over ranks; kernel table / roofline from a second, profiled pass of K eagerly issued steps",
           "graph": bool(graphed[0] is not None)}

    # e2e: pinned host batch in, loss scalar out, every step (the runner's own train_step does exactly this)
    if e2e:
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            l = step(x_host.to(dev, non_blocking=True), e_host.to(dev, non_blocking=True), t_host.to(dev, non_blocking=True))
            lv = float(l.item())
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        out["e2e"] = {"value": world * B * K / dt, "unit": "samples/s", "h2d_bytes_per_step": 2 * x_host.numel() * 4 + 8 * B, "d2h_bytes_per_step": 4,
                      "ms_per_step": 1e3 * dt / K, "api": "full train step on a pinned host batch + loss.item()", "last_loss": lv}

    # fused tail next to the same arithmetic as stock torch objects (torch AdamW for both groups: the reference's AdaBelief is a
    # Python loop of ~12 ops per tensor and does not travel to this box)
    if tail and rank == 0 and world == 1:
        tr = [p for n, p in model.named_parameters() if n.startswith("transformer.") and p.grad is not None]
        de = [p for n, p in model.named_parameters() if not n.startswith("transformer.") and p.grad is not None]
        saved = [p.data.clone() for p in params]

        def wall_ms(fn, reps=5):
            fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                fn()
            torch.cuda.synchronize()
            return 1e3 * (time.perf_counter() - t0) / reps

        t_opt = [torch.optim.AdamW(tr, lr=5e-4, betas=(0.9, 0.998), eps=1e-6, weight_decay=1e-4),
                 torch.optim.AdamW(de, lr=3e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-5)]
        shadow = {n: p.data.clone() for n, p in model.named_parameters()}

        def torch_tail():
            torch.nn.utils.clip_grad_norm_(tr, 1.0)
            torch.nn.utils.clip_grad_norm_(de, 1.0)
            for o in t_opt:
                o.step()
            for n, p in model.named_parameters():                                   # models/ema.py:16-23 as written
                shadow[n].data = (1.0 - 0.9999) * p.data + 0.9999 * shadow[n].data

        out["optimizer_tail"]["torch_ms"] = wall_ms(torch_tail)
        out["optimizer_tail"]["torch_note"] = "clip_grad_norm_ + torch.optim.AdamW (foreach) for both groups + the reference's per-tensor EMA loop, wall clock"
        with torch.no_grad():
            for p, v in zip(params, saved):
                p.data.copy_(v)
        model.mark_params_dirty()

    if cpu and rank == 0 and world == 1:
        from oracle import ddim_oracle as O
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        xs, es, ts = x_host[:1].clone(), e_host[:1].clone(), t_host[:1].clone()
        O.loss_and_grads(sd, xs[:, :, :32], ts, es[:, :, :32], alphas_cumprod())
        t0 = time.perf_counter()
        O.loss_and_grads(sd, xs, ts, es, alphas_cumprod())
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 1.0 / dt, "unit": "samples/s", "cores": cores, "kind": "port",
                               "sample": f"1 of the {B} samples, x[1,2,{T},{F_SIZE}], one loss + backward (torch CPU autograd over the oracle), {dt:.1f} s"}
    del model, optimizers, ema, x0, e, t
    torch.cuda.empty_cache()
    return out


def run_train(args):
    """``--mode train``: the train leg alone as the JSON line (value = full-step samples/s)."""
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 arm has no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    sd = synthetic_state_dict()
    tr = train_leg(sd, dev, world, rank, local, args.batch, args.t_size, args.steps, max(args.warmup, 0), prof_kernels=True, tail=True, train_graph=not args.no_train_graph,
                   cpu=not args.no_cpu_baseline, e2e=not args.no_e2e)
    if rank == 0:
        B, T = args.batch, args.t_size
        line = {
            "metric": "train_samples_per_sec", "value": tr["samples_per_s"], "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": tr["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": tr["workload"], "batch_per_gpu": B, "t_size": T, "f_size": F_SIZE,
                       "storage": "bf16 NHWC activations and activation gradients, fp32 accumulate, fp32 weight gradients",
                       "parallelism": f"data-parallel x{world}" + (", one NCCL all-reduce of the flat fp32 gradient buffer per step" if world > 1 else ""),
                       "l2": "inputs larger than L2 (level-0 activations are 537 MB each)" if B * T >= 16384 else "small debug shape"},
            "clocks": tr["clocks"], "e2e": tr.get("e2e"), "gpu_launches": tr["gpu_launches"], "roofline": tr["roofline"], "cpu_baseline": tr.get("cpu_baseline"),
            "step_roofline": tr["step_roofline"], "kernels": tr["kernels"], "finite": tr["finite"], "train": tr,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_parity(args):
    """BASELINE configs[0]: x[4,2,1024,256], seq = range(0,1000,20) (50 steps), eta = 0, select_index
    None.  The CPU oracle (checker) walks its own trajectory; at every step the CUDA denoiser is ALSO evaluated on the oracle's
    xt (teacher-forced per-step eps error), and the free-running CUDA trajectory through the public ``generalized_steps`` is
    compared snapshot by snapshot (final-sample max-abs / rel-L2).  One JSON line; not a throughput bench."""
    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --mode parity: no CUDA device")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.functions.denoising import generalized_steps
    from ddim_audio_b200.models.diffusion import Model
    from oracle import ddim_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = synthetic_state_dict()
    model = Model(audio_config("torch.FloatTensor")).to(dev)
    model.load_state_dict(sd, strict=True)
    model.eval()
    omodel = O.OracleModel(sd)
    al = alphas_cumprod()
    B, T, n = args.parity_batch, args.parity_t_size, args.parity_steps
    seq = list(range(0, 1000, 1000 // 50))[:n] if n <= 50 else list(range(0, 1000, max(1, 1000 // n)))
    x = torch.randn(B, 2, T, F_SIZE, generator=torch.Generator().manual_seed(1234))

    with torch.no_grad():
        generalized_steps(x.to(dev), seq[:2], model, al.to(dev), None, eta=0.0)       # warm-up (weight packing, descriptors)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        xs_g, x0_g = generalized_steps(x.to(dev), seq, model, al.to(dev), None, eta=0.0)
        torch.cuda.synchronize()
        gpu_s = time.perf_counter() - t0

    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
    xt = x.clone()
    t = torch.zeros(B, dtype=torch.long)
    eps_rel, eps_max_over_std, traj_rel, x0_rel = [], [], [], []
    cpu_s = 0.0
    for k, c in enumerate(O.ddim_coefficients(al, seq, 0.0)):
        t[...] = c["i"]
        t0 = time.perf_counter()
        et = omodel(xt, t)
        cpu_s += time.perf_counter() - t0
        with torch.no_grad():
            et_g = model(xt.to(dev), t.to(dev)).float().cpu()
        eps_rel.append(rel(et_g, et))
        eps_max_over_std.append(float((et_g - et).abs().max() / et.std().clamp_min(1e-30)))
        t0 = time.perf_counter()
        xt.add_(et, alpha=-((1 - c["at"]) ** 0.5)).div_(c["at"] ** 0.5)                # functions/denoising.py:27
        x0_rel.append(rel(x0_g[k].cpu(), xt))
        xt.mul_(c["at_next"] ** 0.5).add_(et, alpha=c["c2"])                           # :41-43, eta = 0
        cpu_s += time.perf_counter() - t0
        traj_rel.append(rel(xs_g[k + 1].cpu(), xt))
    final_g = xs_g[-1].cpu()
    line = {
        "mode": "parity", "impl": "b200-vs-oracle",
        "config": {"workload": f"audio.yml DDIM generalized_steps, x[{B},2,{T},{F_SIZE}], {len(seq)} steps of the uniform 50-step schedule, "
                               f"eta=0, select_index=None (BASELINE configs[0])", "weights": "seeded random init, norm.2.weight <- 0.5 randn"},
        "tolerance": {"per_step_eps_rel_l2": 2e-2, "per_step_eps_max_abs_over_std": 0.114,
                      "note": "bf16 activation storage, fp32 accumulate (DESIGN.md section 2)"},
        "per_step_eps": {"teacher_forced": True, "rel_l2_max": max(eps_rel), "rel_l2_mean": sum(eps_rel) / len(eps_rel),
                         "max_abs_over_std_max": max(eps_max_over_std), "rel_l2_first_last": [eps_rel[0], eps_rel[-1]]},
        "free_running": {"xs_rel_l2_by_step": [round(v, 6) for v in traj_rel[:: max(1, len(traj_rel) // 10)]] + [round(traj_rel[-1], 6)],
                         "x0_pred_rel_l2_last": x0_rel[-1]},
        "final_sample": {"max_abs": float((final_g - xt).abs().max()), "rel_l2": traj_rel[-1], "ref_std": float(xt.std()),
                         "ref_abs_max": float(xt.abs().max())},
        "cpu": {"kind": "port", "cores": cores, "seconds": cpu_s, "sample_steps_per_s": B * len(seq) / cpu_s},
        "gpu": {"seconds": gpu_s, "sample_steps_per_s": B * len(seq) / gpu_s,
                "api": "generalized_steps(x_cuda, seq, model, alphas, None, eta=0): every snapshot copied to the host, wall clock"},
        "pass": bool(max(eps_rel) <= 2e-2 and max(eps_max_over_std) <= 0.114),
    }
    emit(line)


def main():
    args = parse_args()
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    elif args.mode == "parity":
        run_parity(args)
    elif args.mode == "train":
        run_train(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
