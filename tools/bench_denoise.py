"""CUDA-event timing of the denoise_2d post-filter at the sampling shape, next to the reference's
own torch formula (External/UPU/signal/denoise.py:45-70 as torch CUDA ops -- what the reference runs on a GPU) on the same box.
    python tools/bench_denoise.py [--B 64 --T 8192 --F 256 --iters 5] > gpurun_out/denoise_bench.json
Per-kernel times come from the library profiler (DDIMB_PROF brackets the whole call) -- run under ncu for the launch list."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddim_audio_b200 import _C, ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--T", type=int, default=8192)
ap.add_argument("--F", type=int, default=256)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--torch-batch", type=int, default=8, help="samples of the torch-eager comparison (it needs ~10 tensor-sized temporaries)")
ap.add_argument("--no-torch", action="store_true")
a = ap.parse_args()
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(a.B, 2, a.T, a.F, device=dev, generator=g)
x += 0.7 * torch.randn(a.B, 2, a.T, 1, device=dev, generator=g) + 1.3 * torch.randn(a.B, 2, 1, a.F, device=dev, generator=g)


def timed(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def torch_eager(x):
    r = x - x.median(dim=-1, keepdim=True)[0]
    r -= r.median(dim=-2, keepdim=True)[0]
    for _ in range(5):
        denom = 1 / (1e-9 + torch.linalg.norm(r, ord=2, dim=-3, keepdim=True))
        num = r * denom
        for i in (-1, -2):
            A = num.mean(dim=i, keepdim=True)
            A /= denom.mean(dim=i, keepdim=True)
            r -= A
    return r


n0 = _C.launch_count()
REPS = 5
ms = timed(lambda: ops.denoise_2d(x, a.iters), REPS)
launches = (_C.launch_count() - n0) // (REPS + 2)
n = x.numel()
out = {"op": "denoise_2d", "shape": list(x.shape), "iters": a.iters, "ms": ms, "launches_per_call": launches,
       "algorithmic_GBps": 8.0 * n / ms / 1e6, "actual_passes_GBps": 4.0 * n * (7 + a.iters) / ms / 1e6,
       "samples_per_s": a.B / ms * 1e3}
if not a.no_torch:
    xb = x[: a.torch_batch]
    err = float((ops.denoise_2d(xb, 5) - torch_eager(xb)).abs().max())
    tms = timed(lambda: torch_eager(xb), 3, warm=1)
    out.update({"torch_eager_ms_per_sample": tms / xb.shape[0], "ours_ms_per_sample": ms / a.B, "max_abs_vs_torch_eager": err,
                "speedup_vs_torch_eager": (tms / xb.shape[0]) / (ms / a.B)})
print(json.dumps(out))
