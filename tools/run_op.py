"""Run one C-ABI op a few times on synthetic data (for ncu captures and quick CUDA-event timings).
    python tools/run_op.py conv3x3 --C 32 --W 256 --T 1024 --B 2 [--iters 5]"""
import argparse
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddim_audio_b200 import _C, ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("op")
ap.add_argument("--C", type=int, default=32)
ap.add_argument("--Cout", type=int, default=0)
ap.add_argument("--W", type=int, default=256)
ap.add_argument("--T", type=int, default=1024)
ap.add_argument("--B", type=int, default=2)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--nostats", action="store_true")
a = ap.parse_args()
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
bf = torch.bfloat16
C, W, T, B = a.C, a.W, a.T, a.B
Co = a.Cout or C
x = torch.randn(B, T, W, C, device=dev, generator=g).to(bf)
st = None if a.nostats else torch.zeros(B, 8, 2, dtype=torch.float64, device=dev)
if a.op == "conv3x3":
    w = (torch.randn(C, 9 * C, device=dev, generator=g) / math.sqrt(9 * C)).to(bf)
    temb = torch.randn(B, 4416, device=dev, generator=g)
    fn = lambda: ops.conv3x3(x, w, bias_b=temb, bias_b_stride=4416, act=1, stats=st)
elif a.op == "conv3x3_gn":
    w = (torch.randn(C, 9 * C, device=dev, generator=g) / math.sqrt(9 * C)).to(bf)
    temb = torch.randn(B, 4416, device=dev, generator=g)
    gam = torch.randn(C, device=dev, generator=g)
    bet = torch.randn(C, device=dev, generator=g)
    sti = torch.stack([x.double().reshape(B, -1, 8, C // 8).sum(dim=(1, 3)), (x.double() ** 2).reshape(B, -1, 8, C // 8).sum(dim=(1, 3))], -1).contiguous()
    fn = lambda: ops.conv3x3_gn(x, sti, gam, bet, 1e-6, int(os.environ.get("GN_MODE", "2")), w, bias_b=temb, bias_b_stride=4416, act=1, stats=st)
elif a.op == "wgrad":
    dy = torch.randn(B, T, W, C, device=dev, generator=g).to(bf)
    dw = torch.zeros(C, C, 3, 3, device=dev)
    fn = lambda: ops.conv3x3_wgrad(x, dy, dw)
elif a.op == "wgrad_down":                     # x [B,T,W,C] -> dy [B,T/2,W/2,Co]
    dy = torch.randn(B, T // 2, W // 2, Co, device=dev, generator=g).to(bf)
    dw = torch.zeros(Co, C, 4, 4, device=dev)
    fn = lambda: ops.conv_down4x4_wgrad(x, dy, dw)
elif a.op == "wgrad_up":                       # x [B,T,W,C] -> dy [B,2T,2W,Co]
    dy = torch.randn(B, 2 * T, 2 * W, Co, device=dev, generator=g).to(bf)
    dw = torch.zeros(C, Co, 4, 4, device=dev)
    fn = lambda: ops.conv_up4x4_wgrad(x, dy, dw)
elif a.op == "down":
    w = (torch.randn(Co, 16 * C, device=dev, generator=g) / math.sqrt(16 * C)).to(bf)
    b = torch.randn(Co, device=dev, generator=g)
    fn = lambda: ops.conv_down4x4(x, w, b, st)
elif a.op == "up":
    w = (torch.randn(16 * Co, C, device=dev, generator=g) / math.sqrt(4 * C)).to(bf)
    b = torch.randn(Co, device=dev, generator=g)
    sk = torch.randn(B, 2 * T, 2 * W, Co, device=dev, generator=g).to(bf)
    fn = lambda: ops.conv_up4x4(x, w, b, skip=sk, stats=st)
elif a.op == "gn":
    gam = torch.randn(C, device=dev, generator=g)
    bet = torch.randn(C, device=dev, generator=g)
    sti = torch.stack([x.double().reshape(B, -1, 8, C // 8).sum(dim=(1, 3)), (x.double() ** 2).reshape(B, -1, 8, C // 8).sum(dim=(1, 3))], -1).contiguous()
    fn = lambda: ops.gn_apply(x, sti, gam, bet, 1, 1e-6)
elif a.op == "gn2":
    gam = torch.randn(C, device=dev, generator=g)
    sti = torch.stack([x.double().reshape(B, -1, 8, C // 8).sum(dim=(1, 3)), (x.double() ** 2).reshape(B, -1, 8, C // 8).sum(dim=(1, 3))], -1).contiguous()
    res = torch.randn(B, T, W, C, device=dev, generator=g).to(bf)
    yb = torch.empty_like(x)
    fn = lambda: ops.gn_apply(x, sti, gam, None, 2, 1e-6, res=res, stats_out=st, out=yb)
elif a.op == "conv_out":
    w = torch.randn(2, 32, 3, 3, device=dev, generator=g) / 17
    b = torch.randn(2, device=dev, generator=g)
    sk = torch.randn(B, T, W, 32, device=dev, generator=g).to(bf)
    fn = lambda: ops.conv_out(x, sk, w, b)
elif a.op == "conv_out1":                       # one input plane (the denoiser folds the last skip into the preceding gn_apply)
    w = torch.randn(2, 32, 3, 3, device=dev, generator=g) / 17
    b = torch.randn(2, device=dev, generator=g)
    fn = lambda: ops.conv_out(x, None, w, b)
elif a.op == "conv_in":
    xi = torch.randn(B, 2, T, W, device=dev, generator=g)
    w = torch.randn(32, 2, 3, 3, device=dev, generator=g) / 4
    b = torch.randn(32, device=dev, generator=g)
    fn = lambda: ops.conv_in(xi, w, b, st)
else:
    raise SystemExit("unknown op")
fn()
torch.cuda.synchronize()
_C.prof_enable(True)
_C.prof_report()
for _ in range(a.iters):
    fn()
for r in _C.prof_report():
    ms = r["ms"] / r["calls"]
    print(f"{r['name']}: {ms:.4f} ms/call  {r['bytes'] / r['calls'] / ms * 1e-6:.1f} GB/s  {r['flops'] / r['calls'] / ms * 1e-9:.1f} TFLOP/s")
