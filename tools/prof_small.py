"""Per-kernel device time of one denoiser forward at a small batch (CUDA events around every launch; serialises nothing).
    python tools/prof_small.py [--B 4 --T 1024]"""
import argparse, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddim_audio_b200 import _C
from ddim_audio_b200.config import audio_config
from ddim_audio_b200.models.diffusion import Model

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=4)
ap.add_argument("--T", type=int, default=1024)
a = ap.parse_args()
torch.manual_seed(0)
torch.set_grad_enabled(False)      # the sampling path (with gradients enabled Model.forward is the training node)
m = Model(audio_config()).cuda().eval()
x = torch.randn(a.B, 2, a.T, 256, device="cuda")
t = torch.full((a.B,), 500, device="cuda", dtype=torch.int64)
for _ in range(3):
    m(x, t)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    m(x, t)
e1.record()
torch.cuda.synchronize()
print(f"forward wall {e0.elapsed_time(e1) / 10:.3f} ms (eager launches)")
_C.prof_enable(True)
for _ in range(5):
    m(x, t)
torch.cuda.synchronize()
rep = _C.prof_report()
_C.prof_enable(False)
rows = sorted(rep, key=lambda r: -r["ms"])
tot = sum(r["ms"] for r in rows) / 5
print(f"sum of kernel times {tot:.3f} ms / forward, {sum(r['calls'] for r in rows) // 5} launches")
for r in rows[:40]:
    print(f"{r['name']:>26}  calls {r['calls'] // 5:3d}  {r['ms'] / r['calls'] * 1e3:8.1f} us/call  {r['ms'] / 5:7.3f} ms/fwd  {r['bytes'] / r['ms'] / 1e6 if r['ms'] else 0:7.0f} GB/s {r['flops'] / r['ms'] / 1e9 if r['ms'] else 0:7.1f} TF/s")
