"""Where a train step's wall time goes: host time (launch issue) against device time of each phase at x[32,2,1024,256].
    python tools/train_breakdown.py"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddim_audio_b200 import _C
from ddim_audio_b200.config import audio_config
from ddim_audio_b200.models.diffusion import Model
from ddim_audio_b200.functions.losses import loss_registry

torch.manual_seed(0)
dev = "cuda"
m = Model(audio_config()).to(dev).train()
B, T = 32, 1024
x = torch.randn(B, 2, T, 256, device=dev)
e = torch.randn_like(x)
t = torch.randint(0, 1000, (B,), device=dev)
a = torch.linspace(0.999, 0.01, 1000, device=dev)


def phase(name, fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host = 0.0
    e0.record()
    for _ in range(n):
        h0 = time.perf_counter(); r = fn(); host += time.perf_counter() - h0
    e1.record(); torch.cuda.synchronize()
    print(f"{name:>28}: device span {e0.elapsed_time(e1) / n:7.2f} ms   host issue {host / n * 1e3:7.2f} ms   launches {(_C.launch_count() - phase.l0) // (n + 1)}")
    phase.l0 = _C.launch_count()
    return r
phase.l0 = _C.launch_count()

def zero():
    for p in m.parameters(): p.grad = None
def repack():
    m.mark_params_dirty(); m._packed_for(True)
def fwd():
    with torch.enable_grad():
        return loss_registry["simple"](m, x, t, e, a)
def fwd_bwd():
    zero()
    l = fwd(); l.backward(); return l
phase("zero_grad", zero)
phase("re-pack (train)", repack)
phase("loss forward", fwd)
phase("zero + forward + backward", fwd_bwd)
