"""cProfile of the host side of loss forward + backward at x[32,2,1024,256] (the train step is host-bound): top functions by own time."""
import cProfile, os, pstats, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddim_audio_b200.config import audio_config
from ddim_audio_b200.models.diffusion import Model
from ddim_audio_b200.functions.losses import loss_registry

torch.manual_seed(0)
m = Model(audio_config()).cuda().train()
B, T = 32, 1024
x = torch.randn(B, 2, T, 256, device="cuda"); e = torch.randn_like(x)
t = torch.randint(0, 1000, (B,), device="cuda"); a = torch.linspace(0.999, 0.01, 1000, device="cuda")
def step():
    for p in m.parameters(): p.grad = None
    loss_registry["simple"](m, x, t, e, a).backward()
for _ in range(2): step()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(3): step()
pr.disable(); torch.cuda.synchronize()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(28)
