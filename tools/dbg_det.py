import os, sys, math
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddim_audio_b200 import ops
dev = "cuda"; bf = torch.bfloat16
g = torch.Generator(device=dev).manual_seed(0)
for (C, W, T, B) in [(32, 256, 64, 3), (64, 128, 32, 3), (96, 64, 16, 3), (128, 32, 8, 3), (256, 8, 2, 3)]:
    x = torch.randn(B, T, W, C, device=dev, generator=g).to(bf)
    w = (torch.randn(C, 9 * C, device=dev, generator=g) / math.sqrt(9 * C)).to(bf)
    temb = torch.randn(B, 4416, device=dev, generator=g)
    outs = []
    for it in range(3):
        st = torch.zeros(B, 8, 2, dtype=torch.float64, device=dev)
        o = ops.conv3x3(x, w, bias_b=temb, bias_b_stride=4416, act=1, stats=st)
        outs.append((o.clone(), st.clone()))
    d_o = max(float((outs[0][0].float() - o.float()).abs().max()) for o, _ in outs[1:])
    d_s = max(float(((outs[0][1] - s) / outs[0][1].abs().clamp_min(1e-9)).abs().max()) for _, s in outs[1:])
    print(f"conv C={C}: out maxdiff {d_o:.3e}  stats reldiff {d_s:.3e}")
    gam = torch.randn(C, device=dev, generator=g); bet = torch.randn(C, device=dev, generator=g)
    res = torch.randn(B, T, W, C, device=dev, generator=g).to(bf)
    sti = outs[0][1]
    r = []
    for it in range(3):
        so = torch.zeros(B, 8, 2, dtype=torch.float64, device=dev)
        y = ops.gn_apply(outs[0][0], sti, gam, None, 2, 1e-6, res=res, stats_out=so)
        r.append((y.clone(), so.clone()))
    d_o = max(float((r[0][0].float() - o.float()).abs().max()) for o, _ in r[1:])
    d_s = max(float(((r[0][1] - s) / r[0][1].abs().clamp_min(1e-9)).abs().max()) for _, s in r[1:])
    print(f"  gn mode2: out maxdiff {d_o:.3e}  stats reldiff {d_s:.3e}")
