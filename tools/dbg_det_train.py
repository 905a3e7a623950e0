import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ddim_oracle as O
from ddim_audio_b200.config import audio_config
from ddim_audio_b200.models.diffusion import Model
from ddim_audio_b200.models import training as TR
from ddim_audio_b200 import ops
sd = O.make_state_dict(0)
m = Model(audio_config()); m.load_state_dict(sd); m.eval()
al = O.alphas_cumprod().cuda()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
x0, e, t = O.synthetic_inputs(B, 32, 61)
xt = ops.qsample(x0.cuda(), e.cuda(), al, t.cuda())
svs = []
for i in range(6):
    with torch.no_grad():
        o, sv = TR.forward_train(m, xt, t.cuda())
    sv.out = o
    svs.append(sv)
def flat(sv):
    items = []
    for li, rec in enumerate(sv.down):
        for bi, blk in enumerate(rec["blocks"]):
            for k in ("h", "a0", "u0", "n1", "u1"):
                items.append((f"down{li}.b{bi}.{k}", blk[k]))
    items.append(("hidden6", sv.hidden[-1])); items.append(("tok", sv.tok))
    for li, r in enumerate(sv.layers):
        for k in ("xr", "y", "hr", "u", "v"):
            items.append((f"fnet{li}.{k}", r[k]))
    items.append(("xbf_final", sv.xbf_final))
    for li, rec in enumerate(sv.up):
        for bi, blk in enumerate(rec["blocks"]):
            for k in ("h", "a0", "u0", "n1", "u1"):
                items.append((f"up{li}.b{bi}.{k}", blk[k]))
    items.append(("out", sv.out))
    return items
ref = flat(svs[0])
for i in range(1, 6):
    first = None
    for (n, a), (_, b) in zip(flat(svs[i]), ref):
        d = float((a.float() - b.float()).abs().max())
        if d > 0:
            first = (n, d, a.data_ptr() % 1024, b.data_ptr() % 1024, tuple(a.shape)); break
    print("run", i, "first differing tensor:", first)
