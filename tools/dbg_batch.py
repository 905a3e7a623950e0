import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ddim_oracle as O
from ddim_audio_b200.config import audio_config
from ddim_audio_b200.models.diffusion import Model
def rel(a, b): return float((a.double()-b.double()).norm()/b.double().norm())
m = Model(audio_config()); m.load_state_dict(O.make_state_dict(0)); m.eval()
x, _, t = O.synthetic_inputs(3, 64, 5)
x, t = x.cuda(), t.cuda()
a1 = m(x, t); a2 = m(x, t)
print("run-to-run", rel(a1, a2))
for i in range(3):
    b = m(x[i:i+1], t[i:i+1])
    print("sample", i, "B=3 vs B=1", rel(a1[i:i+1], b))
ref = O.OracleModel(O.make_state_dict(0))(x.cpu(), t.cpu())
print("vs oracle B=3", rel(a1.cpu(), ref))
