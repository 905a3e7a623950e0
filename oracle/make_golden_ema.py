"""Mint tests/golden/ema_steps.npz from the UNMODIFIED reference EMAHelper (models/ema.py; build container only).

    python oracle/make_golden_ema.py

A small module with the seeded parameter trajectory of ``ddim_oracle.ema_case`` is registered, its parameters are overwritten
three times with ``update`` after each; the final shadow tensors are stored."""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ddim_oracle as O  # noqa: E402

REF = os.environ.get("DDIM_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
MU, SEED = 0.9999, 61                       # audio.yml ema_rate


class Holder(nn.Module):
    def __init__(self, tensors):
        super().__init__()
        self.keys = list(tensors)
        for i, k in enumerate(self.keys):
            self.register_parameter(f"p{i}", nn.Parameter(tensors[k].clone()))

    def named_parameters(self, *a, **k):
        for i, key in enumerate(self.keys):
            yield key, getattr(self, f"p{i}")


def main():
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location("ref_ema", os.path.join(REF, "models", "ema.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _, traj = O.ema_case(SEED)
    m = Holder(traj[0])
    helper = mod.EMAHelper(mu=MU)
    helper.register(m)
    shadow = {k: v.clone() for k, v in traj[0].items()}
    for step in traj[1:]:
        with torch.no_grad():
            for k, p in m.named_parameters():
                p.copy_(step[k])
        helper.update(m)
        shadow = O.ema_update(shadow, step, MU)
    worst = max(float((helper.shadow[k] - shadow[k]).abs().max()) for k in shadow)
    print(f"ema: oracle vs reference max-abs {worst:.3e}")
    np.savez_compressed(os.path.join(OUT, "ema_steps.npz"), mu=np.array(MU), seed=np.array(SEED),
                        **{k.replace(".", "__"): v.detach().numpy() for k, v in helper.shadow.items()})


if __name__ == "__main__":
    main()
