"""Mint tests/golden/optim_steps.npz from the reference's own optimizers (build container only).

    python oracle/make_golden_optim.py

Three train-step tails (runners/diffusion.py:155-173: clip_grad_norm_ with max_norm 1, then optimizer.step()) on the seeded
parameters / gradients of ``ddim_oracle.optim_case``, once with the UNMODIFIED ``clip_opt.AdaBelief`` imported from
/root/reference/External/step-clip-optimizer (configured as functions/__init__.py:24-42 with audio.yml's default group) and once
with ``torch.optim.AdamW`` (audio.yml's transformer group), a LambdaLR warm-up in between as the runner has.  Stores the final
parameters and optimizer state (tensors above 20 K elements as every 7th element, ``ddim_oracle.optim_probe``); prints the oracle's agreement."""
from __future__ import annotations

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
SEED = 71
CFG = {  # audio.yml:84-104
    "adabelief": dict(lr=0.0003, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.00001, warmup=1000),
    "adamw": dict(lr=0.0005, betas=(0.9, 0.998), eps=1e-6, weight_decay=0.0001, warmup=10000),
}


def lr_at(cfg, step):   # functions/__init__.py:53-60 LambdaLR, `step` optimizer steps already taken
    w = cfg["warmup"]
    return cfg["lr"] * min(((1 + step) / w) ** -0.5, (1 + step) / w)


def main():
    sys.dont_write_bytecode = True
    sys.path.append(os.path.join(REF, "External", "step-clip-optimizer"))
    from clip_opt import AdaBelief

    out = {}
    for name, cfg in CFG.items():
        params0, grads = O.optim_case(SEED)
        ps = {k: nn.Parameter(v.clone()) for k, v in params0.items()}
        kw = dict(lr=cfg["lr"], betas=cfg["betas"], eps=cfg["eps"], weight_decay=cfg["weight_decay"], amsgrad=False)
        if name == "adabelief":
            opt = AdaBelief(list(ps.values()), weight_decouple=True, fixed_decay=False, rectify=False, clip_step=None, norm_ord=2, **kw)
            second = "exp_avg_var"
        else:
            opt = torch.optim.AdamW(list(ps.values()), **kw)
            second = "exp_avg_sq"
        sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s, w=cfg["warmup"]: min(((1 + s) / w) ** -0.5, (1 + s) / w))
        # oracle alongside
        op = {k: v.clone() for k, v in params0.items()}
        om = {k: torch.zeros_like(v) for k, v in params0.items()}
        ov = {k: torch.zeros_like(v) for k, v in params0.items()}
        norms = []
        for i, g in enumerate(grads):
            for k, p in ps.items():
                p.grad = g[k].clone()
            norms.append(float(torch.nn.utils.clip_grad_norm_(list(ps.values()), 1.0)))
            opt.step()
            sched.step()
            _, cg = O.clip_grad_norm(g, 1.0)
            fn = O.adabelief_step if name == "adabelief" else O.adamw_step
            for k in op:
                op[k], om[k], ov[k] = fn(op[k], cg[k], om[k], ov[k], i + 1, lr_at(cfg, i), cfg["betas"], cfg["eps"], cfg["weight_decay"])
        worst = 0.0
        for k, p in ps.items():
            st = opt.state[p]
            worst = max(worst, float((p.data - op[k]).abs().max()), float((st["exp_avg"] - om[k]).abs().max()),
                        float(((st[second] - ov[k]).abs() / st[second].abs().clamp_min(1e-30)).max()) * 1e-3)
            kk = k.replace(".", "__")
            out[f"{name}__p__{kk}"] = O.optim_probe(p.data).numpy().copy()
            out[f"{name}__m__{kk}"] = O.optim_probe(st["exp_avg"]).numpy().copy()
            out[f"{name}__v__{kk}"] = O.optim_probe(st[second]).numpy().copy()
        out[f"{name}__norms"] = np.array(norms)
        print(f"{name}: total norms {norms}; oracle vs reference worst abs (p, m; v relative x1e-3) {worst:.3e}")
    np.savez_compressed(os.path.join(OUT, "optim_steps.npz"), seed=np.array(SEED), **out)


if __name__ == "__main__":
    main()
