"""Mint the BASELINE-shape fixtures from the UNMODIFIED reference (run in the build container only):

    python oracle/make_golden_fullshape.py

  model_b1_t8192      one ``Model.forward`` at the sampling geometry x[1,2,8192,256] (audio.yml:77; models/diffusion.py:237-294)
  ddim_b4_t1024_s3    ``generalized_steps`` at BASELINE configs[0] geometry x[4,2,1024,256], the last 3 steps of the 50-step
                      uniform schedule range(0,1000,20), eta = 0, select_index None (functions/denoising.py:10-52)

Full outputs are 4-16 MB each, so the fixtures keep strided probes (every PROBE_T-th frame, every PROBE_F-th bin -- co-prime
with the 128-wide kernel tiles so every tile column phase is hit) plus float64 norms / sums of the full tensors; inputs are
regenerated from seeds at test time.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ddim_oracle as O  # noqa: E402
from oracle import reference_loader as R  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
PROBE_T, PROBE_F = 61, 7


def probe(v: torch.Tensor) -> np.ndarray:
    return v[..., ::PROBE_T, ::PROBE_F].contiguous().numpy()


def summary(v: torch.Tensor) -> np.ndarray:
    d = v.double()
    return np.array([float(d.norm()), float(d.sum()), float(d.abs().max()), float(d.std())])


def main():
    assert R.available(), "reference tree not found"
    torch.set_num_threads(os.cpu_count())
    sd = O.make_state_dict(seed=0)
    model, _ = R.load_model(sd)
    ref_steps, _ = R.reference_functions()
    omodel = O.OracleModel(sd)
    alphas = O.alphas_cumprod()

    b, T, seed = 1, 8192, 51
    x, _, t = O.synthetic_inputs(b, T, seed)
    model.transformer.embedding.te = None
    with torch.no_grad():
        ref = model(x, t)
    got = omodel(x, t)
    print(f"model_b1_t8192: oracle vs reference max-abs {float((got - ref).abs().max()):.3e} rel-l2 {O.rel_l2(got, ref):.3e}")
    np.savez_compressed(os.path.join(OUT, "model_b1_t8192.npz"), probe=probe(ref), summary=summary(ref), t=t.numpy(), meta=np.array([b, T, seed]))

    b, T, seed, nsteps = 4, 1024, 52, 3
    x, _, _ = O.synthetic_inputs(b, T, seed)
    seq = list(range(0, 1000, 20))[-nsteps:]
    model.transformer.embedding.te = None
    with R.cpu_shims(noise=None):
        xs, x0 = ref_steps(x.clone(), seq, model, alphas, eta=0.0, select_index=None)
    oxs, ox0 = O.generalized_steps(x, seq, omodel, alphas, None, eta=0.0)
    err = max(float((p - q).abs().max()) for p, q in zip(xs[1:] + x0, oxs[1:] + ox0))
    print(f"ddim_b4_t1024_s3: oracle vs reference max-abs over {len(xs) - 1}+{len(x0)} snapshots {err:.3e}")
    np.savez_compressed(os.path.join(OUT, "ddim_b4_t1024_s3.npz"), xs_probe=np.stack([probe(v) for v in xs[1:]]), x0_probe=np.stack([probe(v) for v in x0]),
                        xs_summary=np.stack([summary(v) for v in xs[1:]]), x0_summary=np.stack([summary(v) for v in x0]),
                        seq=np.array(seq), meta=np.array([b, T, seed, nsteps]))
    print("wrote model_b1_t8192.npz, ddim_b4_t1024_s3.npz")


if __name__ == "__main__":
    main()
