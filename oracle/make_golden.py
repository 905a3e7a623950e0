"""Mint tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    python oracle/make_golden.py            # writes tests/golden/, prints oracle-vs-reference errors

The reference has no golden vectors of its own, so these fixtures -- outputs
of the reference's Model / generalized_steps / noise_estimation_loss on seeded inputs and the
seeded synthetic weights of ``ddim_oracle.make_state_dict`` -- are what pins the oracle, and
through it the CUDA path.  Inputs are regenerated from seeds at test time; only outputs (and
a few strided probes of large intermediates) are stored, so the fixtures stay small.
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

# (name, batch, T, seed) -- all at audio.yml geometry (F=256, 6 levels), short T
MODEL_CASES = [("model_b2_t64", 2, 64, 11), ("model_b1_t32", 1, 32, 12), ("model_b3_t96", 3, 96, 13)]
SAMPLER_CASES = [("ddim_b2_t32_eta0", 2, 32, 21, 0.0, 5, None), ("ddim_b1_t64_eta1", 1, 64, 22, 1.0, 4, {0, 2, -1})]
LOSS_CASES = [("loss_b4_t32", 4, 32, 31)]
GRAD_CASES = [("grad_b2_t32", 2, 32, 41)]


def main():
    assert R.available(), "reference tree not found"
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    sd = O.make_state_dict(seed=0)
    model, config = R.load_model(sd)
    ref_steps, ref_loss = R.reference_functions()
    omodel = O.OracleModel(sd)
    alphas = O.alphas_cumprod()

    # schedule: runners/diffusion.py:91-128 needs args/config plumbing; reproduce its tensor ops on
    # the reference's own get_beta_schedule instead
    with R._ref_cwd():
        from runners.diffusion import get_beta_schedule
    betas = get_beta_schedule("linear", beta_start=config.diffusion.beta_start, beta_end=config.diffusion.beta_end,
                              num_diffusion_timesteps=config.diffusion.num_diffusion_timesteps)
    a = torch.from_numpy(np.concatenate([[1], 1.0 - betas])).type("torch.FloatTensor").cumprod(dim=0)[1:]
    assert torch.equal(a, alphas), "oracle schedule != reference schedule"
    np.savez_compressed(os.path.join(OUT, "schedule.npz"), alphas=a.numpy())

    def fresh():
        # models/diffusion.py:132 keeps a positional table across calls and (inverted test) reuses a
        # SMALLER cached one for a longer input; every case starts from the state a new process has.
        model.transformer.embedding.te = None

    for name, b, T, seed in MODEL_CASES:
        fresh()
        x, _, t = O.synthetic_inputs(b, T, seed)
        with torch.no_grad():
            ref = model(x, t)
        got = omodel(x, t)
        print(f"{name}: oracle vs reference max-abs {float((got-ref).abs().max()):.3e} rel-l2 {O.rel_l2(got, ref):.3e}")
        np.savez_compressed(os.path.join(OUT, name + ".npz"), out=ref.numpy(), t=t.numpy(),
                            meta=np.array([b, T, seed]))

    for name, b, T, seed, eta, nsteps, sel in SAMPLER_CASES:
        x, _, _ = O.synthetic_inputs(b, T, seed)
        seq = O.uniform_seq(1000, nsteps)
        g = torch.Generator().manual_seed(seed + 1000)
        noise = [torch.randn(x.shape, generator=g) for _ in seq]
        fresh()
        with R.cpu_shims(noise=noise):
            xs, x0 = ref_steps(x.clone(), seq, model, alphas, eta=eta, select_index=sel)
        oxs, ox0 = O.generalized_steps(x, seq, omodel, alphas, sel, noise=noise, eta=eta)
        assert len(xs) == len(oxs) and len(x0) == len(ox0), (len(xs), len(oxs))
        # xs[0] of the reference is the caller's tensor; on CPU xt is a type() copy so it is untouched
        err = max(float((p - q).abs().max()) for p, q in zip(xs[1:] + x0, oxs[1:] + ox0))
        print(f"{name}: oracle vs reference max-abs over {len(xs)-1}+{len(x0)} snapshots {err:.3e}")
        np.savez_compressed(os.path.join(OUT, name + ".npz"),
                            xs=np.stack([v.numpy() for v in xs[1:]]), x0=np.stack([v.numpy() for v in x0]),
                            meta=np.array([b, T, seed, nsteps]), eta=np.array(eta),
                            sel=np.array(sorted(sel) if sel is not None else [], dtype=np.int64), has_sel=np.array(sel is not None))

    for name, b, T, seed in LOSS_CASES:
        x0, e, t = O.synthetic_inputs(b, T, seed)
        fresh()
        with torch.no_grad():
            ref = ref_loss(model, x0, t, e, alphas)
            refk = ref_loss(model, x0, t, e, alphas, keepdim=True)
        got = O.noise_estimation_loss(omodel, x0, t, e, alphas)
        print(f"{name}: oracle {float(got):.6f} reference {float(ref):.6f}")
        np.savez_compressed(os.path.join(OUT, name + ".npz"), loss=ref.numpy(), per_sample=refk.numpy(), t=t.numpy(),
                            meta=np.array([b, T, seed]))

    # gradients of the training loss through the UNMODIFIED reference (eval mode: dropout off),
    # stored as per-parameter fingerprints (norm, sum, seeded projection) + the oracle's agreement with them
    for name, b, T, seed in GRAD_CASES:
        x0, e, t = O.synthetic_inputs(b, T, seed)
        fresh()
        model.zero_grad(set_to_none=True)
        loss = ref_loss(model, x0, t, e, alphas)
        loss.backward()
        oloss, ograds = O.loss_and_grads(sd, x0, t, e, alphas)
        keys, fps, worst = [], [], 0.0
        for k, p in model.named_parameters():
            keys.append(k)
            fps.append(O.grad_fingerprint(k, p.grad))
            worst = max(worst, O.rel_l2(ograds[k], p.grad))
        print(f"{name}: loss oracle {float(oloss):.6f} reference {float(loss):.6f}; worst per-parameter grad rel-l2 oracle vs reference {worst:.3e}")
        np.savez_compressed(os.path.join(OUT, name + ".npz"), loss=loss.detach().numpy(), keys=np.array(keys), fp=np.array(fps, dtype=np.float64),
                            t=t.numpy(), meta=np.array([b, T, seed]))
        model.zero_grad(set_to_none=True)

    # the reference's own default init under torch.manual_seed(0): the drop-in Model must consume the RNG identically
    torch.manual_seed(0)
    fresh_model, _ = R.load_model(None)
    fsd = fresh_model.state_dict()
    np.savez_compressed(os.path.join(OUT, "init_fingerprint.npz"), keys=np.array(list(fsd.keys())),
                        sums=np.array([float(v.double().sum()) for v in fsd.values()]),
                        abs_sums=np.array([float(v.double().abs().sum()) for v in fsd.values()]),
                        shapes=np.array([str(tuple(v.shape)) for v in fsd.values()]))

    # weight fingerprint so a torch upgrade that changes the generator stream is caught, not silently absorbed
    fp = {k: float(v.double().sum()) for k, v in list(sd.items())[:: max(1, len(sd) // 24)]}
    np.savez_compressed(os.path.join(OUT, "weights_fingerprint.npz"), keys=np.array(list(fp.keys())), sums=np.array(list(fp.values())))
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
