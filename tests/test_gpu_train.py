"""-m gpu: the training path -- noise_estimation_loss forward + backward through the CUDA denoiser -- against fp32 autograd
of the CPU oracle on the same seeded inputs, and against the gradient fingerprint minted from the UNMODIFIED reference
(tests/golden/grad_b2_t32.npz).  Tolerance (bf16 activations AND bf16 activation gradients, fp32 accumulation): loss rel
<= 5e-3; per-parameter gradient rel-L2 <= 5e-2 for tensors that carry at least 1e-4 of the total gradient norm (smaller
ones are compared on the scale of the total), global rel-L2 over all 47.16 M values <= 3e-2."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import DEV, rel_l2  # noqa: E402
from oracle import ddim_oracle as O  # noqa: E402

LOSS_REL, PARAM_REL, GLOBAL_REL = 5e-3, 5e-2, 3e-2


@pytest.fixture(scope="module")
def model(oracle_sd):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    m = Model(audio_config())
    m.load_state_dict(oracle_sd, strict=True)
    m.eval()                       # parity runs with dropout off; gradients still flow
    return m


def _cuda_grads(model, x0, t, e):
    from ddim_audio_b200.functions.losses import noise_estimation_loss
    model.zero_grad(set_to_none=True)
    loss = noise_estimation_loss(model, x0.to(DEV), t.to(DEV), e.to(DEV), O.alphas_cumprod().to(DEV))
    loss.backward()
    return loss.detach().cpu(), {k: p.grad.detach().cpu() for k, p in model.named_parameters()}


def _compare(grads, ref, tag):
    tot = sum(float(v.double().pow(2).sum()) for v in ref.values()) ** 0.5
    err2, worst = 0.0, (0.0, "")
    for k, r in ref.items():
        g = grads[k]
        assert g.shape == r.shape and torch.isfinite(g).all(), k
        d = float((g.double() - r.double()).norm())
        err2 += d * d
        rn = float(r.double().norm())
        rel = d / rn if rn >= 1e-4 * tot else d / (1e-4 * tot)
        if rel > worst[0]:
            worst = (rel, k)
    print(f"{tag}: global grad rel-l2 {err2 ** 0.5 / tot:.3e}; worst parameter {worst[1]} rel {worst[0]:.3e}")
    assert err2 ** 0.5 / tot <= GLOBAL_REL
    assert worst[0] <= PARAM_REL, worst


def test_gradients_vs_reference_fingerprint_and_oracle(model, oracle_sd, golden_dir):
    g = np.load(os.path.join(golden_dir, "grad_b2_t32.npz"))
    b, T, seed = (int(v) for v in g["meta"])
    x0, e, t = O.synthetic_inputs(b, T, seed)
    loss, grads = _cuda_grads(model, x0, t, e)
    assert abs(float(loss) - float(g["loss"])) <= LOSS_REL * float(g["loss"])
    # reference fingerprint: per-parameter norms
    tot = float(np.sqrt((g["fp"][:, 0] ** 2).sum()))
    for k, fp in zip(g["keys"], g["fp"]):
        norm, _, _ = O.grad_fingerprint(str(k), grads[str(k)])
        assert abs(norm - fp[0]) <= PARAM_REL * max(fp[0], 1e-4 * tot), (str(k), norm, fp[0])
    # full comparison with oracle autograd (bit-identical to the reference's gradients in the build container)
    _, ref = O.loss_and_grads(oracle_sd, x0, t, e, O.alphas_cumprod())
    _compare(grads, ref, "b2 t32")


def test_gradients_vs_oracle_t64_and_accumulation(model, oracle_sd):
    x0, e, t = O.synthetic_inputs(3, 64, 52)
    loss, grads = _cuda_grads(model, x0, t, e)
    oloss, ref = O.loss_and_grads(oracle_sd, x0, t, e, O.alphas_cumprod())
    assert abs(float(loss) - float(oloss)) <= LOSS_REL * float(oloss)
    _compare(grads, ref, "b3 t64")
    # every gradient of a fresh backward is a view into ONE flat buffer (single-collective data parallelism)
    flat = model.flat_grad
    lo, hi = flat.data_ptr(), flat.data_ptr() + flat.numel() * 4
    assert all(lo <= p.grad.data_ptr() < hi for p in model.parameters())
    # a second backward without zero_grad accumulates (torch semantics)
    from ddim_audio_b200.functions.losses import noise_estimation_loss
    l2 = noise_estimation_loss(model, x0.to(DEV), t.to(DEV), e.to(DEV), O.alphas_cumprod().to(DEV))
    l2.backward()
    k = "down_modules.3.1.conv.0.weight"
    assert rel_l2(dict(model.named_parameters())[k].grad.cpu(), 2 * grads[k]) < 1e-3


def test_data_parallel_shards_sum_to_full_batch(model):
    """mean-over-batch loss: the 1/world-scaled gradients of two half batches add up to the full-batch gradient (what the
    NCCL all-reduce computes; GroupNorm / LayerNorm / FFT are per sample, so this holds up to summation order)."""
    x0, e, t = O.synthetic_inputs(4, 32, 61)
    _, full = _cuda_grads(model, x0, t, e)
    _, a = _cuda_grads(model, x0[:2], t[:2], e[:2])
    _, b = _cuda_grads(model, x0[2:], t[2:], e[2:])
    tot = sum(float(v.double().pow(2).sum()) for v in full.values()) ** 0.5
    err = sum(float((0.5 * (a[k].double() + b[k].double()) - full[k].double()).pow(2).sum()) for k in full) ** 0.5
    print(f"dp shards vs full batch: rel {err / tot:.3e}")
    # not bitwise: the fp64 statistics atomics commute only to ~1e-16, which occasionally flips a bf16 rounding that the
    # network then amplifies (the same run-to-run jitter the forward has); well below the bf16 gradient noise floor (~1e-2)
    assert err / tot < 1e-2


def test_train_mode_dropout_runs_and_is_seeded(model):
    from ddim_audio_b200.functions.losses import noise_estimation_loss
    x0, e, t = O.synthetic_inputs(2, 32, 71)
    a = O.alphas_cumprod().to(DEV)
    model.train()
    try:
        torch.manual_seed(5)
        l1 = noise_estimation_loss(model, x0.to(DEV), t.to(DEV), e.to(DEV), a)
        model.zero_grad(set_to_none=True)
        l1.backward()
        g1 = model.transformer.compute_out.weight.grad.clone()
        torch.manual_seed(5)
        l2 = noise_estimation_loss(model, x0.to(DEV), t.to(DEV), e.to(DEV), a)
        torch.manual_seed(6)
        l3 = noise_estimation_loss(model, x0.to(DEV), t.to(DEV), e.to(DEV), a)
        assert float(l1) == float(l2) and float(l1) != float(l3)
        assert torch.isfinite(g1).all() and float(g1.abs().max()) > 0
    finally:
        model.eval()
        model.zero_grad(set_to_none=True)


def test_runner_train_step_updates_parameters(oracle_sd):
    """Diffusion.train over an injected loader with audio.yml's own optimizers: loss/backward, gradient clipping, AdaBelief
    (default group) + AdamW (transformer group) and the EMA all in library kernels on the fp32 masters."""
    import argparse
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200.runners.diffusion import Diffusion
    cfg = audio_config()
    cfg.training.snapshot_freq = 10 ** 9
    m = Model(cfg)
    m.load_state_dict(oracle_sd, strict=True)
    before = m.transformer.compute_out.weight.detach().clone()
    r = Diffusion(argparse.Namespace(log_path=None), cfg)
    x0, _, _ = O.synthetic_inputs(2, 32, 81)
    _, log = r.train(loader=[(x0, None)], model=m, max_steps=2)
    assert log["step"] == 2 and np.isfinite(log["loss"])
    assert not torch.equal(before, m.transformer.compute_out.weight.detach())
    # config.model.ema: the runner registered the fused EMAHelper (runners/diffusion.py:232-236) and updated it every step
    sh = r.ema_helper.shadow["transformer.compute_out.weight"]
    assert torch.isfinite(sh).all() and not torch.equal(sh, before) and not torch.equal(sh, m.transformer.compute_out.weight.detach())


def _train_cfg(lr_scale=1.0):
    """audio.yml's optimizers with the LambdaLR warm-up off (its first factors are 1e-4 / 1e-3: nothing would move in three
    steps) and dropout off (parity runs; torch's Philox stream cannot be matched), learning rates scaled so that three steps
    change the weights by far more than the parity tolerance."""
    from ddim_audio_b200.config import audio_config
    cfg = audio_config()
    cfg.training.snapshot_freq = 10 ** 9
    cfg.model.transformers.kwargs.hidden_dropout_prob = 0.0
    for g in (cfg.optimization.optimizer.transformer, cfg.optimization.optimizer.default):
        g.warmup = 0
        g.lr = g.lr * lr_scale
    return cfg


def test_packed_weights_follow_optimizer_step_and_ema(oracle_sd):
    """the fused optimizers and ``EMAHelper.ema`` write the fp32 masters through raw pointers / ``.data``;
    the bf16 GEMM operand copies must be rebuilt.  After two optimizer steps (a) every packed operand equals the bf16 image
    of its CURRENT master and (b) the forward equals the oracle loaded with ``model.state_dict()``; the same after the EMA
    weights are copied in."""
    import argparse
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200.runners.diffusion import Diffusion
    cfg = _train_cfg(lr_scale=10.0)
    m = Model(cfg)
    m.load_state_dict(oracle_sd, strict=True)
    r = Diffusion(argparse.Namespace(log_path=None), cfg)
    x0, _, _ = O.synthetic_inputs(2, 32, 83)
    w_before = m.down_modules[1][0].conv[0].weight.detach().clone()
    r.train(loader=[(x0, None)], model=m, max_steps=2)
    w_after = m.down_modules[1][0].conv[0].weight.detach()
    assert float((w_after - w_before).norm() / w_before.norm()) > 0.05       # the test's premise: the weights really moved

    def check(tag):
        m.eval()
        pk = m._packed_for(False)
        want = m.down_modules[1][0].conv[0].weight.detach().permute(0, 2, 3, 1).reshape(32, -1).to(torch.bfloat16)
        assert torch.equal(pk["down"][0]["blocks"][0]["w0"], want), tag
        want = m.transformer.encoder.layer[3].output.dense.weight.detach().to(torch.bfloat16)
        assert torch.equal(pk["fnet"][3]["w2"], want), tag
        x, _, t = O.synthetic_inputs(2, 32, 84)
        ref = O.OracleModel({k: v.detach().cpu() for k, v in m.state_dict().items()})(x, t)
        with torch.no_grad():
            out = m(x.to(DEV), t.to(DEV)).cpu()
        rel = rel_l2(out, ref)
        print(f"{tag}: forward vs oracle(model.state_dict()) rel-l2 {rel:.3e}")
        assert rel <= 2e-2, tag

    check("after 2 optimizer steps")
    with torch.no_grad():
        out_trained = m(x0.to(DEV), torch.tensor([10, 500], device=DEV))
    r.ema_helper.ema(m)                       # shadow = 0.9999-average, still close to the initial weights
    check("after EMAHelper.ema")
    with torch.no_grad():
        out_ema = m(x0.to(DEV), torch.tensor([10, 500], device=DEV))
    assert rel_l2(out_ema, out_trained) > 1e-2  # the EMA weights differ from the trained ones, and the forward saw it


def test_three_train_steps_vs_oracle_optimizers(oracle_sd):
    """N full train steps -- loss, backward, clip_grad_norm_ (runners/diffusion.py:155-160), AdaBelief
    (default group) + AdamW (transformer group) (:161-173), EMA (:182-183) -- against fp32 autograd of the oracle followed by
    the oracle's restatement of the reference optimizers.  Tolerance: loss rel <= 5e-3 at step 1 (same weights), <= 5e-2 at
    steps 2-3 (bf16 gradient noise flips the sign-like AdaBelief update of near-zero gradients); the parameter and EMA
    updates must point the same way (cosine >= 0.8 over all 47 M values).  A run on frozen weights must NOT match."""
    import argparse
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200.models.ema import EMAHelper
    from ddim_audio_b200.runners.diffusion import Diffusion
    cfg = _train_cfg(lr_scale=3.0)
    m = Model(cfg)
    m.load_state_dict(oracle_sd, strict=True)
    r = Diffusion(argparse.Namespace(log_path=None), cfg)
    opts, scheds, groups = r.build_training(m)
    ema = EMAHelper(mu=cfg.model.ema_rate)
    ema.register(m)
    al = O.alphas_cumprod()
    oc = cfg.optimization.optimizer
    sd = {k: v.clone() for k, v in oracle_sd.items()}
    trainable = [k for k, _ in m.named_parameters()]
    om = {k: torch.zeros_like(sd[k]) for k in trainable}
    ov = {k: torch.zeros_like(sd[k]) for k in trainable}
    shadow = {k: sd[k].clone() for k in trainable}
    losses, olosses, frozen = [], [], []
    for step in range(1, 4):
        x0, e, t = O.synthetic_inputs(2, 32, 90 + step)
        log = r.train_step(m, x0, opts, scheds, groups, ema, step, 0, t=t, e=e)
        losses.append(log["loss"])
        fl, _ = O.loss_and_grads(oracle_sd, x0, t, e, al)
        frozen.append(float(fl))
        ol, grads = O.loss_and_grads(sd, x0, t, e, al)
        olosses.append(float(ol))
        _, grads = O.clip_grad_norm(grads, 1.0)
        for k in trainable:
            if k.startswith("transformer."):
                c = oc.transformer
                sd[k], om[k], ov[k] = O.adamw_step(sd[k], grads[k], om[k], ov[k], step, c.lr, tuple(c.beta), c.eps, c.weight_decay)
            else:
                c = oc.default
                sd[k], om[k], ov[k] = O.adabelief_step(sd[k], grads[k], om[k], ov[k], step, c.lr, tuple(c.beta), c.eps, c.weight_decay)
        shadow = O.ema_update(shadow, sd, cfg.model.ema_rate)
    print("losses cuda", losses, "oracle", olosses, "oracle on frozen weights", frozen)
    assert abs(losses[0] - olosses[0]) <= 5e-3 * olosses[0]
    for a, b, f in zip(losses[1:], olosses[1:], frozen[1:]):
        assert abs(a - b) <= 5e-2 * b
        assert abs(b - f) > 2 * abs(a - b), "the updated-weight loss must be distinguishable from the frozen-weight one"
    dot = na = nb = 0.0
    for k, p in m.named_parameters():
        dg = (p.detach().cpu() - oracle_sd[k]).double().flatten()
        do = (sd[k] - oracle_sd[k]).double().flatten()
        dot += float(dg @ do); na += float(dg @ dg); nb += float(do @ do)
    cos = dot / (na * nb) ** 0.5
    print(f"parameter update cosine vs oracle optimizers: {cos:.4f}; |delta| cuda {na ** 0.5:.4e} oracle {nb ** 0.5:.4e}")
    assert cos >= 0.8 and 0.8 <= (na / nb) ** 0.5 <= 1.25
    dot = na = nb = 0.0
    for k in trainable:
        dg = (ema.shadow[k].cpu() - oracle_sd[k]).double().flatten()
        do = (shadow[k] - oracle_sd[k]).double().flatten()
        dot += float(dg @ do); na += float(dg @ dg); nb += float(do @ do)
    assert dot / (na * nb) ** 0.5 >= 0.8
