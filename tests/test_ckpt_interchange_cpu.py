"""CPU, build container only (needs /root/reference; skipped where it is absent, e.g. on the GPU box): checkpoint interchange with
the reference.  The reference's train_step saves ``[model_sd, optimizer_sd, epoch, step, ema_sd]``
(runners/diffusion.py:185-199; ``optimizer`` is its leaked loop variable = the last group's object, the AdaBelief of the default group)
and ``sample()`` loads ``states[0]`` + copies the EMA shadow in (:289-316).

  * a file written by the reference's OWN objects (its Model, ``clip_opt.AdaBelief`` with populated state, its ``EMAHelper``) loads into
    ``Diffusion.load_model`` (parameters == the reference's EMA weights) and into the fused ``AdaBelief`` (same per-parameter state);
  * a file written by THIS package's objects loads into the reference's Model / AdaBelief / EMAHelper, and the reference optimizer
    can take a step from it.
"""
import argparse
import os

import pytest
import torch

from oracle import reference_loader as R

pytestmark = pytest.mark.skipif(not R.available(), reason="reference tree not present")


def _ref_objects(seed):
    torch.manual_seed(seed)
    model, config = R.load_model(None, eval_mode=False)
    with R._ref_cwd():
        from functions import get_optimizer
        from models.ema import EMAHelper
    from ddim_audio_b200.runners.diffusion import classify_group    # same grouping rule as runners/diffusion.py:71-87 (CPU-tested elsewhere)
    groups = classify_group(config.optimization.optimizer, model)
    with R._ref_cwd():
        opts = {name: get_optimizer(g.config, g.params) for name, g in groups.items()}
    ema = EMAHelper(mu=config.model.ema_rate)
    ema.register(model)
    return model, config, opts, ema


def _fake_step(model, opts, ema, seed):
    g = torch.Generator().manual_seed(seed)
    for p in model.parameters():
        p.grad = 1e-3 * torch.randn(p.shape, generator=g)
    for o in opts.values():
        o.step()
    ema.update(model)


def test_reference_checkpoint_loads_here_and_back(tmp_path):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.functions import optim as fused
    from ddim_audio_b200.models.ema import EMAHelper
    from ddim_audio_b200.runners.diffusion import Diffusion, classify_group

    # ---- written by the reference's own objects
    rmodel, rconfig, ropts, rema = _ref_objects(3)
    _fake_step(rmodel, ropts, rema, 4)
    _fake_step(rmodel, ropts, rema, 5)
    names = list(ropts)
    assert names[-1] == "default" and type(ropts["default"]).__name__ == "AdaBelief"
    optimizer = ropts[names[-1]]                                    # the reference's leaked loop variable
    path = os.path.join(tmp_path, "ckpt.pth")
    torch.save([rmodel.state_dict(), optimizer.state_dict(), 7, 1234, rema.state_dict()], path)

    # ---- loaded here
    cfg = audio_config("torch.FloatTensor")
    runner = Diffusion(argparse.Namespace(log_path=str(tmp_path)), cfg, device="cpu")
    model = runner.load_model()                                     # ckpt.pth from args.log_path, EMA weights copied in (:309-313)
    assert list(model.state_dict().keys()) == list(rmodel.state_dict().keys())
    shadow = rema.state_dict()
    for name, p in model.named_parameters():
        assert torch.equal(p.detach(), shadow[name]), name
    states = torch.load(path, map_location="cpu")
    assert states[2:4] == [7, 1234]
    groups = classify_group(cfg.optimization.optimizer, model)
    c = groups["default"].config
    opt = fused.AdaBelief(groups["default"].params, lr=c.lr, betas=tuple(c.beta), eps=c.eps, weight_decay=c.weight_decay)
    opt.load_state_dict(states[1])
    ref_state = optimizer.state_dict()["state"]
    assert len(opt.state_dict()["state"]) == len(ref_state) == len(groups["default"].params)
    for i, st in opt.state_dict()["state"].items():
        assert set(st) == {"step", "exp_avg", "exp_avg_var"} and int(st["step"]) == 2
        assert torch.equal(st["exp_avg"], ref_state[i]["exp_avg"]) and torch.equal(st["exp_avg_var"], ref_state[i]["exp_avg_var"])

    # ---- written here, loaded by the reference's classes
    ema = EMAHelper(mu=cfg.model.ema_rate)
    ema.register(model)
    with torch.no_grad():
        for p in model.parameters():
            p.add_(0.01)                                            # make model weights and shadow differ
    back = os.path.join(tmp_path, "ckpt_back.pth")
    torch.save([model.state_dict(), opt.state_dict(), 8, 1235, ema.state_dict()], back)
    r2model, _, r2opts, r2ema = _ref_objects(9)
    st2 = torch.load(back, map_location="cpu")
    r2model.load_state_dict(st2[0], strict=True)
    r2opts["default"].load_state_dict(st2[1])
    r2ema.load_state_dict(st2[-1])
    r2ema.ema(r2model)                                              # what the reference's sample() does (:309-313)
    for name, p in r2model.named_parameters():
        assert torch.equal(p.detach(), ema.shadow[name]), name
    _fake_step(r2model, r2opts, r2ema, 6)                           # the reference optimizer continues from the loaded state
    assert all(int(s["step"]) == 3 for s in r2opts["default"].state_dict()["state"].values())
