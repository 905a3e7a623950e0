"""-m gpu: the captured train step (models/training.py::GraphedLossBackward) against the eagerly issued one on the same batch:
loss and every parameter gradient must agree to the path's own run-to-run noise (fp32 atomics in the weight-gradient kernels), the
re-pack inside the graph must follow an optimizer step, and dropout must draw a new mask per replay."""
import pytest
import torch

from oracle import ddim_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _setup(p_drop):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model

    torch.manual_seed(3)
    cfg = audio_config()
    m = Model(cfg).to(DEV).train()
    sd = O.make_state_dict(seed=5)
    m.load_state_dict(sd, strict=True)
    m.transformer.embedding.dropout_p = p_drop       # the library's counter-based dropout reads this (models/training.py)
    x, e, t = O.synthetic_inputs(2, 64, 17)
    a = O.alphas_cumprod().to(DEV)
    return m, x.to(DEV), e.to(DEV), t.to(DEV), a


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20))


def test_graphed_step_equals_eager_and_follows_the_optimizer():
    from ddim_audio_b200.functions.losses import loss_registry
    from ddim_audio_b200.models.training import GraphedLossBackward

    m, x, e, t, a = _setup(0.0)
    fn = loss_registry["simple"]
    for p in m.parameters():
        p.grad = None
    le = fn(m, x, t, e, a)
    le.backward()
    ge = [p.grad.clone() for p in m.parameters()]
    gl = GraphedLossBackward(m, a, x, t, e)
    lg = gl(x, t, e)
    le = le.detach()
    assert abs(float(lg) - float(le)) <= 1e-4 * abs(float(le))
    num = sum(float(((p.grad - g).float() ** 2).sum()) for p, g in zip(m.parameters(), ge))
    den = sum(float((g.float() ** 2).sum()) for g in ge)
    assert (num / den) ** 0.5 < 1e-3
    # change the fp32 masters the way the fused optimizers do (raw writes + epoch bump): the next replay must see them
    from ddim_audio_b200 import _C
    with torch.no_grad():
        for p in m.parameters():
            p.data.mul_(0.9)
    _C.bump_param_epoch()
    l2 = float(gl(x, t, e))
    for p in m.parameters():
        p.grad = None
    gl_e = fn(m, x, t, e, a)
    assert abs(l2 - float(gl_e)) <= 1e-4 * abs(float(gl_e)) and abs(l2 - float(le)) > 1e-3 * abs(float(le))


def test_graphed_step_draws_a_new_dropout_mask_per_replay():
    from ddim_audio_b200.functions.losses import loss_registry
    from ddim_audio_b200.models.training import GraphedLossBackward

    m, x, e, t, a = _setup(0.1)
    gl = GraphedLossBackward(m, a, x, t, e)
    l1 = float(gl(x, t, e))
    l2 = float(gl(x, t, e))
    assert l1 != l2 and abs(l1 - l2) < 0.2 * abs(l1)
