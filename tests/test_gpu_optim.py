"""-m gpu: the fused optimizer tail -- ``clip_grad_norm_``, ``AdamW``, ``AdaBelief`` of
``ddim_audio_b200.functions.optim`` -- against the golden parameters / state of the reference's own optimizers after three
steps (oracle/make_golden_optim.py), against torch's CUDA AdamW + clip on the full denoiser, and for state_dict compatibility.

Bar (floating point): parameters max-abs <= 2e-7 (one fp32 ulp at |p| ~ 1; the update itself is ~1e-3 so this is <= 2e-4 of a
step), first moment rtol 1e-5, second moment rtol 2e-5 -- torch's own kernels differ from its formula by the same amounts
(fused multiply-adds)."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import ddim_oracle as O

pytestmark = pytest.mark.gpu

CFG = {"adabelief": dict(lr=0.0003, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.00001, warmup=1000),
       "adamw": dict(lr=0.0005, betas=(0.9, 0.998), eps=1e-6, weight_decay=0.0001, warmup=10000)}


def _make(name, params):
    from ddim_audio_b200.functions import optim as fused

    cfg = CFG[name]
    kw = dict(lr=cfg["lr"], betas=cfg["betas"], eps=cfg["eps"], weight_decay=cfg["weight_decay"], amsgrad=False)
    if name == "adabelief":
        return fused.AdaBelief(params, weight_decouple=True, fixed_decay=False, rectify=False, clip_step=None, norm_ord=2, **kw), "exp_avg_var"
    return fused.AdamW(params, **kw), "exp_avg_sq"


@pytest.mark.parametrize("name", ["adabelief", "adamw"])
def test_vs_reference_golden(golden_dir, name):
    from ddim_audio_b200 import _C
    from ddim_audio_b200.functions import optim as fused

    g = np.load(os.path.join(golden_dir, "optim_steps.npz"))
    params0, grads = O.optim_case(int(g["seed"]))
    ps = {k: nn.Parameter(v.clone().cuda()) for k, v in params0.items()}
    opt, second = _make(name, list(ps.values()))
    w = CFG[name]["warmup"]
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: min(((1 + s) / w) ** -0.5, (1 + s) / w))
    for i, gr in enumerate(grads):
        for k, p in ps.items():
            p.grad = gr[k].clone().cuda()
        n0 = _C.launch_count()
        total = fused.clip_grad_norm_(list(ps.values()), 1.0)
        out = opt.step()
        assert _C.launch_count() - n0 == 3                     # sum of squares, scale, update -- whatever the tensor count
        sched.step()
        assert float(total) == pytest.approx(float(g[f"{name}__norms"][i]), rel=1e-6)
        if name == "adabelief":
            assert out == {"loss": None, "norm": None}         # AdaBelief.py:162 with clip_step None
    for k, p in ps.items():
        kk = k.replace(".", "__")
        st = opt.state[p]
        np.testing.assert_allclose(O.optim_probe(p.data.cpu()).numpy(), g[f"{name}__p__{kk}"], rtol=0, atol=2e-7)
        np.testing.assert_allclose(O.optim_probe(st["exp_avg"].cpu()).numpy(), g[f"{name}__m__{kk}"], rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(O.optim_probe(st[second].cpu()).numpy(), g[f"{name}__v__{kk}"], rtol=2e-5, atol=1e-12)
    assert int(opt.state[next(iter(ps.values()))]["step"]) == len(grads)


def test_full_model_vs_torch_adamw_and_state_dict(oracle_sd):
    """Every denoiser parameter, 2 steps: fused clip + AdamW against torch's own CUDA clip_grad_norm_ + AdamW; then the fused
    optimizer's state_dict loads into torch.optim.AdamW and back (checkpoint compatibility, runners/diffusion.py:185-199)."""
    from ddim_audio_b200.functions import optim as fused

    g = torch.Generator(device="cuda").manual_seed(5)
    a = [nn.Parameter(v.clone().cuda()) for k, v in oracle_sd.items() if k != "temb.te"]
    b = [nn.Parameter(p.data.clone()) for p in a]
    kw = dict(lr=5e-4, betas=(0.9, 0.998), eps=1e-6, weight_decay=1e-4)
    fo, to = fused.AdamW(a, **kw), torch.optim.AdamW(b, **kw)
    for _ in range(2):
        for p, q in zip(a, b):
            p.grad = torch.randn(p.shape, device="cuda", generator=g) * 0.05
            q.grad = p.grad.clone()
        n1 = fused.clip_grad_norm_(a, 1.0)
        n2 = torch.nn.utils.clip_grad_norm_(b, 1.0)
        assert float(n1) == pytest.approx(float(n2), rel=1e-5)
        fo.step()
        to.step()
    worst = max(float((p.data - q.data).abs().max()) for p, q in zip(a, b))
    assert worst <= 2e-7, worst
    sd = fo.state_dict()
    to2 = torch.optim.AdamW([nn.Parameter(p.data.clone()) for p in a], **kw)
    to2.load_state_dict(sd)
    fo2 = fused.AdamW([nn.Parameter(p.data.clone()) for p in a], **kw)
    fo2.load_state_dict(to.state_dict())
    assert int(fo2.state[fo2.param_groups[0]["params"][0]]["step"]) == 2


def test_errors():
    from ddim_audio_b200.functions import optim as fused

    p = nn.Parameter(torch.zeros(8))
    p.grad = torch.ones(8)
    with pytest.raises(RuntimeError):
        fused.clip_grad_norm_([p], 1.0)                        # CPU: no CPU path
    with pytest.raises(RuntimeError):
        fused.AdamW([p]).step()
    with pytest.raises(NotImplementedError):
        fused.AdaBelief([p], clip_step=0.1)
    with pytest.raises(NotImplementedError):
        fused.AdaBelief([p], rectify=True)


def test_misaligned_views_take_the_scalar_path_with_identical_results():
    """Gradients (and EMA shadows) that are views of a foreign flat buffer at odd 4-byte offsets: same bits as the aligned run."""
    from ddim_audio_b200.functions import optim as fused
    from ddim_audio_b200.models.ema import EMAHelper

    g = torch.Generator(device="cuda").manual_seed(11)
    shapes = [(5, 7), (70001,), (3,)]
    base = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    grads = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    res = []
    for shift in (0, 2):
        flat = torch.zeros(sum(t.numel() for t in grads) + 16, device="cuda")
        ps, off = [], shift
        for b, gr in zip(base, grads):
            p = nn.Parameter(b.clone())
            view = flat[off:off + gr.numel()].view(gr.shape)
            view.copy_(gr)
            p.grad = view
            off += gr.numel()
            ps.append(p)
        assert shift == 0 or all(p.grad.data_ptr() % 16 != 0 for p in ps[:2])
        norm = fused.clip_grad_norm_(ps, 1.0)
        opt = fused.AdaBelief(ps, lr=1e-3, weight_decay=1e-5)
        opt.step()
        res.append((float(norm), [p.data.clone() for p in ps], [p.grad.clone() for p in ps]))
    assert res[0][0] == pytest.approx(res[1][0], rel=1e-6)
    for a, b in zip(res[0][1] + res[0][2], res[1][1] + res[1][2]):
        assert float((a - b).abs().max()) <= 1e-7               # the norm's fp64 atomics may differ in the last bit -> coef
    m = nn.Module()
    m.w = nn.Parameter(base[1].clone())
    h = EMAHelper(mu=0.99)
    h.register(m)
    flat = torch.zeros(70001 + 8, device="cuda")
    flat[1:70002].copy_(h.shadow["w"])
    h.shadow["w"] = flat[1:70002]
    want = (1.0 - 0.99) * m.w.data + 0.99 * h.shadow["w"]
    h._table = None
    h.update(m)
    assert torch.equal(h.shadow["w"], want)
