"""Training loss behind the reference's ``noise_estimation_loss(model, x0, t, e, a, keepdim=False)``
(functions/losses.py:4-18): forward diffusion and the per-sample squared-error reduction are one kernel each.
"""
from __future__ import annotations

import torch

from .. import ops


class _SqErr(torch.autograd.Function):
    """per[b] = sum_{c,t,f} (e - out)^2 (functions/losses.py:15-18); d per[b] / d out = -2 (e - out)."""

    @staticmethod
    def forward(ctx, e, output):
        ctx.save_for_backward(e, output)
        return ops.sqerr_loss(e, output).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        e, output = ctx.saved_tensors
        return None, ops.sqerr_loss_bwd(e, output, g.to(torch.float32).contiguous())


def noise_estimation_loss(model, x0: torch.Tensor, t: torch.LongTensor, e: torch.Tensor, a: torch.Tensor, keepdim=False):
    if not x0.is_cuda:
        raise RuntimeError("ddim_audio_b200.noise_estimation_loss needs CUDA tensors: there is no CPU path")
    x0 = x0.contiguous().float()
    e = e.to(x0.device).contiguous().float()
    t = t.to(x0.device).long().contiguous()
    a = a.to(device=x0.device, dtype=torch.float32).contiguous()
    x = ops.qsample(x0, e, a, t)                                                      # :12-13
    output = model(x, t)                                                              # :14
    per = _SqErr.apply(e, output.contiguous())
    return per if keepdim else per.mean(dim=0)


loss_registry = {
    "simple": noise_estimation_loss,
}
