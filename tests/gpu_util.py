"""Shared helpers for the -m gpu tests (torch fp32 CUDA ops are the per-kernel reference; the oracle is the
model-level one)."""
import torch

# the per-kernel references must be true fp32: torch lets cuDNN/cuBLAS use TF32 for convolutions by default
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

DEV = "cuda"


def rel_l2(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def assert_close_bf16(got, ref, name="", rel=4e-3, max_rel=1.6e-2):
    """got is bf16-stored output of an fp32-accumulated kernel, ref the fp32 torch result on the same bf16 inputs:
    rel-L2 <= ~bf16 rounding noise (2^-9 rms) and max-abs <= 2 bf16 ulps of the largest value."""
    got, ref = got.float(), ref.float()
    assert got.shape == ref.shape, (name, got.shape, ref.shape)
    assert torch.isfinite(got).all(), name
    r = rel_l2(got, ref)
    m = float((got - ref).abs().max()) / max(float(ref.abs().max()), 1e-30)
    assert r <= rel and m <= max_rel, f"{name}: rel-l2 {r:.3e} (<= {rel}), max-abs/absmax {m:.3e} (<= {max_rel})"


def group_stats(x_nhwc, groups=8):
    """float64 [B, 8, 2] (sum, sumsq) per (sample, group) of an NHWC tensor."""
    B, C = x_nhwc.shape[0], x_nhwc.shape[-1]
    v = x_nhwc.double().reshape(B, -1, groups, C // groups)
    return torch.stack([v.sum(dim=(1, 3)), (v * v).sum(dim=(1, 3))], dim=-1)


def assert_stats(got, x_nhwc, name=""):
    want = group_stats(x_nhwc)
    n = x_nhwc.numel() / (x_nhwc.shape[0] * 8)
    # sums are compared on the scale of n * rms
    scale = (want[..., 1] / n).sqrt().clamp_min(1e-6) * n
    err = ((got[..., 0] - want[..., 0]).abs() / scale).max()
    err2 = ((got[..., 1] - want[..., 1]).abs() / want[..., 1].clamp_min(1e-12)).max()
    assert float(err) < 2e-3 and float(err2) < 2e-3, f"{name}: stats mismatch sum {float(err):.3e} sumsq {float(err2):.3e}"
