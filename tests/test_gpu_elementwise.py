"""-m gpu: HBM-bound kernels of libddim_b200 against plain fp32 torch ops on the same inputs."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import DEV, assert_close_bf16, assert_stats, group_stats, rel_l2  # noqa: E402


@pytest.fixture(scope="module")
def ops():
    from ddim_audio_b200 import ops as o
    return o


@pytest.mark.parametrize("with_noise,snap", [(False, False), (True, True), (False, True)])
def test_ddim_step_matches_reference_op_order(ops, with_noise, snap):
    # functions/denoising.py:27,41-43 replayed with torch in-place ops (fp32)
    g = torch.Generator(device=DEV).manual_seed(3)
    xt = torch.randn(3, 2, 64, 256, device=DEV, generator=g)
    et = torch.randn_like(xt)
    z = torch.randn_like(xt) if with_noise else None
    at, at_next = 0.4321, 0.6789
    eta = 1.0 if with_noise else 0.0
    c1 = eta * ((1 - at / at_next) * (1 - at_next) / (1 - at)) ** 0.5
    c2 = ((1 - at_next) - c1 ** 2) ** 0.5
    ref = xt.clone()
    ref.add_(et, alpha=-((1 - at) ** 0.5)).div_(at ** 0.5)
    ref_x0 = ref.clone()
    ref.mul_(at_next ** 0.5).add_(et, alpha=c2)
    if with_noise:
        ref.add_(z, alpha=c1)
    x0b = torch.empty_like(xt) if snap else None
    xsb = torch.empty_like(xt) if snap else None
    out = ops.ddim_step(xt, et, z, at, at_next, c1, c2, x0_out=x0b, xs_out=xsb)
    assert out.data_ptr() == xt.data_ptr()
    torch.testing.assert_close(xt, ref, rtol=2e-6, atol=2e-6)
    if snap:
        torch.testing.assert_close(x0b, ref_x0, rtol=2e-6, atol=2e-6)
        assert torch.equal(xsb, xt)


def test_qsample_and_loss(ops):
    g = torch.Generator(device=DEV).manual_seed(4)
    B = 5
    x0 = torch.randn(B, 2, 32, 256, device=DEV, generator=g)
    e = torch.randn_like(x0)
    a = torch.linspace(0.999, 0.01, 1000, device=DEV)
    t = torch.tensor([0, 999, 500, 17, 333], device=DEV)
    av = a.index_select(0, t).view(-1, 1, 1, 1)
    ref = x0 * av.sqrt() + e * (1.0 - av).sqrt()
    torch.testing.assert_close(ops.qsample(x0, e, a, t), ref, rtol=1e-6, atol=1e-6)
    out = torch.randn_like(e)
    per = ops.sqerr_loss(e, out)
    want = (e.double() - out.double()).square().sum(dim=(1, 2, 3))
    torch.testing.assert_close(per, want, rtol=1e-5, atol=0)
    gvec = torch.randn(B, device=DEV)
    go = ops.sqerr_loss_bwd(e, out, gvec)
    torch.testing.assert_close(go, -2 * (e - out) * gvec.view(-1, 1, 1, 1), rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("C,P", [(32, 4096), (64, 1000), (96, 512), (128, 300), (192, 96), (256, 8)])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_gn_apply(ops, C, P, mode):
    g = torch.Generator(device=DEV).manual_seed(C + mode)
    B = 3
    x = (torch.randn(B, P, 1, C, device=DEV, generator=g) * 1.5 + 0.3).to(torch.bfloat16)
    res = torch.randn(B, P, 1, C, device=DEV, generator=g).to(torch.bfloat16) if mode == 2 else None
    gamma = 1 + 0.1 * torch.randn(C, device=DEV, generator=g)
    beta = 0.1 * torch.randn(C, device=DEV, generator=g) if mode != 2 else None
    st = group_stats(x)
    xn = x.float().permute(0, 3, 1, 2)
    ref = torch.nn.functional.group_norm(xn, 8, gamma, beta, 1e-6)
    if mode == 1:
        ref = torch.nn.functional.silu(ref)
    if mode == 2:
        ref = ref + res.float().permute(0, 3, 1, 2)
    ref = ref.permute(0, 2, 3, 1)
    so = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV) if mode == 2 else None
    y = ops.gn_apply(x, st, gamma, beta, mode, 1e-6, res=res, stats_out=so)
    assert_close_bf16(y, ref, f"gn C={C} mode={mode}")
    if mode == 2:
        assert_stats(so, ref, f"gn stats C={C}")
        # second residual (the last block also adds conv_out's skip operand): y = res + res2 + gn(x), statistics of that sum
        res2 = torch.randn(B, P, 1, C, device=DEV, generator=g).to(torch.bfloat16)
        so2 = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
        y2 = ops.gn_apply(x, st, gamma, beta, mode, 1e-6, res=res, stats_out=so2, res2=res2)
        ref2 = ref + res2.float()
        assert_close_bf16(y2, ref2, f"gn C={C} mode=2 + res2")
        assert_stats(so2, ref2, f"gn stats C={C} + res2")


@pytest.mark.parametrize("B,T,F", [(1, 32, 256), (3, 64, 256), (2, 1, 256), (1, 77, 256), (2, 300, 256), (2, 45, 128), (70, 9, 256), (2, 33, 96), (1, 2500, 256)])
def test_conv_in_out(ops, B, T, F):
    """F % 128 == 0 runs the tensor-core kernels (conv_wide.cuh prologue 3 / conv_out_tc.cuh), ragged F the strip kernels."""
    g = torch.Generator(device=DEV).manual_seed(7)
    x = torch.randn(B, 2, T, F, device=DEV, generator=g)
    w = torch.randn(32, 2, 3, 3, device=DEV, generator=g) * 0.3
    b = torch.randn(32, device=DEV, generator=g) * 0.1
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    y = ops.conv_in(x, w, b, st)
    ref = torch.nn.functional.conv2d(x, w, b, padding=1).permute(0, 2, 3, 1)
    assert_close_bf16(y, ref, "conv_in")
    assert_stats(st, ref, "conv_in stats")
    h = torch.randn(B, T, F, 32, device=DEV, generator=g).to(torch.bfloat16)
    sk = torch.randn(B, T, F, 32, device=DEV, generator=g).to(torch.bfloat16)
    w2 = torch.randn(2, 32, 3, 3, device=DEV, generator=g) * 0.1
    b2 = torch.randn(2, device=DEV, generator=g) * 0.1
    for skip in (None, sk):
        hin = h.float() + (skip.float() if skip is not None else 0)
        ref2 = torch.nn.functional.conv2d(hin.permute(0, 3, 1, 2), w2, b2, padding=1)
        out = ops.conv_out(h, skip, w2, b2)
        assert rel_l2(out, ref2) < 1e-5


def test_temb_mlp(ops):
    g = torch.Generator(device=DEV).manual_seed(8)
    te = torch.randn(1000, 128, device=DEV, generator=g)
    w0 = torch.randn(512, 128, device=DEV, generator=g) * 0.1
    b0 = torch.randn(512, device=DEV, generator=g) * 0.1
    w2 = torch.randn(4416, 512, device=DEV, generator=g) * 0.05
    b2 = torch.randn(4416, device=DEV, generator=g) * 0.1
    t = torch.tensor([0, 999, 5, 5, 77], device=DEV)
    h = ops.linear_small(te, w0, b0, act=1, idx=t)
    ref = torch.nn.functional.silu(torch.nn.functional.linear(te.index_select(0, t), w0, b0))
    torch.testing.assert_close(h, ref, rtol=2e-5, atol=2e-5)
    o = ops.linear_small(h, w2, b2, act=0)
    torch.testing.assert_close(o, torch.nn.functional.linear(ref, w2, b2), rtol=2e-5, atol=5e-5)


@pytest.mark.parametrize("B,S", [(1, 1), (2, 3), (3, 32)])
def test_token_repacks(ops, B, S):
    g = torch.Generator(device=DEV).manual_seed(9)
    C, W = 256, 8
    D = C * W
    x = torch.randn(B, S, W, C, device=DEV, generator=g).to(torch.bfloat16)
    te = torch.randn(S, D, device=DEV, generator=g)
    gamma = 1 + 0.1 * torch.randn(D, device=DEV, generator=g)
    beta = 0.1 * torch.randn(D, device=DEV, generator=g)
    tok = x.float().permute(0, 1, 3, 2).reshape(B, S, D)          # feature index c*W + f (models/diffusion.py:273-275)
    ref = torch.nn.functional.layer_norm(tok + te, (D,), gamma, beta, 1e-6).reshape(B * S, D)
    assert_close_bf16(ops.tokens_ln(x, te, gamma, beta, 1e-6), ref, "tokens_ln")
    tk = torch.randn(B * S, D, device=DEV, generator=g)
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    y = ops.tokens_to_nhwc(tk, x, st)
    ref2 = tk.reshape(B, S, C, W).permute(0, 1, 3, 2) + x.float()
    assert_close_bf16(y, ref2, "tokens_to_nhwc")
    assert_stats(st, ref2, "tokens_to_nhwc stats")


@pytest.mark.parametrize("rows", [1, 7, 96])
def test_layernorm512_and_split(ops, rows):
    g = torch.Generator(device=DEV).manual_seed(10)
    a = torch.randn(rows, 512, device=DEV, generator=g) * 3
    b = torch.randn(rows, 512, device=DEV, generator=g)
    gamma = 1 + 0.1 * torch.randn(512, device=DEV, generator=g)
    beta = 0.1 * torch.randn(512, device=DEV, generator=g)
    ybf = torch.zeros(rows, 1024, dtype=torch.bfloat16, device=DEV)
    y = ops.layernorm512(a, b, gamma, beta, 1e-6, ybf, True)
    ref = torch.nn.functional.layer_norm(a + b, (512,), gamma, beta, 1e-6)
    torch.testing.assert_close(y, ref, rtol=1e-5, atol=1e-5)
    hi, lo = ybf[:, :512].float(), ybf[:, 512:].float()
    assert torch.equal(hi, y.to(torch.bfloat16).float())
    assert float((hi + lo - y).abs().max()) <= float(y.abs().max()) * 2.0 ** -15
    y2 = ops.layernorm512(a, None, gamma, beta, 1e-6, ybf, False)
    torch.testing.assert_close(y2, torch.nn.functional.layer_norm(a, (512,), gamma, beta, 1e-6), rtol=1e-5, atol=1e-5)
    buf = torch.zeros(rows, 1024, dtype=torch.bfloat16, device=DEV)
    ops.split_bf16(a, buf, True)
    assert torch.equal(buf[:, :512].float(), a.to(torch.bfloat16).float())
    assert float((buf[:, :512].float() + buf[:, 512:].float() - a).abs().max()) <= float(a.abs().max()) * 2.0 ** -15
