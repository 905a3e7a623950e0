"""-m gpu: the backward-pass kernels (csrc/gemm_tn.cuh, csrc/train_elementwise.cu) one by one against fp32 torch autograd
(TF32 off) on the same bf16-rounded operands.  Weight gradients are fp32 sums of bf16 products: rel-L2 <= 1e-3 (the order
of the split-K atomics is the only noise); bf16-stored gradients: the bf16 bounds of gpu_util.assert_close_bf16."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from gpu_util import DEV, assert_close_bf16, group_stats, rel_l2  # noqa: E402


@pytest.fixture(scope="module")
def ops():
    from ddim_audio_b200 import ops as o
    return o


def _bf(t):
    return t.to(torch.bfloat16)


def _nchw(t):
    return t.float().permute(0, 3, 1, 2)


WG_CASES = [(32, 256, 8, 1), (32, 256, 5, 2), (32, 256, 1, 1), (32, 256, 2, 1), (32, 256, 131, 2), (64, 128, 1, 2), (64, 128, 77, 2), (64, 256, 9, 1), (32, 128, 300, 1), (64, 128, 16, 3), (96, 64, 8, 1), (128, 32, 12, 2), (192, 16, 10, 1), (256, 8, 4, 3), (256, 8, 1, 1)]


@pytest.mark.parametrize("C,W,T,B", WG_CASES)
def test_conv3x3_wgrad(ops, C, W, T, B):
    g = torch.Generator(device=DEV).manual_seed(C + W + T)
    x = _bf(torch.randn(B, T, W, C, device=DEV, generator=g))
    dy = _bf(torch.randn(B, T, W, C, device=DEV, generator=g))
    ref = torch.nn.grad.conv2d_weight(_nchw(x), (C, C, 3, 3), _nchw(dy), padding=1)
    dw = torch.zeros(C, C, 3, 3, device=DEV)
    ops.conv3x3_wgrad(x, dy, dw)
    assert rel_l2(dw, ref) < 1e-3, rel_l2(dw, ref)
    ops.conv3x3_wgrad(x, dy, dw)                         # accumulates
    assert rel_l2(dw, 2 * ref) < 1e-3


@pytest.mark.parametrize("Cin,Cout,W,T,B", [(32, 64, 256, 8, 2), (32, 64, 256, 2, 1), (32, 64, 256, 70, 2), (32, 64, 256, 4, 3), (64, 96, 128, 16, 1), (96, 128, 64, 8, 2), (128, 192, 32, 4, 1), (192, 256, 16, 2, 3)])
def test_down_up_wgrad(ops, Cin, Cout, W, T, B):
    g = torch.Generator(device=DEV).manual_seed(Cin + W)
    # Downsample: x [B,T,W,Cin] -> y [B,T/2,W/2,Cout]
    x = _bf(torch.randn(B, T, W, Cin, device=DEV, generator=g))
    dy = _bf(torch.randn(B, T // 2, W // 2, Cout, device=DEV, generator=g))
    ref = torch.nn.grad.conv2d_weight(_nchw(x), (Cout, Cin, 4, 4), _nchw(dy), stride=2, padding=1)
    dw = torch.zeros(Cout, Cin, 4, 4, device=DEV)
    ops.conv_down4x4_wgrad(x, dy, dw)
    assert rel_l2(dw, ref) < 1e-3, rel_l2(dw, ref)
    # Upsample: xs [B,T/2,W/2,Cout] -> ys [B,T,W,Cin], ConvTranspose2d weight [Cout, Cin, 4, 4] (in, out)
    xs = _bf(torch.randn(B, T // 2, W // 2, Cout, device=DEV, generator=g))
    dys = _bf(torch.randn(B, T, W, Cin, device=DEV, generator=g))
    wt = torch.zeros(Cout, Cin, 4, 4, device=DEV, requires_grad=True)
    F.conv_transpose2d(_nchw(xs), wt, stride=2, padding=1).backward(_nchw(dys))
    dwt = torch.zeros(Cout, Cin, 4, 4, device=DEV)
    ops.conv_up4x4_wgrad(xs, dys, dwt)
    assert rel_l2(dwt, wt.grad) < 1e-3, rel_l2(dwt, wt.grad)


@pytest.mark.parametrize("rows,N,K", [(1024, 2048, 512), (1024, 512, 2048), (32, 512, 512), (300, 2048, 512), (96, 512, 2048)])
def test_linear_wgrad(ops, rows, N, K):
    g = torch.Generator(device=DEV).manual_seed(rows + N)
    a = _bf(torch.randn(rows, K, device=DEV, generator=g))
    gy = _bf(torch.randn(rows, N, device=DEV, generator=g))
    dw = torch.zeros(N, K, device=DEV)
    ops.linear_wgrad(a, gy, dw)
    ref = gy.float().t() @ a.float()
    assert rel_l2(dw, ref) < 1e-3, rel_l2(dw, ref)
    # strided operand: only the first K columns of a [rows, 2K] buffer (hi | lo planes)
    buf = torch.cat([a, _bf(torch.randn(rows, K, device=DEV, generator=g))], dim=1)
    dw2 = torch.zeros(N, K, device=DEV)
    ops.linear_wgrad(buf, gy, dw2, K=K)
    assert rel_l2(dw2, ref) < 1e-3


@pytest.mark.parametrize("C,P,B", [(32, 2048, 2), (64, 1024, 3), (96, 640, 1), (128, 384, 2), (192, 160, 2), (256, 32, 3)])
@pytest.mark.parametrize("gtype", [0, 1])
def test_gn_backward(ops, C, P, B, gtype):
    eps = 1e-6
    g = torch.Generator(device=DEV).manual_seed(C + P + gtype)
    z = _bf(torch.randn(B, P, 1, C, device=DEV, generator=g) * 1.5 + 0.3)
    gy = _bf(torch.randn(B, P, 1, C, device=DEV, generator=g))
    gamma = torch.randn(C, device=DEV, generator=g)
    beta = torch.randn(C, device=DEV, generator=g) * 0.3
    add = _bf(torch.randn(B, P, 1, C, device=DEV, generator=g))
    zf = z.float().requires_grad_(True)
    gm, bt = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    zin = zf if gtype == 0 else F.silu(zf)
    y = F.group_norm(zin.permute(0, 3, 1, 2), 8, gm, bt, eps)
    if gtype == 0:
        y = F.silu(y)
    y.backward(gy.float().permute(0, 3, 1, 2))
    stats = group_stats(zin.detach())
    R = ops.gn_bwd_reduce(gy, z, stats, gamma, beta, gtype, eps)
    out, sums = ops.gn_bwd_apply(gy, z, stats, R, gamma, beta, gtype, eps, add=add if gtype == 0 else None, sums=(gtype == 1))
    ref = zf.grad + (add.float() if gtype == 0 else 0)
    assert_close_bf16(out, ref, f"gn_bwd type {gtype} C={C}", rel=6e-3, max_rel=2.5e-2)
    dg, db = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.gn_bwd_param(R, stats, dg, db, P, eps)
    assert rel_l2(dg, gm.grad) < 5e-3 and rel_l2(db, bt.grad) < 5e-3, (rel_l2(dg, gm.grad), rel_l2(db, bt.grad))
    if sums is not None:
        want = out.float().sum(dim=(1, 2))
        assert float((sums.float() - want).abs().max()) <= 1e-3 * float(want.abs().max()) + 1e-3
        tb = torch.zeros(B, 100 + C, device=DEV)
        bias = torch.zeros(C, device=DEV)
        ops.sums_scatter(sums, out_b=tb[:, 100:], ldb=100 + C, out_c=bias)
        assert rel_l2(tb[:, 100:], want) < 1e-4 and rel_l2(bias, want.sum(0)) < 1e-4 and float(tb[:, :100].abs().max()) == 0.0


@pytest.mark.parametrize("C,W,T,B", [(32, 256, 4, 2), (96, 64, 8, 1), (128, 32, 4, 2)])
def test_training_forward_modes(ops, C, W, T, B):
    """conv epilogue act=3 (store the pre-activation, statistics of its SiLU) and gn_apply modes 3 / 4."""
    g = torch.Generator(device=DEV).manual_seed(C)
    x = _bf(torch.randn(B, T, W, C, device=DEV, generator=g))
    w = _bf(torch.randn(C, C, 3, 3, device=DEV, generator=g) / math.sqrt(9 * C))
    bias = torch.randn(C, device=DEV, generator=g) * 0.2
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    u = ops.conv3x3(x, w.permute(0, 2, 3, 1).reshape(C, -1).contiguous(), bias=bias, act=3, stats=st)
    ref_u = F.conv2d(_nchw(x), w.float(), bias, padding=1).permute(0, 2, 3, 1)
    assert_close_bf16(u, ref_u, "pre-activation")
    want = group_stats(F.silu(u.float()))
    assert float(((st - want).abs() / want.abs().clamp_min(1.0)).max()) < 2e-3
    gamma = torch.randn(C, device=DEV, generator=g)
    beta = torch.randn(C, device=DEV, generator=g) * 0.3
    res = _bf(torch.randn(B, T, W, C, device=DEV, generator=g))
    ref = F.group_norm(F.silu(u.float()).permute(0, 3, 1, 2), 8, gamma, beta, 1e-6).permute(0, 2, 3, 1)
    assert_close_bf16(ops.gn_apply(u, want, gamma, beta, 3, 1e-6), ref, "gn(silu(u))", rel=5e-3, max_rel=2e-2)
    so = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    y = ops.gn_apply(u, want, gamma, None, 4, 1e-6, res=res, stats_out=so)
    ref2 = F.group_norm(F.silu(u.float()).permute(0, 3, 1, 2), 8, gamma, None, 1e-6).permute(0, 2, 3, 1) + res.float()
    assert_close_bf16(y, ref2, "res + gn(silu(u))", rel=5e-3, max_rel=2e-2)


@pytest.mark.parametrize("rows,two", [(8, True), (1024, True), (257, False)])
def test_layernorm512_bwd(ops, rows, two):
    g = torch.Generator(device=DEV).manual_seed(rows)
    a = torch.randn(rows, 512, device=DEV, generator=g)
    b = torch.randn(rows, 512, device=DEV, generator=g) * 3 if two else None
    dy1 = torch.randn(rows, 512, device=DEV, generator=g)
    dy2 = torch.randn(rows, 512, device=DEV, generator=g) if two else None
    gamma = torch.randn(512, device=DEV, generator=g)
    s = (a + b if two else a).clone().requires_grad_(True)
    gm, bt = gamma.clone().requires_grad_(True), torch.zeros(512, device=DEV, requires_grad=True)
    F.layer_norm(s, (512,), gm, bt, 1e-6).backward(dy1 + dy2 if two else dy1)
    dg, db = torch.zeros(512, device=DEV), torch.zeros(512, device=DEV)
    planes = torch.zeros(rows, 1024, dtype=torch.bfloat16, device=DEV)
    ds = ops.layernorm512_bwd(a, b, dy1, dy2, gamma, dg, db, 1e-6, dsbf=planes, write_lo=True)
    assert rel_l2(ds, s.grad) < 1e-5 and rel_l2(dg, gm.grad) < 1e-4 and rel_l2(db, bt.grad) < 1e-4
    assert float((planes[:, :512].float() + planes[:, 512:].float() - ds).abs().max()) <= float(ds.abs().max()) * 2.0 ** -14


def test_gelu_add_split_dropout_tokens(ops):
    g = torch.Generator(device=DEV).manual_seed(3)
    u = _bf(torch.randn(96, 2048, device=DEV, generator=g) * 2)
    assert_close_bf16(ops.gelu(u), F.gelu(u.float(), approximate="tanh"), "gelu_new")
    dact = _bf(torch.randn(96, 2048, device=DEV, generator=g))
    uf = u.float().requires_grad_(True)
    F.gelu(uf, approximate="tanh").backward(dact.float())
    assert_close_bf16(ops.gelu(u, dact), uf.grad, "gelu_new backward", rel=5e-3, max_rel=2e-2)
    # fp32 a + b -> fp32 sum + hi | lo planes
    a, b = torch.randn(64, 512, device=DEV, generator=g), torch.randn(64, 512, device=DEV, generator=g)
    planes = torch.zeros(64, 1024, dtype=torch.bfloat16, device=DEV)
    s = torch.empty_like(a)
    ops.add_split_bf16(a, b, planes, True, sum_out=s)
    assert torch.equal(s, a + b)
    assert float((planes[:, :512].float() + planes[:, 512:].float() - s).abs().max()) <= float(s.abs().max()) * 2.0 ** -14
    # dropout: deterministic in (seed, index), keeps ~1-p, scales by 1/(1-p); p = 0 is the identity
    x = torch.ones(1 << 20, device=DEV)
    y1, y2, y3 = ops.dropout(x, 0.1, 1234), ops.dropout(x, 0.1, 1234), ops.dropout(x, 0.1, 99)
    assert torch.equal(y1, y2) and not torch.equal(y1, y3)
    keep = float((y1 != 0).float().mean())
    assert abs(keep - 0.9) < 2e-3 and abs(float(y1.max()) - 1 / 0.9) < 1e-6
    assert torch.equal(ops.dropout(x, 0.0, 5), x)
    # token repack backward
    B, S, W, C = 2, 5, 8, 256
    gn = _bf(torch.randn(B, S, W, C, device=DEV, generator=g))
    tok = ops.nhwc_to_tokens(gn)
    assert torch.equal(tok, gn.permute(0, 1, 3, 2).reshape(B * S, C * W))


def test_tokens_ln_bwd(ops):
    B, S, W, C = 2, 6, 8, 256
    D = C * W
    g = torch.Generator(device=DEV).manual_seed(11)
    x = _bf(torch.randn(B, S, W, C, device=DEV, generator=g))
    te = torch.randn(8, D, device=DEV, generator=g)
    gamma = torch.randn(D, device=DEV, generator=g)
    dtok = torch.randn(B * S, D, device=DEV, generator=g)
    add = _bf(torch.randn(B, S, W, C, device=DEV, generator=g))
    xf = x.float().requires_grad_(True)
    gm, bt = gamma.clone().requires_grad_(True), torch.zeros(D, device=DEV, requires_grad=True)
    v = xf.permute(0, 1, 3, 2).reshape(B, S, D) + te[:S]
    F.layer_norm(v, (D,), gm, bt, 1e-6).reshape(B * S, D).backward(dtok)
    dg, db = torch.zeros(D, device=DEV), torch.zeros(D, device=DEV)
    dx = ops.tokens_ln_bwd(x, te, gamma, dtok, add, dg, db, 1e-6)
    assert_close_bf16(dx, xf.grad + add.float(), "tokens_ln_bwd")
    assert rel_l2(dg, gm.grad) < 1e-4 and rel_l2(db, bt.grad) < 1e-4


def test_temb_mlp_backward(ops):
    B = 5
    g = torch.Generator(device=DEV).manual_seed(2)
    te = torch.randn(1000, 128, device=DEV, generator=g)
    t = torch.randint(0, 1000, (B,), device=DEV, generator=g)
    ws = [torch.randn(512, 128, device=DEV, generator=g) / 11, torch.randn(512, 512, device=DEV, generator=g) / 22, torch.randn(4416, 512, device=DEV, generator=g) / 22]
    bs = [torch.randn(n, device=DEV, generator=g) * 0.1 for n in (512, 512, 4416)]
    wr = [w.clone().requires_grad_(True) for w in ws]
    br = [b.clone().requires_grad_(True) for b in bs]
    z1r = F.linear(te[t], wr[0], br[0]); z2r = F.linear(F.silu(z1r), wr[1], br[1]); outr = F.linear(F.silu(z2r), wr[2], br[2])
    gout = torch.randn(B, 4416, device=DEV, generator=g)
    outr.backward(gout)
    # forward with stored pre-activations (act bit 1 = SiLU on the input)
    z1 = ops.linear_small(te, ws[0], bs[0], act=0, idx=t)
    z2 = ops.linear_small(z1, ws[1], bs[1], act=2)
    out = ops.linear_small(z2, ws[2], bs[2], act=2)
    assert rel_l2(out, outr) < 1e-5
    dw = [torch.zeros_like(w) for w in ws]
    db = [torch.zeros_like(b) for b in bs]
    ops.linear_small_wgrad(gout, z2, dw[2], db[2], True)
    dz2 = ops.linear_small_dgrad(gout, ws[2], z2)
    ops.linear_small_wgrad(dz2, z1, dw[1], db[1], True)
    dz1 = ops.linear_small_dgrad(dz2, ws[1], z1)
    ops.linear_small_wgrad(dz1, te, dw[0], db[0], False, idx=t)
    for i in range(3):
        assert rel_l2(dw[i], wr[i].grad) < 1e-4 and rel_l2(db[i], br[i].grad) < 1e-4, i


def test_conv_in_out_wgrad_and_colsum(ops):
    B, T, Fq = 2, 12, 256
    g = torch.Generator(device=DEV).manual_seed(9)
    x = torch.randn(B, 2, T, Fq, device=DEV, generator=g)
    g1 = _bf(torch.randn(B, T, Fq, 32, device=DEV, generator=g))
    g2 = _bf(torch.randn(B, T, Fq, 32, device=DEV, generator=g))
    gs = _nchw(g1) + _nchw(g2)
    ref = torch.nn.grad.conv2d_weight(x, (32, 2, 3, 3), gs, padding=1)
    dw, db = torch.zeros(32, 2, 3, 3, device=DEV), torch.zeros(32, device=DEV)
    ops.conv_in_wgrad(x, g1, g2, dw, db)
    assert rel_l2(dw, ref) < 1e-4 and rel_l2(db, gs.sum(dim=(0, 2, 3))) < 1e-4
    h = _bf(torch.randn(B, T, Fq, 32, device=DEV, generator=g))
    skip = _bf(torch.randn(B, T, Fq, 32, device=DEV, generator=g))
    dout = torch.randn(B, 2, T, Fq, device=DEV, generator=g)
    ref = torch.nn.grad.conv2d_weight(_nchw(h) + _nchw(skip), (2, 32, 3, 3), dout, padding=1)
    dw, db = torch.zeros(2, 32, 3, 3, device=DEV), torch.zeros(2, device=DEV)
    ops.conv_out_wgrad(h, skip, dout, dw, db)
    assert rel_l2(dw, ref) < 1e-4 and rel_l2(db, dout.sum(dim=(0, 2, 3))) < 1e-4
    out = torch.zeros(32, device=DEV)
    ops.colsum(g1, out)
    assert rel_l2(out, g1.float().sum(dim=(0, 1, 2))) < 1e-4
    for C in (64, 96, 192, 256):                                   # flat 16-byte-chunk kernel: block size keeps a thread on the same 8 columns
        gc = _bf(torch.randn(3, 7, 33, C, device=DEV, generator=g))
        oc = torch.zeros(C, device=DEV)
        ops.colsum(gc, oc)
        assert rel_l2(oc, gc.float().sum(dim=(0, 1, 2))) < 1e-4, C
    f = torch.randn(1000, 512, device=DEV, generator=g)
    o2 = torch.zeros(512, device=DEV)
    ops.colsum(f, o2)
    assert rel_l2(o2, f.sum(0)) < 1e-5
