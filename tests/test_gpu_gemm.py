"""-m gpu: the tcgen05/TMEM/TMA implicit-GEMM kernel (csrc/umma_gemm.cuh) in each of its host programs --
Linear, Conv2d 3x3, Conv2d 4x4 s2, ConvTranspose2d 4x4 s2, FNet Fourier mixing -- against fp32 torch ops on the
same bf16-rounded operands."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from gpu_util import DEV, assert_close_bf16, assert_stats, rel_l2  # noqa: E402


@pytest.fixture(scope="module")
def ops():
    from ddim_audio_b200 import ops as o
    return o


def _bf(t):
    return t.to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K,act,f32out", [
    (128, 256, 512, 0, True), (96, 512, 2048, 0, True), (32, 2048, 512, 2, False), (1, 512, 2048, 0, True),
    (300, 2048, 512, 0, True), (257, 512, 512, 1, False), (1024, 2048, 512, 2, False), (8192, 512, 2048, 0, True),
])
def test_linear(ops, M, N, K, act, f32out):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    a = _bf(torch.randn(M, K, device=DEV, generator=g))
    w = _bf(torch.randn(N, K, device=DEV, generator=g) / math.sqrt(K))
    b = torch.randn(N, device=DEV, generator=g) * 0.2
    ref = F.linear(a.float(), w.float(), b)
    if act == 1:
        ref = F.silu(ref)
    elif act == 2:
        ref = F.gelu(ref, approximate="tanh")
    out = ops.linear(a, w, b, act=act, out_f32=f32out)
    if f32out:
        assert out.dtype == torch.float32
        assert rel_l2(out, ref) < 2e-4, rel_l2(out, ref)      # tanh.approx / bf16 products, fp32 accumulate
    else:
        assert_close_bf16(out, ref, f"linear {M}x{N}x{K}")


def test_linear_strided_operand_and_lo_plane(ops):
    g = torch.Generator(device=DEV).manual_seed(5)
    M, K, N = 64, 512, 2048
    buf = _bf(torch.randn(M, 2 * K, device=DEV, generator=g))            # hi | lo planes: only [:, :K] is the operand
    w = _bf(torch.randn(N, K, device=DEV, generator=g) / math.sqrt(K))
    out = ops.linear(buf, w, None, out_f32=True, M=M, K=K)
    assert rel_l2(out, F.linear(buf[:, :K].float(), w.float())) < 1e-4
    # bf16 output with the rounding residual written next to it
    o2 = torch.zeros(M, 2 * N, dtype=torch.bfloat16, device=DEV)
    ops.linear(buf, w, None, out=o2, out_lo_offset=N, M=M, K=K)
    full = F.linear(buf[:, :K].float(), w.float())
    assert float((o2[:, :N].float() + o2[:, N:].float() - full).abs().max()) <= float(full.abs().max()) * 2.0 ** -13


CONV_CASES = [  # (C, W, T, B)
    (32, 256, 8, 1), (32, 256, 5, 2), (64, 128, 16, 3), (96, 64, 8, 1), (128, 32, 12, 2), (192, 16, 10, 1), (256, 8, 4, 3),
    (256, 8, 1, 1), (128, 32, 1, 1),
    # C = 96 / 128 run the flattened-window kernel (csrc/conv_window.cuh): unit boundaries inside a row, ragged last unit, tiny images
    (96, 64, 33, 3), (96, 64, 1, 2), (96, 64, 130, 2), (96, 32, 7, 2), (96, 16, 5, 1),
    (128, 32, 33, 3), (128, 32, 130, 2), (128, 64, 9, 1), (128, 128, 5, 1),
]


@pytest.mark.parametrize("C,W,T,B", CONV_CASES)
@pytest.mark.parametrize("variant", ["temb_silu", "bias_silu_add"])
def test_conv3x3(ops, C, W, T, B, variant):
    g = torch.Generator(device=DEV).manual_seed(C + W + T)
    x = _bf(torch.randn(B, T, W, C, device=DEV, generator=g))
    w = _bf(torch.randn(C, C, 3, 3, device=DEV, generator=g) / math.sqrt(9 * C))
    wp = w.permute(0, 2, 3, 1).reshape(C, -1).contiguous()
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), None, padding=1)
    if variant == "temb_silu":
        temb = torch.randn(B, 4416, device=DEV, generator=g) * 0.5
        off = 64
        ref = F.silu(ref + temb[:, off:off + C, None, None])
        out = ops.conv3x3(x, wp, bias_b=temb[:, off:], bias_b_stride=4416, act=1, stats=st)
    else:
        bias = torch.randn(C, device=DEV, generator=g) * 0.3
        add = _bf(torch.randn(B, T, W, C, device=DEV, generator=g))
        ref = F.silu(ref + bias[None, :, None, None]) + add.float().permute(0, 3, 1, 2)
        out = ops.conv3x3(x, wp, bias=bias, act=1, add=add, stats=st)
    ref = ref.permute(0, 2, 3, 1)
    assert_close_bf16(out, ref, f"conv3x3 C={C} W={W} T={T} B={B} {variant}")
    assert_stats(st, ref, f"conv3x3 stats C={C}")


@pytest.mark.parametrize("Cin,Cout,W,T,B", [(32, 64, 256, 8, 1), (64, 96, 128, 4, 2), (96, 128, 64, 8, 3), (128, 192, 32, 6, 1),
                                            (192, 256, 16, 2, 2), (192, 256, 16, 64, 1),
                                            # 32 -> 64 at W = 256 runs csrc/conv_down_wide.cuh: one output row, ragged units, several samples
                                            (32, 64, 256, 2, 2), (32, 64, 256, 70, 3)])
def test_conv_down4x4(ops, Cin, Cout, W, T, B):
    g = torch.Generator(device=DEV).manual_seed(Cin + W)
    x = _bf(torch.randn(B, T, W, Cin, device=DEV, generator=g))
    w = _bf(torch.randn(Cout, Cin, 4, 4, device=DEV, generator=g) / math.sqrt(16 * Cin))
    bias = torch.randn(Cout, device=DEV, generator=g) * 0.3
    wp = w.permute(0, 2, 3, 1).reshape(Cout, -1).contiguous()
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    out = ops.conv_down4x4(x, wp, bias, st)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, stride=2, padding=1).permute(0, 2, 3, 1)
    assert_close_bf16(out, ref, f"down {Cin}->{Cout} W={W}")
    assert_stats(st, ref, "down stats")


@pytest.mark.parametrize("Cin,Cout,W,T,B", [(256, 192, 8, 1, 1), (256, 192, 8, 4, 2), (192, 128, 16, 6, 1), (128, 96, 32, 4, 3),
                                            (96, 64, 64, 8, 1), (64, 32, 128, 4, 2),
                                            # 64 -> 32 at W = 128 runs csrc/conv_up_wide.cuh: single row, ragged units, ring wrap over many rows
                                            (64, 32, 128, 1, 1), (64, 32, 128, 37, 3), (64, 32, 128, 300, 1)])
@pytest.mark.parametrize("with_skip", [False, True])
def test_conv_up4x4(ops, Cin, Cout, W, T, B, with_skip):
    g = torch.Generator(device=DEV).manual_seed(Cin + W + 1)
    x = _bf(torch.randn(B, T, W, Cin, device=DEV, generator=g))
    w = _bf(torch.randn(Cin, Cout, 4, 4, device=DEV, generator=g) / math.sqrt(4 * Cin))
    bias = torch.randn(Cout, device=DEV, generator=g) * 0.3
    wp = w.permute(2, 3, 1, 0).reshape(-1, Cin).contiguous()
    skip = _bf(torch.randn(B, 2 * T, 2 * W, Cout, device=DEV, generator=g)) if with_skip else None
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    out = ops.conv_up4x4(x, wp, bias, skip=skip, stats=st)
    ref = F.conv_transpose2d(x.float().permute(0, 3, 1, 2), w.float(), bias, stride=2, padding=1).permute(0, 2, 3, 1)
    if with_skip:
        ref = ref + skip.float()
    assert_close_bf16(out, ref, f"up {Cin}->{Cout} W={W}")
    assert_stats(st, ref, "up stats")


@pytest.mark.parametrize("B,S", [(1, 1), (2, 2), (1, 3), (3, 32), (2, 256)])
def test_fourier_mix(ops, B, S):
    from ddim_audio_b200.models.diffusion import _fourier_operands
    H = 512
    g = torch.Generator(device=DEV).manual_seed(S)
    x = torch.randn(B * S, H, device=DEV, generator=g) * 2
    xbf = torch.zeros(B * S, 2 * H, dtype=torch.bfloat16, device=DEV)
    ops.split_bf16(x, xbf, True)
    wh, a2, Sp = _fourier_operands(S, H, DEV)
    ut = torch.zeros(B, 2 * H, 2 * Sp, dtype=torch.bfloat16, device=DEV)
    y = ops.fourier_mix(xbf, 2 * H, wh, a2, ut, B, S, H, 3)
    ref = torch.fft.fftn(x.double().reshape(B, S, H), dim=(1, 2)).real.reshape(B * S, H)
    r = rel_l2(y, ref)
    assert r < 2e-4, f"fourier split-bf16 rel-l2 {r:.3e}"


# ---------------------------------------------------------------------------------------------- wide 3x3 conv (+ fused GroupNorm)
WIDE_CASES = [(32, 256, 8, 1), (32, 256, 70, 2), (32, 128, 33, 3), (64, 128, 16, 2), (64, 256, 9, 1), (64, 128, 130, 2), (32, 256, 1, 1),
              # long strips (many ring revolutions per CTA): the shape at which a ring / transform-warp phase aliasing once deadlocked
              (64, 128, 4096, 6), (32, 256, 2048, 5)]


# the window kernel (C = 96 / 128): ragged last unit, T = 1, several samples, long images (many units per CTA), other widths
WINDOW_GN_CASES = [(96, 64, 8, 1), (96, 64, 33, 3), (96, 64, 1, 2), (128, 32, 12, 2), (128, 32, 1, 1), (128, 32, 257, 3), (96, 64, 2048, 4),
                   (128, 32, 1024, 6), (96, 32, 7, 2), (128, 64, 9, 1)]


@pytest.mark.parametrize("C,W,T,B", WIDE_CASES + WINDOW_GN_CASES)
@pytest.mark.parametrize("gn_mode", [1, 2])
def test_conv3x3_gn_fused(ops, C, W, T, B, gn_mode):
    """models/diffusion.py:47-52: conv(silu(GN(x))) / conv(GN(x)) in one kernel vs torch group_norm + conv2d in fp32."""
    from gpu_util import group_stats
    assert ops.conv3x3_gn_supported(C, W)
    g = torch.Generator(device=DEV).manual_seed(C + W + T + gn_mode)
    x = _bf(torch.randn(B, T, W, C, device=DEV, generator=g) * 1.5 + 0.3)
    w = _bf(torch.randn(C, C, 3, 3, device=DEV, generator=g) / math.sqrt(9 * C))
    wp = w.permute(0, 2, 3, 1).reshape(C, -1).contiguous()
    gamma = 1.0 + 0.2 * torch.randn(C, device=DEV, generator=g)
    beta = 0.2 * torch.randn(C, device=DEV, generator=g)
    bias = torch.randn(C, device=DEV, generator=g) * 0.3
    sti = group_stats(x).contiguous()
    st = torch.zeros(B, 8, 2, dtype=torch.float64, device=DEV)
    out = ops.conv3x3_gn(x, sti, gamma, beta, 1e-6, gn_mode, wp, bias=bias, act=1, stats=st)
    a = F.group_norm(x.float().permute(0, 3, 1, 2), 8, gamma, beta, eps=1e-6)
    if gn_mode == 2:
        a = F.silu(a)
    a = a.to(torch.bfloat16).float()          # the kernel rounds the activated operand to bf16 before the tensor core
    ref = F.silu(F.conv2d(a, w.float(), bias, padding=1)).permute(0, 2, 3, 1)
    assert_close_bf16(out, ref, f"conv3x3_gn C={C} W={W} T={T} B={B} mode={gn_mode}", rel=6e-3, max_rel=3e-2)
    assert_stats(st, ref, f"conv3x3_gn stats C={C}")
