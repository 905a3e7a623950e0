"""Tensor-level wrappers over the C ABI (one per entry point of include/ddim_b200.h).

Layout conventions: network activations are NHWC bf16 ``[B, T, W, C]``; GroupNorm statistics are
``float64 [B, 8, 2]`` (sum, sum of squares) accumulated by the producing kernel into a zeroed buffer.
Allocation uses torch's caching allocator (plumbing); all arithmetic happens in the library.
"""
from __future__ import annotations

import torch

from . import _C
from ._C import call, ptr

bf16, f32, f64, i64 = torch.bfloat16, torch.float32, torch.float64, torch.int64


# ---------------------------------------------------------------------------------------- sampler / loss
def ddim_step(xt, et, z, at, at_next, c1, c2, x0_out=None, xs_out=None):
    """functions/denoising.py:27,41-43 in place on ``xt`` (fp32)."""
    n = xt.numel()
    call("ddimb_ddim_step", ptr(xt, f32), ptr(et, f32), ptr(z, f32) if z is not None else None,
         ptr(x0_out, f32) if x0_out is not None else None, ptr(xs_out, f32) if xs_out is not None else None,
         float(at), float(at_next), float(c1), float(c2), n)
    return xt


def qsample(x0, e, alphas, t):
    """functions/losses.py:12-13."""
    out = torch.empty_like(x0)
    B = x0.shape[0]
    call("ddimb_qsample", ptr(x0, f32), ptr(e, f32), ptr(alphas, f32), ptr(t, i64), ptr(out, f32), B, x0.numel() // B)
    return out


def sqerr_loss(e, output):
    """functions/losses.py:15-18 -> float64 [B] per-sample sums."""
    B = e.shape[0]
    per = torch.empty(B, dtype=f64, device=e.device)
    call("ddimb_sqerr_loss", ptr(e, f32), ptr(output, f32), ptr(per, f64), B, e.numel() // B)
    return per


def sqerr_loss_bwd(e, output, g):
    B = e.shape[0]
    go = torch.empty_like(output)
    call("ddimb_sqerr_loss_bwd", ptr(e, f32), ptr(output, f32), ptr(g, f32), ptr(go, f32), B, e.numel() // B)
    return go


def denoise_2d(x, iters=5):
    """External/UPU/signal/denoise.py:45-70 on fp32 ``x[*, 2, T, F]`` -> new tensor."""
    if x.dim() < 3:
        raise RuntimeError("denoise_2d: expected x[*, D, H, W]")
    C, T, F = x.shape[-3:]
    B = x.numel() // max(1, C * T * F)
    if B == 0:
        return torch.empty_like(x)
    nbytes = int(_C.lib.ddimb_denoise2d_workspace(B, C, T, F, int(iters)))
    if nbytes < 0:
        raise RuntimeError(f"denoise_2d: unsupported shape {tuple(x.shape)} (D must be 2, W one of 128/256/512, iters <= 8); there is no CPU path")
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    out = torch.empty_like(x)
    call("ddimb_denoise2d", ptr(x, f32), ptr(out, f32), ptr(ws), nbytes, B, C, T, F, int(iters))
    return out


# ---------------------------------------------------------------------------------------- convolutions
KIND_CONV3X3, KIND_DOWN, KIND_UP, KIND_CONV_IN, KIND_CONV_OUT = 0, 1, 2, 3, 4


def conv_wimg(kind, w, Cin, Cout, W):
    """Kernel-ready weight image of one convolution (include/ddim_b200.h ``ddimb_conv_wimg``): a uint8 device tensor, or None
    when the kernel for this shape reads ``w`` directly.  ``Model._pack`` builds these once per weight version; the op wrappers
    below build one on the fly when the caller passes none (tests, one-off calls)."""
    nbytes = int(_C.lib.ddimb_conv_wimg_bytes(int(kind), int(Cin), int(Cout), int(W)))
    if nbytes < 0:
        raise RuntimeError(f"ddim_audio_b200: no convolution kernel for kind {kind}, Cin {Cin}, Cout {Cout}, W {W}")
    if nbytes == 0:
        return None
    img = torch.empty(nbytes, dtype=torch.uint8, device=w.device)
    call("ddimb_conv_wimg", int(kind), ptr(w), ptr(img), int(Cin), int(Cout), int(W))
    return img


def conv_in(x, w, bias, stats=None, wimg=None):
    B, Cin, T, F = x.shape
    Cout = w.shape[0]
    if wimg is None:
        wimg = conv_wimg(KIND_CONV_IN, w, Cin, Cout, F)
    out = torch.empty((B, T, F, Cout), dtype=bf16, device=x.device)
    call("ddimb_conv_in", ptr(x, f32), ptr(w, f32), ptr(wimg), ptr(bias, f32), ptr(out), ptr(stats, f64), B, Cin, T, F, Cout)
    return out


def conv_out(x, skip, w, bias, wimg=None):
    B, T, F, Cin = x.shape
    Cout = w.shape[0]
    if wimg is None:
        wimg = conv_wimg(KIND_CONV_OUT, w, Cin, Cout, F)
    out = torch.empty((B, Cout, T, F), dtype=f32, device=x.device)
    call("ddimb_conv_out", ptr(x, bf16), ptr(skip, bf16), ptr(w, f32), ptr(wimg), ptr(bias, f32), ptr(out), B, T, F, Cin, Cout)
    return out


def conv3x3(x, w_packed, bias=None, bias_b=None, bias_b_stride=0, act=0, add=None, stats=None, out=None, wimg=None):
    """x [B,T,W,Cin] bf16; w_packed bf16 [Cout, 9*Cin]; bias_b is a (possibly offset) fp32 view whose row b starts at
    ``bias_b.data_ptr() + 4*b*bias_b_stride``."""
    B, T, W, Cin = x.shape
    Cout = w_packed.shape[0]
    if wimg is None:
        wimg = conv_wimg(KIND_CONV3X3, w_packed, Cin, Cout, W)
    if out is None:
        out = torch.empty((B, T, W, Cout), dtype=bf16, device=x.device)
    call("ddimb_conv3x3", ptr(x, bf16), ptr(w_packed, bf16), ptr(wimg), ptr(bias, f32), bias_b.data_ptr() if bias_b is not None else None,
         int(bias_b_stride), int(act), ptr(add, bf16), ptr(out, bf16), ptr(stats, f64), B, T, W, Cin, Cout)
    return out


def conv3x3_gn_supported(C, W) -> bool:
    return bool(_C.lib.ddimb_conv3x3_gn_supported(int(C), int(W)))


def conv3x3_gn(x, gn_stats, gamma, beta, eps, gn_mode, w_packed, bias=None, bias_b=None, bias_b_stride=0, act=0, add=None, stats=None, out=None,
               wimg=None):
    """GroupNorm(+SiLU) of ``x`` from ``gn_stats`` fused into the 3x3 convolution (models/diffusion.py:47-52)."""
    B, T, W, C = x.shape
    if wimg is None:
        wimg = conv_wimg(KIND_CONV3X3, w_packed, C, C, W)
    if out is None:
        out = torch.empty((B, T, W, C), dtype=bf16, device=x.device)
    call("ddimb_conv3x3_gn", ptr(x, bf16), ptr(gn_stats, f64), ptr(gamma, f32), ptr(beta, f32), float(eps), int(gn_mode), ptr(w_packed, bf16), ptr(wimg),
         ptr(bias, f32), bias_b.data_ptr() if bias_b is not None else None, int(bias_b_stride), int(act), ptr(add, bf16), ptr(out, bf16),
         ptr(stats, f64), B, T, W, C)
    return out


def conv_down4x4(x, w_packed, bias, stats=None, wimg=None):
    B, T, W, Cin = x.shape
    Cout = w_packed.shape[0]
    if wimg is None:
        wimg = conv_wimg(KIND_DOWN, w_packed, Cin, Cout, W)
    out = torch.empty((B, T // 2, W // 2, Cout), dtype=bf16, device=x.device)
    call("ddimb_conv_down4x4", ptr(x, bf16), ptr(w_packed, bf16), ptr(wimg), ptr(bias, f32), ptr(out), ptr(stats, f64), B, T, W, Cin, Cout)
    return out


def conv_up4x4(x, w_packed, bias, skip=None, stats=None, wimg=None):
    B, T, W, Cin = x.shape
    Cout = w_packed.shape[0] // 16
    if wimg is None:
        wimg = conv_wimg(KIND_UP, w_packed, Cin, Cout, W)
    out = torch.empty((B, 2 * T, 2 * W, Cout), dtype=bf16, device=x.device)
    call("ddimb_conv_up4x4", ptr(x, bf16), ptr(w_packed, bf16), ptr(wimg), ptr(bias, f32), ptr(skip, bf16), ptr(out), ptr(stats, f64),
         B, T, W, Cin, Cout)
    return out


# ---------------------------------------------------------------------------------------- GEMMs
def linear(a, w, bias=None, act=0, out_f32=False, out=None, out_lo_offset=0, M=None, K=None):
    """out[M,N] = act(a[M,K] w[N,K]^T + bias); a, w bf16 2-D (row stride = leading dimension)."""
    M = a.shape[0] if M is None else M
    K = w.shape[1] if K is None else K
    N = w.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=f32 if out_f32 else bf16, device=a.device)
    assert a.stride(-1) == 1 and w.stride(-1) == 1 and out.stride(-1) == 1
    call("ddimb_linear", a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), ptr(bias, f32), int(act), out.data_ptr(),
         out.stride(0), 1 if out.dtype == f32 else 0, int(out_lo_offset), M, N, K)
    return out


def fourier_mix(xbf, ldx, wh, a2, ut, B, S, H, split):
    """hf modeling_fnet.py:142-177 -> fp32 [B*S, H]."""
    y = torch.empty((B * S, H), dtype=f32, device=xbf.device)
    call("ddimb_fourier_mix", xbf.data_ptr(), int(ldx), ptr(wh, bf16), ptr(a2, bf16), ptr(ut, bf16), ptr(y), B, S, H, int(split))
    return y


# ---------------------------------------------------------------------------------------- norms / repacks
def gn_apply(x, stats_in, gamma, beta, mode, eps, res=None, stats_out=None, out=None, res2=None):
    B, C = x.shape[0], x.shape[-1]
    positions = x.numel() // (B * C)
    if out is None:
        out = torch.empty_like(x)
    call("ddimb_gn_apply", ptr(x, bf16), ptr(res, bf16), ptr(res2, bf16), ptr(out, bf16), ptr(stats_in, f64), ptr(stats_out, f64), ptr(gamma, f32),
         ptr(beta, f32), int(mode), B, positions, C, float(eps))
    return out


def linear_small(inp, w, bias, act=0, idx=None):
    rows = idx.numel() if idx is not None else inp.shape[0]
    N, K = w.shape
    out = torch.empty((rows, N), dtype=f32, device=w.device)
    call("ddimb_linear_small", ptr(inp, f32), ptr(idx, i64), ptr(w, f32), ptr(bias, f32), ptr(out), rows, K, N, int(act))
    return out


def gather_rows(table, idx):
    """out[r] = table[idx[r]] (fp32)."""
    rows, width = idx.numel(), table.shape[1]
    out = torch.empty((rows, width), dtype=f32, device=table.device)
    call("ddimb_gather_rows", ptr(table, f32), ptr(idx, i64), ptr(out), rows, width, table.shape[0])
    return out


def tokens_ln(x, te, gamma, beta, eps):
    B, S, W, Cc = x.shape
    out = torch.empty((B * S, Cc * W), dtype=bf16, device=x.device)
    call("ddimb_tokens_ln", ptr(x, bf16), ptr(te, f32), ptr(gamma, f32), ptr(beta, f32), ptr(out), B, S, W, Cc, float(eps))
    return out


def tokens_to_nhwc(tok, skip, stats=None):
    B, S, W, Cc = skip.shape
    out = torch.empty_like(skip)
    call("ddimb_tokens_to_nhwc", ptr(tok, f32), ptr(skip, bf16), ptr(out), ptr(stats, f64), B, S, W, Cc)
    return out


def layernorm512(a, b, gamma, beta, eps, ybf, write_lo):
    rows = a.shape[0]
    y = torch.empty_like(a)
    call("ddimb_layernorm512", ptr(a, f32), ptr(b, f32), ptr(gamma, f32), ptr(beta, f32), ptr(y), ptr(ybf, bf16), ybf.stride(0),
         1 if write_lo else 0, rows, float(eps))
    return y


def split_bf16(a, ybf, write_lo):
    rows, H = a.shape
    call("ddimb_split_bf16", ptr(a, f32), ptr(ybf, bf16), ybf.stride(0), H, 1 if write_lo else 0, rows)
    return ybf


# ---------------------------------------------------------------------------------------- training: weight gradients
def conv3x3_wgrad(x, dy, dw):
    """dw fp32 [Cout, Cin, 3, 3] += autograd of Conv2d 3x3 (x, dy bf16 NHWC)."""
    B, T, W, Cin = x.shape
    call("ddimb_conv3x3_wgrad", ptr(x, bf16), ptr(dy, bf16), ptr(dw, f32), B, T, W, Cin, dy.shape[-1])


def conv_down4x4_wgrad(x, dy, dw):
    B, T, W, Cin = x.shape
    call("ddimb_conv_down4x4_wgrad", ptr(x, bf16), ptr(dy, bf16), ptr(dw, f32), B, T, W, Cin, dy.shape[-1])


def conv_up4x4_wgrad(x, dy, dw):
    B, T, W, Cin = x.shape
    call("ddimb_conv_up4x4_wgrad", ptr(x, bf16), ptr(dy, bf16), ptr(dw, f32), B, T, W, Cin, dy.shape[-1])


def linear_wgrad(a, g, dw, rows=None, K=None):
    """dw fp32 [N, K] += g[rows, N]^T a[rows, K] (bf16 operands, row stride = leading dimension)."""
    rows = a.shape[0] if rows is None else rows
    N, Kw = dw.shape
    K = Kw if K is None else K
    assert a.stride(-1) == 1 and g.stride(-1) == 1
    call("ddimb_linear_wgrad", a.data_ptr(), a.stride(0), g.data_ptr(), g.stride(0), ptr(dw, f32), rows, N, K)


# ---------------------------------------------------------------------------------------- training: GroupNorm backward
def gn_bwd_reduce(g, z, stats, gamma, beta, gtype, eps):
    B, C = z.shape[0], z.shape[-1]
    R = torch.zeros((B, C, 2), dtype=f64, device=z.device)
    call("ddimb_gn_bwd_reduce", ptr(g, bf16), ptr(z, bf16), ptr(stats, f64), ptr(gamma, f32), ptr(beta, f32), ptr(R, f64), int(gtype), B,
         z.numel() // (B * C), C, float(eps))
    return R


def gn_bwd_apply(g, z, stats, R, gamma, beta, gtype, eps, add=None, sums=False, out=None):
    B, C = z.shape[0], z.shape[-1]
    if out is None:
        out = torch.empty_like(z)
    s = torch.zeros((B, C), dtype=f64, device=z.device) if sums else None
    call("ddimb_gn_bwd_apply", ptr(g, bf16), ptr(z, bf16), ptr(add, bf16), ptr(out, bf16), ptr(stats, f64), ptr(R, f64), ptr(gamma, f32),
         ptr(beta, f32), ptr(s, f64), int(gtype), B, z.numel() // (B * C), C, float(eps))
    return out, s


def gn_bwd_param(R, stats, dgamma, dbeta, positions, eps):
    B, C = R.shape[0], R.shape[1]
    call("ddimb_gn_bwd_param", ptr(R, f64), ptr(stats, f64), ptr(dgamma, f32), ptr(dbeta, f32), B, int(positions), C, float(eps))


def colsum(x, out):
    """out fp32 [C] += column sums of x viewed as [rows, C] (bf16 or fp32; last dim contiguous)."""
    x2 = x.reshape(-1, x.shape[-1])
    call("ddimb_colsum", x2.data_ptr(), 1 if x.dtype == bf16 else 0, x2.stride(0), x2.shape[0], x2.shape[1], ptr(out, f32))


def sums_scatter(sums, out_b=None, ldb=0, out_c=None):
    B, C = sums.shape
    call("ddimb_sums_scatter", ptr(sums, f64), out_b.data_ptr() if out_b is not None else None, int(ldb), ptr(out_c, f32), B, C)


# ---------------------------------------------------------------------------------------- training: FNet / tokens / temb / boundary
def layernorm512_bwd(a, b, dy1, dy2, gamma, dgamma, dbeta, eps, dsbf=None, write_lo=False):
    rows = a.shape[0]
    ds = torch.empty_like(a)
    call("ddimb_layernorm512_bwd", ptr(a, f32), ptr(b, f32), ptr(dy1, f32), ptr(dy2, f32), ptr(gamma, f32), ptr(ds), ptr(dsbf, bf16),
         dsbf.stride(0) if dsbf is not None else 0, 1 if write_lo else 0, ptr(dgamma, f32), ptr(dbeta, f32), rows, float(eps))
    return ds


def gelu(u, dact=None):
    out = torch.empty_like(u)
    call("ddimb_gelu", ptr(u, bf16), ptr(dact, bf16), ptr(out, bf16), u.numel())
    return out


def add_split_bf16(a, b, ybf, write_lo, sum_out=None):
    rows, H = a.shape
    call("ddimb_add_split_bf16", ptr(a, f32), ptr(b, f32), ptr(sum_out, f32), ptr(ybf, bf16), ybf.stride(0), H, 1 if write_lo else 0, rows)
    return ybf


def dropout(x, p, seed, out=None, seed_dev=None):
    """``seed_dev``: optional int64 device tensor added to ``seed`` inside the kernel (CUDA-graph replays: models/training.py)."""
    if out is None:
        out = torch.empty_like(x)
    call("ddimb_dropout", ptr(x, f32), ptr(out, f32), x.numel(), float(p), int(seed) & 0xFFFFFFFFFFFFFFFF, ptr(seed_dev))
    return out


def tokens_ln_bwd(x, te, gamma, dtok, add, dgamma, dbeta, eps):
    B, S, W, Cc = x.shape
    dx = torch.empty_like(x)
    call("ddimb_tokens_ln_bwd", ptr(x, bf16), ptr(te, f32), ptr(gamma, f32), ptr(dtok, f32), ptr(add, bf16), ptr(dx, bf16), ptr(dgamma, f32),
         ptr(dbeta, f32), B, S, W, Cc, float(eps))
    return dx


def nhwc_to_tokens(g):
    B, S, W, Cc = g.shape
    tok = torch.empty((B * S, Cc * W), dtype=bf16, device=g.device)
    call("ddimb_nhwc_to_tokens", ptr(g, bf16), ptr(tok, bf16), B, S, W, Cc)
    return tok


def linear_small_wgrad(g, inp, dw, db, in_act, idx=None):
    rows, N = g.shape
    K = dw.shape[1]
    call("ddimb_linear_small_wgrad", ptr(g, f32), ptr(inp, f32), ptr(idx, i64), ptr(dw, f32), ptr(db, f32), rows, K, N, 1 if in_act else 0)


def linear_small_dgrad(g, w, z):
    rows, N = g.shape
    K = w.shape[1]
    out = torch.empty((rows, K), dtype=f32, device=g.device)
    call("ddimb_linear_small_dgrad", ptr(g, f32), ptr(w, f32), ptr(z, f32), ptr(out), rows, K, N)
    return out


def conv_in_wgrad(x, g1, g2, dw, db):
    B, Cin, T, F = x.shape
    call("ddimb_conv_in_wgrad", ptr(x, f32), ptr(g1, bf16), ptr(g2, bf16), ptr(dw, f32), ptr(db, f32), B, Cin, T, F, dw.shape[0])


def conv_out_wgrad(h, skip, dout, dw, db):
    B, T, F, Cin = h.shape
    call("ddimb_conv_out_wgrad", ptr(h, bf16), ptr(skip, bf16), ptr(dout, f32), ptr(dw, f32), ptr(db, f32), B, T, F, Cin, dw.shape[0])
