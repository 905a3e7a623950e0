"""ctypes binding of libddim_b200.so (include/ddim_b200.h) -- the only door from Python to the sm_100a kernels.

There is no CPU path and no fallback: importing this module without the built library raises, and every
wrapper raises ``RuntimeError`` (with the library's own message) on a non-zero return code.  Tensors are passed
as raw device pointers plus sizes; launches go to ``torch.cuda.current_stream()``.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DDIMB_LIB_PATH") or os.path.join(_HERE, "lib", "libddim_b200.so")   # override: kernel-variant A/B runs (tools/)

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc); "
        "ddim_audio_b200 has no CPU fallback"
    )
lib = C.CDLL(LIB_PATH)

_vp, _i, _ll, _d, _f = C.c_void_p, C.c_int, C.c_longlong, C.c_double, C.c_float

# name -> argtypes (the trailing cudaStream_t is appended below); must mirror include/ddim_b200.h
SIGNATURES = {
    "ddimb_ddim_step": [_vp, _vp, _vp, _vp, _vp, _d, _d, _d, _d, _ll],
    "ddimb_qsample": [_vp, _vp, _vp, _vp, _vp, _i, _ll],
    "ddimb_sqerr_loss": [_vp, _vp, _vp, _i, _ll],
    "ddimb_sqerr_loss_bwd": [_vp, _vp, _vp, _vp, _i, _ll],
    "ddimb_repack": [_vp, _i, _ll],
    "ddimb_conv_wimg_batch": [_vp, _i],
    "ddimb_denoise2d": [_vp, _vp, _vp, _ll, _i, _i, _i, _i, _i],
    "ddimb_denoise2d_static": [_vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_pfft2wav": [_vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _i, _i, _d, _d, _d, _d, _d],
    "ddimb_ema_update": [_vp, _i, _ll, _d],
    "ddimb_grad_clip": [_vp, _i, _ll, _d, _vp, _vp],
    "ddimb_optim_step": [_vp, _i, _ll, _i, _d, _d, _d, _d, _d, _ll],
    "ddimb_conv_wimg": [_i, _vp, _vp, _i, _i, _i],
    "ddimb_conv_in": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv_out": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv3x3": [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv3x3_gn": [_vp, _vp, _vp, _vp, _f, _i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i],
    "ddimb_conv_down4x4": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv_up4x4": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_linear": [_vp, _ll, _vp, _ll, _vp, _i, _vp, _ll, _i, _ll, _i, _i, _i],
    "ddimb_fourier_mix": [_vp, _ll, _vp, _vp, _vp, _vp, _i, _i, _i, _i],
    "ddimb_gn_apply": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f],
    "ddimb_linear_small": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i],
    "ddimb_gather_rows": [_vp, _vp, _vp, _i, _i, _i],
    "ddimb_tokens_ln": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f],
    "ddimb_tokens_to_nhwc": [_vp, _vp, _vp, _vp, _i, _i, _i, _i],
    "ddimb_layernorm512": [_vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _f],
    "ddimb_split_bf16": [_vp, _vp, _ll, _i, _i, _ll],
    # ---- training
    "ddimb_conv3x3_wgrad": [_vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv_down4x4_wgrad": [_vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv_up4x4_wgrad": [_vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_linear_wgrad": [_vp, _ll, _vp, _ll, _vp, _i, _i, _i],
    "ddimb_gn_bwd_reduce": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f],
    "ddimb_gn_bwd_apply": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f],
    "ddimb_gn_bwd_param": [_vp, _vp, _vp, _vp, _i, _ll, _i, _f],
    "ddimb_colsum": [_vp, _i, _ll, _ll, _i, _vp],
    "ddimb_sums_scatter": [_vp, _vp, _ll, _vp, _i, _i],
    "ddimb_layernorm512_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _vp, _vp, _i, _f],
    "ddimb_gelu": [_vp, _vp, _vp, _ll],
    "ddimb_add_split_bf16": [_vp, _vp, _vp, _vp, _ll, _i, _i, _ll],
    "ddimb_dropout": [_vp, _vp, _ll, _f, C.c_ulonglong, _vp],
    "ddimb_tokens_ln_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f],
    "ddimb_nhwc_to_tokens": [_vp, _vp, _i, _i, _i, _i],
    "ddimb_linear_small_wgrad": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i],
    "ddimb_linear_small_dgrad": [_vp, _vp, _vp, _vp, _i, _i, _i],
    "ddimb_conv_in_wgrad": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i],
    "ddimb_conv_out_wgrad": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i],
}
EXPORTS = ["ddimb_version", "ddimb_last_error", "ddimb_launch_count", "ddimb_prof_enable", "ddimb_prof_report", "ddimb_conv3x3_wide_supported", "ddimb_conv3x3_gn_supported", "ddimb_denoise2d_workspace", "ddimb_denoise2d_static_max_bytes", "ddimb_conv_wimg_bytes", "ddimb_pfft2wav_workspace"] + list(SIGNATURES)

lib.ddimb_version.restype = _i
lib.ddimb_last_error.restype = C.c_char_p
lib.ddimb_launch_count.restype = _ll
lib.ddimb_conv3x3_wide_supported.argtypes = [_i, _i, _i]
lib.ddimb_conv3x3_wide_supported.restype = _i
lib.ddimb_conv3x3_gn_supported.argtypes = [_i, _i]
lib.ddimb_conv3x3_gn_supported.restype = _i
lib.ddimb_pfft2wav_workspace.argtypes = [_i, _i, _i]
lib.ddimb_pfft2wav_workspace.restype = _ll
lib.ddimb_conv_wimg_bytes.argtypes = [_i, _i, _i, _i]
lib.ddimb_conv_wimg_bytes.restype = _ll
lib.ddimb_denoise2d_workspace.argtypes = [_i, _i, _i, _i, _i]
lib.ddimb_denoise2d_workspace.restype = _ll
lib.ddimb_denoise2d_static_max_bytes.argtypes = []
lib.ddimb_denoise2d_static_max_bytes.restype = _ll
lib.ddimb_prof_enable.argtypes = [_i]
lib.ddimb_prof_enable.restype = _i
lib.ddimb_prof_report.argtypes = [C.c_char_p, _i]
lib.ddimb_prof_report.restype = _i
for _name, _args in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.argtypes = _args + [_vp]
    _fn.restype = _i


# Generation of "some parameter tensor was rewritten through a raw device pointer" (fused optimizer step, EMA copy-in):
# such writes bump neither ``Tensor._version`` nor ``data_ptr()``, so Model keys its packed bf16 operand copies on this.
param_epoch = [0]


def bump_param_epoch() -> None:
    param_epoch[0] += 1


# The hot host path: every library call needs the current device and its current stream.  torch.cuda.current_stream() /
# current_device() cost ~13 / ~1 us per call in Python (device-index normalisation, lazy-init checks): at ~1000 launches per train
# step that was 12 ms of a host-bound 58 ms step.  The two C entry points below return the same values in ~0.2 us.
_raw_stream = torch._C._cuda_getCurrentRawStream
_get_device = torch._C._cuda_getDevice


def _stream() -> int:
    return _raw_stream(_get_device())


def ptr(t, dtype=None):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("ddim_audio_b200: expected a CUDA tensor (there is no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"ddim_audio_b200: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise RuntimeError("ddim_audio_b200: expected a contiguous tensor")
    if t.device.index != _get_device():
        # launches go to torch.cuda.current_stream() of the CURRENT device; a tensor of another GPU would be launched on the wrong one
        raise RuntimeError(f"ddim_audio_b200: tensor on {t.device} but the current CUDA device is {torch.cuda.current_device()} "
                           "(wrap the call in torch.cuda.device(tensor.device))")
    return t.data_ptr()


def call(name: str, *args) -> None:
    rc = getattr(lib, name)(*args, _stream())
    if rc != 0:
        msg = lib.ddimb_last_error().decode(errors="replace")
        raise RuntimeError(f"{name} failed (code {rc}): {msg}")


def launch_count() -> int:
    """Kernels launched by the library so far."""
    return int(lib.ddimb_launch_count())


def prof_enable(on: bool) -> None:
    lib.ddimb_prof_enable(1 if on else 0)


def prof_report():
    """Per-kernel-class totals since the last report (synchronises the device): list of dicts."""
    import json

    buf = C.create_string_buffer(1 << 16)
    n = lib.ddimb_prof_report(buf, len(buf))
    if n < 0:
        raise RuntimeError("ddimb_prof_report failed: " + lib.ddimb_last_error().decode(errors="replace"))
    return json.loads(buf.value.decode())


def version() -> int:
    return int(lib.ddimb_version())
