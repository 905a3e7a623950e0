"""``denoise_2d`` behind the reference's ``UPU.signal.denoise.denoise_2d`` (External/UPU/signal/denoise.py:45-70), the
post-filter ``Diffusion.sample_sequence`` applies to every returned x0 (runners/diffusion.py:400-401).

One library call (``ddimb_denoise2d``): exact lower medians by radix select, the five norm / mean iterations as read-only
passes over ``x`` (the per-row and per-column corrections are carried as offsets), one write.  CUDA fp32 tensors only --
the reference's numpy branch and ``generall_fully_denoise`` (unused by the runner) are not provided, and there is no CPU path.
"""
from __future__ import annotations

import torch

from .. import ops


@torch.no_grad()
def denoise_2d(x):
    """``x[*, D, H, W]`` (D = 2 channels, H = frames, W = frequency bins) -> filtered copy, same shape."""
    if not isinstance(x, torch.Tensor):
        raise RuntimeError("ddim_audio_b200.denoise_2d: torch CUDA tensors only (there is no CPU / numpy path)")
    if x.dtype != torch.float32:
        raise RuntimeError(f"ddim_audio_b200.denoise_2d: expected float32, got {x.dtype}")
    return ops.denoise_2d(x.contiguous())
