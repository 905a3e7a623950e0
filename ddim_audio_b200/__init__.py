"""ddim_audio_b200 -- B200-native (sm_100a) implementation of the klae01/ddim-audio hot path.

Mirrors the reference's package layout for the path it replaces:
    ddim_audio_b200.models.diffusion.Model                      (models/diffusion.py)
    ddim_audio_b200.functions.denoising.generalized_steps       (functions/denoising.py)
    ddim_audio_b200.functions.losses.noise_estimation_loss      (functions/losses.py)
    ddim_audio_b200.runners.diffusion.Diffusion                 (runners/diffusion.py)
All arithmetic runs in libddim_b200.so (csrc/, C ABI in include/ddim_b200.h); importing the package without the
built library raises -- there is no CPU fallback.
"""
from . import _C  # noqa: F401  (fails loudly when the CUDA library is missing)

__all__ = ["_C"]
