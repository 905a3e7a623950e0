"""Output codec behind the reference's ``pfft2wav`` (External/SST/utils/wav2img.py:82-127) as it is called after the sampling loop
(runners/diffusion.py:401-420): every predicted x0 ``[B, 2, T, F]`` is permuted ``(0, 3, 2, 1)`` to ``[B, F, T, 2]`` and each sample
becomes int32 PCM (ISO-226 loudness statistic, percentile-window normalisation, scipy's default inverse STFT, clip).

``pfft2wav_batch`` runs the whole batch on the device straight from the NCHW tensor (the permute is an index map inside the
kernels: csrc/pfft2wav.cu); ``pfft2wav`` keeps the reference's per-sample signature.  CUDA only: there is no CPU path.
The constant tables (ISO-226 weights, the irfft basis with the hann synthesis window folded in, the overlap-add normaliser) are
host-side float64 set-up, like the DFT operands of the FNet mixing."""
from __future__ import annotations

import math

import numpy as np
import torch

from .. import _C

# External/SST/utils/wav2img.py:11-27 (ISO 226 equal-loudness contour at 40 phon)
ISO226_BASE = 40
ISO226_FREQ = np.array([20., 25., 31.5, 40., 50., 63., 80., 100., 125., 160., 200., 250., 315., 400., 500., 630., 800., 1000., 1250., 1600.,
                        2000., 2500., 3150., 4000., 5000., 6300., 8000., 10000., 12500., 20000.])
ISO226_ESPL = np.array([99.85392334, 93.94441144, 88.16590253, 82.62867609, 77.78487094, 73.08254532, 68.47788682, 64.37114939, 60.58550325,
                        56.70224677, 53.40873978, 50.3992414, 47.5774866, 44.97662259, 43.05067937, 41.339195, 40.06176083, 40.01004637,
                        41.81945508, 42.50756876, 39.2296391, 36.50900986, 35.60891914, 36.64917709, 40.00774113, 45.82828132, 51.79680693,
                        54.28413025, 51.48590719, 99.85392334])


def get_sri(frequencies) -> np.ndarray:
    """wav2img.py:30-43."""
    frequencies = np.asarray(frequencies, dtype=np.float64)
    upp = np.minimum(np.searchsorted(ISO226_FREQ, frequencies, "left"), len(ISO226_FREQ) - 1)
    low = np.where(upp == 0, 1, upp - 1)
    sri = (ISO226_ESPL[upp] * (frequencies - ISO226_FREQ[low]) + ISO226_ESPL[low] * (ISO226_FREQ[upp] - frequencies)) / (ISO226_FREQ[upp] - ISO226_FREQ[low])
    return 10 ** ((ISO226_BASE - sri) / 10)


def _erfinv(y: float) -> float:
    """scipy.special.erfinv for a scalar in (-1, 1): Newton steps on math.erf from Winitzki's approximation (float64-exact here)."""
    a = 0.147
    ln = math.log(1.0 - y * y)
    t = 2.0 / (math.pi * a) + ln / 2.0
    x = math.copysign(math.sqrt(math.sqrt(t * t - ln / a) - t), y)
    for _ in range(4):
        x -= (math.erf(x) - y) / (2.0 / math.sqrt(math.pi) * math.exp(-x * x))
    return x


def norm_integral(p: float) -> float:
    """wav2img.py:46-49."""
    return -math.exp(-_erfinv(2 * p - 1) ** 2) / (math.pi * 2) ** 0.5


_tables = {}


def _codec_tables(F: int, samplerate: float, device):
    key = (int(F), float(samplerate), str(device))
    if key not in _tables:
        n = 2 * (F - 1)                                       # scipy.signal.istft's default segment length for a one-sided input
        half = n // 2
        NP = (n + 127) // 128 * 128
        freqs = np.arange(F, dtype=np.float64) / ((2 * F - 1) * (1.0 / samplerate))      # scipy.fft.rfftfreq(2F - 1, 1 / samplerate), :105
        sri = get_sri(freqs)
        win = 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)                        # get_window('hann', n)
        win32 = win.astype(np.float32)
        winsum = float(win32.sum())                                                      # float32 sum, as istft scales xsubs
        m = np.arange(n, dtype=np.float64)
        k = np.arange(F, dtype=np.float64)[:, None]
        wk = np.full((F, 1), 2.0)
        wk[0, 0] = wk[F - 1, 0] = 1.0
        ang = 2.0 * np.pi * k * m[None, :] / n
        scale = (win32.astype(np.float64) * winsum / n)[None, :]
        wt = np.zeros((2 * F, NP), dtype=np.float64)
        wt[:F, :n] = wk * np.cos(ang) * scale
        ws = -2.0 * np.sin(ang) * scale
        ws[0, :] = 0.0
        ws[F - 1, :] = 0.0                                                               # c2r ignores the imaginary parts of DC and Nyquist
        wt[F:, :n] = ws
        sq = win32 * win32                                                               # float32, as istft accumulates `norm`
        norm = (np.float32(0.0) + sq[half:]) + sq[:half]
        _tables[key] = (torch.from_numpy(sri).to(device), torch.from_numpy(wt.astype(np.float32)).to(device), torch.from_numpy(norm.astype(np.float32)).to(device),
                        n, half, NP)
    return _tables[key]


def pfft2wav_batch(x0: torch.Tensor, samplerate: int, normalize_percentile=(0.925, 0.975), Base_dB: float = 73, Max_dB: float = 86,
                   HPI: bool = True) -> torch.Tensor:
    """``x0`` fp32 CUDA ``[B, 2, T, F]`` (what ``generalized_steps`` / ``denoise_2d`` return) -> int32 CUDA ``[B, (F - 1) * (T - 1)]``:
    row b is ``pfft2wav(x0.permute(0, 3, 2, 1)[b], samplerate, np.int32, ...)`` of the reference."""
    if not x0.is_cuda:
        raise RuntimeError("ddim_audio_b200.pfft2wav_batch needs a CUDA tensor: there is no CPU path")
    if x0.dim() != 4 or x0.shape[1] != 2:
        raise RuntimeError("pfft2wav_batch: expected x0[B, 2, T, F]")
    x0 = x0.contiguous().float()
    B, _, T, F = x0.shape
    nbytes = int(_C.lib.ddimb_pfft2wav_workspace(B, T, F))
    if nbytes < 0:
        raise RuntimeError(f"pfft2wav_batch: unsupported shape {tuple(x0.shape)} (2 <= T <= 16384, F a multiple of 16)")
    p_low, p_upp = normalize_percentile
    target_norm = (norm_integral(p_upp) - norm_integral(p_low)) / (p_upp - p_low)
    sri, wt, norm, n, half, NP = _codec_tables(F, samplerate, x0.device)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x0.device)
    out = torch.empty((B, half * (T - 1)), dtype=torch.int32, device=x0.device)
    _C.call("ddimb_pfft2wav", _C.ptr(x0, torch.float32), _C.ptr(sri, torch.float64), _C.ptr(wt, torch.float32), _C.ptr(norm, torch.float32),
            _C.ptr(out), _C.ptr(ws), nbytes, B, T, F, 1 if HPI else 0, float(target_norm), float(Base_dB), float(Max_dB), float(p_low), float(p_upp))
    return out


def pfft2wav(spectrogram, samplerate: int, dtype=np.int32, normalize_percentile=(0.925, 0.975), Base_dB: float = 73, Max_dB: float = 86,
             HPI: bool = True) -> np.ndarray:
    """The reference's signature (wav2img.py:82-91): ``spectrogram`` ``[freq, time, 2]`` (numpy or torch) -> int32 numpy PCM."""
    if np.dtype(dtype) != np.int32:
        raise NotImplementedError("ddim_audio_b200.pfft2wav: int32 PCM only (the dtype the reference's sampler passes, runners/diffusion.py:413)")
    s = torch.as_tensor(spectrogram)
    assert s.shape[-1] == 2 and s.dim() == 3
    x0 = s.permute(2, 1, 0).unsqueeze(0).to(device="cuda", dtype=torch.float32).contiguous()      # [1, 2, T, F]
    return pfft2wav_batch(x0, samplerate, normalize_percentile, Base_dB, Max_dB, HPI)[0].cpu().numpy()
