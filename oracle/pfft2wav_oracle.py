"""CPU restatement (numpy) of the reference's output codec after the sampling loop -- TEST INFRASTRUCTURE ONLY (SURVEY.md section 8
row f4): ``runners/diffusion.py:401-420`` hands every predicted x0, permuted ``(0, 3, 2, 1)`` to ``[B, F, T, 2]``, sample by sample to
``External/SST/utils/wav2img.py:82-127 pfft2wav`` (ISO-226 loudness weighting, percentile normalisation, inverse STFT, int32 PCM).
The inverse STFT is ``scipy.signal.istft`` with its defaults (scipy 1.18.1 in this image: hann window, nperseg = nfft = 2 (F - 1) =
510 although the forward transform used 511 -- the reference's own asymmetry, kept), restated here as explicit irfft segments +
overlap-add so the CUDA kernels can be checked stage by stage.  Pinned against the reference's own function by
oracle/make_golden_pfft2wav.py (bit-exact in the build container)."""
from __future__ import annotations

import numpy as np
import scipy.fft
import scipy.special

# External/SST/utils/wav2img.py:11-27
ISO226_BASE = 40
ISO226_FREQ = np.array([20., 25., 31.5, 40., 50., 63., 80., 100., 125., 160., 200., 250., 315., 400., 500., 630., 800., 1000., 1250., 1600.,
                        2000., 2500., 3150., 4000., 5000., 6300., 8000., 10000., 12500., 20000.])
ISO226_ESPL = np.array([99.85392334, 93.94441144, 88.16590253, 82.62867609, 77.78487094, 73.08254532, 68.47788682, 64.37114939, 60.58550325,
                        56.70224677, 53.40873978, 50.3992414, 47.5774866, 44.97662259, 43.05067937, 41.339195, 40.06176083, 40.01004637,
                        41.81945508, 42.50756876, 39.2296391, 36.50900986, 35.60891914, 36.64917709, 40.00774113, 45.82828132, 51.79680693,
                        54.28413025, 51.48590719, 99.85392334])


def get_sri(frequencies: np.ndarray) -> np.ndarray:
    """wav2img.py:30-43 -- "sound recognition intensity": linear interpolation of the ISO-226 equal-loudness contour, as 10^((40 - L)/10)."""
    frequencies = np.asarray(frequencies, dtype=np.float64)
    upp = np.minimum(np.searchsorted(ISO226_FREQ, frequencies, "left"), len(ISO226_FREQ) - 1)
    low = np.where(upp == 0, 1, upp - 1)
    sri = (ISO226_ESPL[upp] * (frequencies - ISO226_FREQ[low]) + ISO226_ESPL[low] * (ISO226_FREQ[upp] - frequencies)) / (ISO226_FREQ[upp] - ISO226_FREQ[low])
    return 10 ** ((ISO226_BASE - sri) / 10)


def norm_integral(p: float) -> float:
    """wav2img.py:46-49."""
    return -np.exp(-scipy.special.erfinv(2 * p - 1) ** 2) / (np.pi * 2) ** 0.5


def rfft_frequencies(n_freq: int, samplerate: float) -> np.ndarray:
    """scipy.fft.rfftfreq(2 F - 1, 1 / samplerate) (wav2img.py:105): k * samplerate / (2 F - 1), k = 0 .. F - 1."""
    n = 2 * n_freq - 1
    return np.arange(n_freq, dtype=np.float64) / (n * (1.0 / samplerate))


def hann_periodic(n: int) -> np.ndarray:
    """scipy.signal.get_window('hann', n) (fftbins=True), float64."""
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


def loudness_rows(spec: np.ndarray, sri: np.ndarray, hpi: bool) -> np.ndarray:
    """s_HPI of wav2img.py:106-110 for complex64 spec [F, T]: per-frame sum over frequency of |spec| (x sri when HPI is off)."""
    a = np.abs(spec)
    return a.sum(axis=0) if hpi else (a * sri[..., None]).sum(axis=0)


def window_mean(s_hpi: np.ndarray, p_low: float, p_upp: float):
    """wav2img.py:112-113: mean of the order statistics int(T p_low) .. int(T p_upp) of s_HPI."""
    lo, hi = int(len(s_hpi) * p_low), int(len(s_hpi) * p_upp)
    return np.partition(s_hpi, [lo, hi])[lo:hi + 1].mean()


def istft_default(spec: np.ndarray) -> np.ndarray:
    """scipy.signal.istft(spec, fs) with every default, for complex64 spec [F, T] (scipy _spectral_py.py istft): irfft of n = 2 (F - 1)
    per frame, x win.sum(), windowed overlap-add with hop n / 2, n / 2 samples trimmed at both ends, divided by the summed win^2."""
    n_freq, nseg = spec.shape
    n = 2 * (n_freq - 1)
    nstep = n - n // 2
    win = hann_periodic(n)
    xsubs = scipy.fft.irfft(spec, axis=-2, n=n)[..., :n, :]             # scipy's pocketfft, as istft calls it: float32 for complex64 input
    win = win.astype(xsubs.dtype)
    xsubs *= win.sum()
    length = n + (nseg - 1) * nstep
    x = np.zeros(length, dtype=xsubs.dtype)
    norm = np.zeros(length, dtype=xsubs.dtype)
    for ii in range(nseg):
        x[ii * nstep:ii * nstep + n] += xsubs[:, ii] * win
        norm[ii * nstep:ii * nstep + n] += win ** 2
    x, norm = x[n // 2:-(n // 2)], norm[n // 2:-(n // 2)]
    x /= np.where(norm > 1e-10, norm, 1.0)
    return x


def pfft2wav(spectrogram: np.ndarray, samplerate: int, dtype=np.int32, normalize_percentile=(0.925, 0.975), Base_dB: float = 73,
             Max_dB: float = 86, HPI: bool = True) -> np.ndarray:
    """wav2img.py:82-127 for spectrogram [F, T, 2] (float32)."""
    assert spectrogram.shape[-1] == 2 and spectrogram.ndim == 3
    p_low, p_upp = normalize_percentile
    target_norm = (norm_integral(p_upp) - norm_integral(p_low)) / (p_upp - p_low)
    spec = spectrogram[..., 0] + 1j * spectrogram[..., 1]
    sri = get_sri(rfft_frequencies(spec.shape[0], samplerate))
    s_hpi = loudness_rows(spec, sri, HPI)
    if HPI:
        spec /= sri[..., None]                                              # in place: stays complex64, as in the reference
    current_norm = window_mean(s_hpi, p_low, p_upp)
    exp_amp = np.log10(current_norm / target_norm + 1e-9) * 10
    amp = min(Base_dB, Max_dB - exp_amp)
    r = istft_default(spec)
    r *= 10 ** (amp / 10)
    info = np.iinfo(dtype)
    return np.clip(r, info.min, info.max).astype(dtype)


def synthetic_spectrogram(batch: int, t_size: int, seed: int, f_size: int = 256, scale: float = 1.0) -> np.ndarray:
    """Seeded x0-like batch [B, 2, T, F] (float32): coloured noise with a per-frame loudness envelope, so the percentile window of
    the loudness statistic is not degenerate."""
    g = np.random.default_rng(seed)
    x = g.standard_normal((batch, 2, t_size, f_size)).astype(np.float32)
    env = (0.2 + np.abs(np.sin(np.arange(t_size)[None, None, :, None] * 0.37 + g.uniform(0, 6.28, (batch, 1, 1, 1))))).astype(np.float32)
    tilt = (1.0 / (1.0 + np.arange(f_size) / 40.0))[None, None, None, :].astype(np.float32)
    return (x * env * tilt * np.float32(scale)).astype(np.float32)
