"""Mint tests/golden/pfft2wav_*.npz from the UNMODIFIED reference codec (run in the build container only):

    python oracle/make_golden_pfft2wav.py

For seeded x0-like batches [B, 2, T, 256] (oracle/pfft2wav_oracle.py synthetic_spectrogram) the fixture holds what the reference's
sampler would write to the WAV file of every sample: ``pfft2wav(x0.permute(0, 3, 2, 1)[b], 48000, dtype=np.int32, HPI=...)``
(runners/diffusion.py:401-420, External/SST/utils/wav2img.py:82-127), plus the oracle's agreement with it (bit-exact here)."""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from oracle import pfft2wav_oracle as P  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
# (name, batch, T, seed, HPI, scale)
CASES = [("pfft2wav_b2_t64", 2, 64, 701, False, 1.0), ("pfft2wav_b3_t97_loud", 3, 97, 702, False, 40.0), ("pfft2wav_b1_t130_hpi", 1, 130, 703, True, 1e-9)]


def main():
    assert os.path.isdir(REF), "reference tree not found"
    cwd = os.getcwd()
    os.chdir(REF)
    sys.path.insert(0, os.path.join(REF, "External"))
    from SST.utils.wav2img import pfft2wav as ref_pfft2wav
    os.chdir(cwd)
    for name, b, T, seed, hpi, scale in CASES:
        x = P.synthetic_spectrogram(b, T, seed, scale=scale)
        wavs = []
        for j in range(b):
            img = np.ascontiguousarray(np.transpose(x[j], (2, 1, 0)))          # y.permute(0, 3, 2, 1)[j] -> [F, T, 2]
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                r = ref_pfft2wav(img.copy(), 48000, dtype=np.int32, HPI=hpi)
                o = P.pfft2wav(img.copy(), 48000, dtype=np.int32, HPI=hpi)
            assert np.array_equal(r, o), f"{name}[{j}]: oracle != reference"
            wavs.append(r)
        wavs = np.stack(wavs)
        sat = int((wavs == np.iinfo(np.int32).min).sum())
        print(f"{name}: wav {wavs.shape}, peak {int(np.abs(wavs.astype(np.int64)).max())}, saturated samples {sat}; oracle == reference")
        np.savez_compressed(os.path.join(OUT, name + ".npz"), wav=wavs, meta=np.array([b, T, seed, int(hpi)]), scale=np.array(scale), samplerate=np.array(48000))


if __name__ == "__main__":
    main()
