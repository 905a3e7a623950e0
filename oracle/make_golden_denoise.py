"""Mint tests/golden/denoise_*.npz from the UNMODIFIED reference post-filter (build container only).

    python oracle/make_golden_denoise.py

``denoise_2d`` (External/UPU/signal/denoise.py:45-70, called at runners/diffusion.py:400-401) is imported read-only from
/root/reference and run on CPU on the seeded inputs of ``ddim_oracle.denoise_inputs``; the inputs are regenerated from the
seed at test time, only the outputs are stored.  Prints the oracle's agreement with the reference.
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ddim_oracle as O  # noqa: E402

REF = os.environ.get("DDIM_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
# (name, batch, T, F, seed): F = 256 is audio.yml; odd T exercises the (n-1)//2 median rank
CASES = [("denoise_b2_t48", 2, 48, 256, 51), ("denoise_b1_t33_f128", 1, 33, 128, 52)]


def main():
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location("ref_denoise", os.path.join(REF, "External", "UPU", "signal", "denoise.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for name, b, T, F_, seed in CASES:
        x = O.denoise_inputs(b, T, seed, f_size=F_)
        ref = mod.denoise_2d(x.clone())
        got = O.denoise_2d(x)
        print(f"{name}: oracle vs reference max-abs {float((got - ref).abs().max()):.3e} rel-l2 {O.rel_l2(got, ref):.3e}")
        np.savez_compressed(os.path.join(OUT, name + ".npz"), out=ref.numpy(), meta=np.array([b, T, F_, seed]))


if __name__ == "__main__":
    main()
