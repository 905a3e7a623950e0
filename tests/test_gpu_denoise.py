"""-m gpu: the denoise_2d post-filter through the reference-facing ``denoise_2d(x)`` and the C ABI
(``ddimb_denoise2d``), against (a) golden outputs of the reference's own function (External/UPU/signal/denoise.py:45-70),
(b) the CPU oracle on fresh seeded inputs, (c) size-independent properties at the sampling shape.

Bar: the two median subtractions are selection + one fp32 subtraction each -> BIT-EXACT (checked with iters = 0); the five
norm / mean iterations differ from torch only in the summation order of the means -> max-abs <= 2e-5 on unit-scale data.
"""
import os

import numpy as np
import pytest
import torch

from oracle import ddim_oracle as O

pytestmark = pytest.mark.gpu

TOL = 2e-5


def _gpu():
    from ddim_audio_b200 import ops
    from ddim_audio_b200.signal.denoise import denoise_2d

    return ops, denoise_2d


@pytest.mark.parametrize("name", ["denoise_b2_t48", "denoise_b1_t33_f128"])
def test_vs_reference_golden(golden_dir, name):
    _, denoise_2d = _gpu()
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    b, T, F_, seed = (int(v) for v in g["meta"])
    x = O.denoise_inputs(b, T, seed, f_size=F_)
    got = denoise_2d(x.cuda()).cpu()
    assert got.shape == x.shape and torch.isfinite(got).all()
    assert float((got - torch.from_numpy(g["out"])).abs().max()) <= TOL


# (B, T, F): T = 1 / 2 (degenerate medians), odd and even ranks, T beyond one histogram chunk (1024) and one iteration chunk
SHAPES = [(1, 1, 256), (1, 2, 256), (3, 7, 256), (2, 64, 128), (1, 40, 512), (2, 1030, 256), (1, 2309, 256), (5, 257, 256)]


@pytest.mark.parametrize("B,T,F_", SHAPES)
def test_medians_bit_exact(B, T, F_):
    """iters = 0 leaves r = (x - median_F(x)) - median_T(.): selection and two fp32 subtractions, no rounding freedom."""
    ops, _ = _gpu()
    x = O.denoise_inputs(B, T, 100 + T, f_size=F_)
    want = O.denoise_2d(x, iters=0)
    got = ops.denoise_2d(x.cuda(), iters=0).cpu()
    assert torch.equal(got, want)


@pytest.mark.parametrize("B,T,F_", SHAPES)
def test_vs_oracle(B, T, F_):
    _, denoise_2d = _gpu()
    x = O.denoise_inputs(B, T, 200 + T, f_size=F_)
    want = O.denoise_2d(x)
    got = denoise_2d(x.cuda()).cpu()
    assert float((got - want).abs().max()) <= TOL, float((got - want).abs().max())


def test_ties_negative_zero_and_leading_dims():
    """Heavily quantised data (every median rank falls inside a run of equal values) and extra leading dims [2, 3, 2, T, F]."""
    ops, denoise_2d = _gpu()
    g = torch.Generator().manual_seed(7)
    x = torch.randint(-3, 4, (2, 3, 2, 50, 256), generator=g).float() * 0.5
    want0 = O.denoise_2d(x, iters=0)
    got0 = ops.denoise_2d(x.cuda(), iters=0).cpu()
    assert torch.equal(got0 + 0.0, want0 + 0.0)                # + 0.0 folds -0.0 into +0.0 (torch's median picks either among equals)
    xr = x + 0.01 * torch.randn(x.shape, generator=g)
    assert float((denoise_2d(xr.cuda()).cpu() - O.denoise_2d(xr)).abs().max()) <= TOL


def test_input_untouched_reproducible_and_per_sample():
    """The reference returns a new tensor; the filter never mixes samples, so a batch equals its samples run alone (what lets
    ``sample_sequence`` shard it by batch)."""
    _, denoise_2d = _gpu()
    x = O.denoise_inputs(3, 96, 5).cuda()
    keep = x.clone()
    a = denoise_2d(x)
    b = denoise_2d(x)
    assert torch.equal(x, keep) and a.data_ptr() != x.data_ptr()
    assert float((a - b).abs().max()) <= 1e-6                  # fp64 atomics: order-dependent in the last bits only
    solo = torch.cat([denoise_2d(x[i:i + 1]) for i in range(3)])
    assert float((a - solo).abs().max()) <= 1e-6


def test_sampling_shape_properties():
    """x[4, 2, 8192, 256] (T of BASELINE configs[2]; the oracle's sorts take a few seconds here): bit-exact medians, the
    iterated result within tolerance, and two size-independent properties of the filter that need no oracle --
    equivariance under a permutation of the frames (medians are order-free, only the summation order of the means moves)
    and under scaling by a power of two (exact in fp32 up to the 1e-9 guard of denoise.py:64)."""
    ops, denoise_2d = _gpu()
    x = O.denoise_inputs(4, 8192, 9)
    xd = x.cuda()
    exact = torch.equal(ops.denoise_2d(xd, iters=0).cpu(), O.denoise_2d(x, iters=0))
    got = denoise_2d(xd)
    err = float((got.cpu() - O.denoise_2d(x)).abs().max())
    perm = torch.randperm(8192, generator=torch.Generator().manual_seed(1)).cuda()
    perm_exact = torch.equal(ops.denoise_2d(xd[:, :, perm].contiguous(), iters=0), ops.denoise_2d(xd, iters=0)[:, :, perm])
    perm_err = float((denoise_2d(xd[:, :, perm].contiguous()) - got[:, :, perm]).abs().max())
    scale_err = float((denoise_2d(xd * 4.0) - got * 4.0).abs().max()) / 4.0
    print(f"sampling shape: medians exact {exact}, max-abs vs oracle {err:.3e}, frame permutation exact(iters=0) {perm_exact} "
          f"err {perm_err:.3e}, x4 scaling err {scale_err:.3e}")
    assert exact and err <= TOL and perm_exact and perm_err <= 5e-5 and scale_err <= 5e-5, (exact, err, perm_exact, perm_err, scale_err)


def test_errors():
    ops, denoise_2d = _gpu()
    with pytest.raises(RuntimeError):
        denoise_2d(torch.zeros(1, 2, 8, 256))                   # CPU tensor: no CPU path
    with pytest.raises(RuntimeError):
        denoise_2d(torch.zeros(1, 3, 8, 256, device="cuda"))    # D != 2
    with pytest.raises(RuntimeError):
        denoise_2d(torch.zeros(1, 2, 8, 200, device="cuda"))    # W not 128 / 256 / 512
    with pytest.raises(RuntimeError):
        denoise_2d(np.zeros((1, 2, 8, 256), np.float32))        # numpy branch of the reference is not provided
    with pytest.raises(RuntimeError):
        denoise_2d(torch.zeros(1, 2, 8, 256, device="cuda", dtype=torch.float64))


def test_runner_applies_filter_to_x0(oracle_sd):
    """runners/diffusion.py:400-401: ``sample_sequence`` post-filters every returned x0 when config.sampling.denoise."""
    import argparse

    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200.runners.diffusion import Diffusion

    cfg = audio_config()
    cfg.sampling.num_samples, cfg.sampling.t_size = 2, 32
    args = argparse.Namespace(timesteps=3, sequence=2, skip_type="uniform", sample_type="generalized", eta=0.0, seed=5)
    model = Model(cfg)
    model.load_state_dict(oracle_sd, strict=True)
    model.cuda().eval()
    out = {}
    for flag in (False, True):
        cfg.sampling.denoise = flag
        r = Diffusion(args, cfg, device="cuda")
        out[flag] = r.sample_sequence(model, generator=torch.Generator().manual_seed(5))
    (xs_raw, x0_raw), (xs_f, x0_f) = out[False], out[True]
    assert len(x0_f) == len(x0_raw) == 2
    for a, b in zip(xs_raw, xs_f):
        assert torch.equal(a, b)
    for raw, f in zip(x0_raw, x0_f):
        assert not f.is_cuda
        assert float((f - O.denoise_2d(raw)).abs().max()) <= TOL * max(1.0, float(raw.abs().max()))   # random-init x0 is O(1e2)
