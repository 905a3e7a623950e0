"""-m gpu: the CUDA denoiser / sampler / loss through the reference-facing Python API, against (a) golden outputs
minted from the UNMODIFIED reference (tests/golden, made by oracle/make_golden.py) and (b) the CPU oracle on fresh
seeded inputs.  Tolerance (bf16 activation storage, fp32 accumulate): per-step epsilon
rel-L2 <= 2e-2 and max-abs <= 0.1 on outputs of std 0.88, i.e. max-abs <= 0.114 * std(reference output) -- the
synthetic weights here give epsilon of std ~4, so the absolute bound is scaled by the output's own std."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import DEV, rel_l2  # noqa: E402
from oracle import ddim_oracle as O  # noqa: E402

EPS_REL, EPS_ABS_PER_STD = 2e-2, 0.114


@pytest.fixture(scope="module")
def model(oracle_sd):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    m = Model(audio_config())
    m.load_state_dict(oracle_sd, strict=True)
    m.eval()
    assert next(m.parameters()).is_cuda
    return m


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


@pytest.mark.parametrize("name", ["model_b1_t32", "model_b2_t64", "model_b3_t96"])
def test_forward_vs_reference_golden(model, golden_dir, name):
    g = _load(golden_dir, name)
    b, T, seed = (int(v) for v in g["meta"])
    x, _, t = O.synthetic_inputs(b, T, seed)
    out = model(x.to(DEV), t.to(DEV))
    assert out.shape == x.shape and out.dtype == torch.float32 and out.is_cuda
    ref = torch.from_numpy(g["out"])
    r, m = rel_l2(out.cpu(), ref), float((out.cpu() - ref).abs().max())
    print(f"{name}: rel-l2 {r:.3e} max-abs {m:.3e} (ref std {float(ref.std()):.3f})")
    assert r <= EPS_REL and m <= EPS_ABS_PER_STD * float(ref.std())


def test_forward_vs_oracle_t128_mixed_t(model, oracle_sd):
    x, _, t = O.synthetic_inputs(2, 128, 77)
    ref = O.OracleModel(oracle_sd)(x, t)
    out = model(x.to(DEV), t.to(DEV)).cpu()
    r, m = rel_l2(out, ref), float((out - ref).abs().max())
    print(f"t128: rel-l2 {r:.3e} max-abs {m:.3e}")
    assert r <= EPS_REL and m <= EPS_ABS_PER_STD * float(ref.std())


def test_forward_is_reproducible_and_batch_independent(model):
    # every reduction on the path has a fixed order except the final double-precision atomics: two runs agree (almost) bitwise
    x, _, t = O.synthetic_inputs(3, 64, 5)
    a = model(x.to(DEV), t.to(DEV))
    a2 = model(x.to(DEV), t.to(DEV))
    assert rel_l2(a2, a) < 1e-3
    # samples never mix.  A different batch size changes which CTA sums which tile, i.e. the fp32
    # summation order of the GroupNorm statistics (~1e-7), and bf16 rounding of the activations amplifies that through 32
    # residual blocks up to the path's own rounding noise -- so the bound is the model tolerance, not zero.
    b = model(x[1:2].to(DEV), t[1:2].to(DEV))
    assert rel_l2(a[1:2], b) < EPS_REL


@pytest.mark.parametrize("name", ["ddim_b2_t32_eta0", "ddim_b1_t64_eta1"])
def test_generalized_steps_vs_reference_golden(model, golden_dir, name):
    from ddim_audio_b200.functions.denoising import generalized_steps
    g = _load(golden_dir, name)
    b, T, seed, nsteps = (int(v) for v in g["meta"])
    eta = float(g["eta"])
    sel = set(int(v) for v in g["sel"]) if bool(g["has_sel"]) else None
    x, _, _ = O.synthetic_inputs(b, T, seed)
    seq = O.uniform_seq(1000, nsteps)
    gen = torch.Generator().manual_seed(seed + 1000)
    noise = [torch.randn(x.shape, generator=gen) for _ in seq]
    xd = x.to(DEV)
    xs, x0 = generalized_steps(xd, seq, model, O.alphas_cumprod().to(DEV), sel, eta=eta, noise=noise)
    assert torch.equal(xd.cpu(), x), "input must not be mutated"
    assert len(xs) - 1 == g["xs"].shape[0] and len(x0) == g["x0"].shape[0]
    assert all(not v.is_cuda for v in xs[1:] + x0)
    worst = 0.0
    for got, want in zip(xs[1:] + x0, list(g["xs"]) + list(g["x0"])):
        want = torch.from_numpy(want)
        worst = max(worst, rel_l2(got, want))
    print(f"{name}: worst snapshot rel-l2 {worst:.3e}")
    assert worst <= 5e-2       # error compounds over the steps; the final-sample error is reported by bench.py


def test_loss_vs_reference_golden(model, golden_dir):
    from ddim_audio_b200.functions.losses import loss_registry
    g = _load(golden_dir, "loss_b4_t32")
    b, T, seed = (int(v) for v in g["meta"])
    x0, e, t = O.synthetic_inputs(b, T, seed)
    a = O.alphas_cumprod().to(DEV)
    with torch.no_grad():
        per = loss_registry["simple"](model, x0.to(DEV), t.to(DEV), e.to(DEV), a, keepdim=True)
        loss = loss_registry["simple"](model, x0.to(DEV), t.to(DEV), e.to(DEV), a)
    np.testing.assert_allclose(per.cpu().numpy(), g["per_sample"], rtol=3e-2)
    assert float(loss) == pytest.approx(float(g["loss"]), rel=3e-2)


@pytest.mark.parametrize("B,chunks", [(9, 4), (8, 3), (2, None)])
def test_generalized_steps_host_input_is_pipelined_by_sample_groups(model, B, chunks):
    """A pinned HOST x runs as independent sample groups (H2D / snapshots overlapped with the denoiser): the lists must equal
    the single-group run on a device tensor sample for sample (samples never interact: SURVEY.md section 8e)."""
    from ddim_audio_b200.functions.denoising import generalized_steps
    x, _, _ = O.synthetic_inputs(B, 32, 321)
    seq = O.uniform_seq(1000, 3)
    al = O.alphas_cumprod().to(DEV)
    kw = {} if chunks is None else {"chunks": chunks}
    xs_h, x0_h = generalized_steps(x.pin_memory(), seq, model, al, {0, 2}, eta=0.0, **kw)
    xs_d, x0_d = generalized_steps(x.to(DEV), seq, model, al, {0, 2}, eta=0.0)
    assert len(xs_h) == len(xs_d) == 3 and len(x0_h) == len(x0_d) == 2
    assert not xs_h[0].is_cuda and torch.equal(xs_h[0], x)
    for a, b in zip(xs_h[1:] + x0_h, xs_d[1:] + x0_d):
        assert a.shape == b.shape == x.shape and not a.is_cuda
        assert rel_l2(a, b) < 1e-3    # identical arithmetic per sample; only the fp64 statistics atomics may reorder


PROBE_T, PROBE_F = 61, 7      # oracle/make_golden_fullshape.py


def test_forward_t8192_vs_reference_golden_and_oracle(model, oracle_sd, golden_dir):
    """The sampling geometry x[1,2,8192,256] (audio.yml:77): strided probe + norm of the UNMODIFIED reference's output
    (tests/golden/model_b1_t8192.npz), and the full tensor against the oracle run on this box's CPU."""
    g = _load(golden_dir, "model_b1_t8192")
    b, T, seed = (int(v) for v in g["meta"])
    x, _, t = O.synthetic_inputs(b, T, seed)
    assert np.array_equal(t.numpy(), g["t"])
    out = model(x.to(DEV), t.to(DEV)).cpu()
    ref_p = torch.from_numpy(g["probe"])
    norm, _, _, std = (float(v) for v in g["summary"])
    got_p = out[..., ::PROBE_T, ::PROBE_F]
    r, m = rel_l2(got_p, ref_p), float((got_p - ref_p).abs().max())
    print(f"t8192 probe: rel-l2 {r:.3e} max-abs {m:.3e} (ref std {std:.3f}); norm {float(out.double().norm()):.4f} vs {norm:.4f}")
    assert r <= EPS_REL and m <= EPS_ABS_PER_STD * std
    assert abs(float(out.double().norm()) - norm) <= EPS_REL * norm
    ref = O.OracleModel(oracle_sd)(x, t)
    r, m = rel_l2(out, ref), float((out - ref).abs().max())
    print(f"t8192 full tensor vs oracle: rel-l2 {r:.3e} max-abs {m:.3e}")
    assert r <= EPS_REL and m <= EPS_ABS_PER_STD * float(ref.std())


def test_generalized_steps_configs0_geometry(model, oracle_sd, golden_dir):
    """BASELINE configs[0] geometry x[4,2,1024,256], last 3 steps of range(0,1000,20), eta 0: every snapshot against the
    reference-minted probes, and the per-step eps teacher-forced on the oracle's own trajectory (run on this box's CPU):
    rel-L2 <= 2e-2, max-abs <= 0.114 std."""
    from ddim_audio_b200.functions.denoising import generalized_steps
    g = _load(golden_dir, "ddim_b4_t1024_s3")
    b, T, seed, nsteps = (int(v) for v in g["meta"])
    seq = [int(v) for v in g["seq"]]
    assert seq == list(range(0, 1000, 20))[-nsteps:]
    x, _, _ = O.synthetic_inputs(b, T, seed)
    al = O.alphas_cumprod()
    xs, x0 = generalized_steps(x.to(DEV), seq, model, al.to(DEV), None, eta=0.0)
    assert len(xs) == nsteps + 1 and len(x0) == nsteps
    worst = 0.0
    for got, want, summ in zip(xs[1:] + x0, list(g["xs_probe"]) + list(g["x0_probe"]), list(g["xs_summary"]) + list(g["x0_summary"])):
        worst = max(worst, rel_l2(got[..., ::PROBE_T, ::PROBE_F], torch.from_numpy(want)))
        assert abs(float(got.double().norm()) - float(summ[0])) <= 5e-2 * float(summ[0])
    print(f"configs[0] geometry: worst snapshot probe rel-l2 {worst:.3e}")
    assert worst <= 5e-2
    # teacher-forced per-step eps
    omodel = O.OracleModel(oracle_sd)
    xt = x.clone()
    t = torch.zeros(b, dtype=torch.long)
    for c in O.ddim_coefficients(al, seq, 0.0):
        t[...] = c["i"]
        et = omodel(xt, t)
        et_g = model(xt.to(DEV), t.to(DEV)).cpu()
        r, m = rel_l2(et_g, et), float((et_g - et).abs().max() / et.std())
        print(f"  step t={c['i']}: eps rel-l2 {r:.3e} max-abs/std {m:.3e}")
        assert r <= EPS_REL and m <= EPS_ABS_PER_STD
        xt.add_(et, alpha=-((1 - c["at"]) ** 0.5)).div_(c["at"] ** 0.5)
        xt.mul_(c["at_next"] ** 0.5).add_(et, alpha=c["c2"])


def test_generalized_steps_device_groups_and_graph_replay(model):
    """Runtime around the loop must not change a result: a device-resident batch walked as two sample groups (what happens when a
    shard does not fit beside its activations) and the CUDA-graph replay of the forward both equal the plain single-group run
    sample for sample (identical kernels per sample; only the fp64 statistics atomics may reorder)."""
    from ddim_audio_b200.functions.denoising import generalized_steps
    x, _, _ = O.synthetic_inputs(5, 32, 411)
    seq = O.uniform_seq(1000, 4)
    al = O.alphas_cumprod().to(DEV)
    ref = generalized_steps(x.to(DEV), seq, model, al, {1, -1}, eta=0.0, graph=False)
    for kw in (dict(chunks=2, graph=False), dict(graph=True), dict(chunks=2, graph=True)):
        got = generalized_steps(x.to(DEV), seq, model, al, {1, -1}, eta=0.0, **kw)
        assert len(got[0]) == len(ref[0]) == 3 and len(got[1]) == len(ref[1]) == 2
        for a, b in zip(got[0][1:] + got[1], ref[0][1:] + ref[1]):
            assert a.shape == b.shape == x.shape and not a.is_cuda
            assert rel_l2(a, b) < 1e-3, kw
    # a second call re-uses the captured graph (same shape, same weight version) and may recycle the pinned snapshot tensors
    again = generalized_steps(x.to(DEV), seq, model, al, {1, -1}, eta=0.0, graph=True, out=got)
    assert all(p.data_ptr() == q.data_ptr() for p, q in zip(again[0][1:] + again[1], got[0][1:] + got[1]))
    for a, b in zip(again[0][1:] + again[1], ref[0][1:] + ref[1]):
        assert rel_l2(a, b) < 1e-3


def test_generalized_steps_match_reference_rng(model):
    """functions/denoising.py:41-43 draws torch.randn_like(x) every step even when its coefficient is zero (eta = 0).  With
    match_reference_rng=True the CUDA generator ends where the reference leaves it: exactly len(seq) draws of x's shape."""
    from ddim_audio_b200.functions.denoising import generalized_steps
    x, _, _ = O.synthetic_inputs(2, 32, 412)
    seq = O.uniform_seq(1000, 3)
    al = O.alphas_cumprod().to(DEV)
    xd = x.to(DEV)
    torch.cuda.manual_seed(99)
    a = generalized_steps(xd, seq, model, al, None, eta=0.0, match_reference_rng=True)
    after = torch.cuda.get_rng_state(xd.device)
    torch.cuda.manual_seed(99)
    for _ in seq:
        torch.randn_like(xd)                                   # what the reference's loop consumes
    assert torch.equal(after, torch.cuda.get_rng_state(xd.device))
    torch.cuda.manual_seed(99)
    b = generalized_steps(xd, seq, model, al, None, eta=0.0)  # default: no draw at eta = 0, same samples
    assert torch.equal(torch.cuda.get_rng_state(xd.device), torch.cuda.get_rng_state(xd.device))
    for p, q in zip(a[0][1:], b[0][1:]):
        assert rel_l2(p, q) < 1e-3
