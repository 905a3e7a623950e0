"""The oracle (oracle/ddim_oracle.py) replayed against fixtures minted from the UNMODIFIED
reference by oracle/make_golden.py.  CPU only; this is what pins the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import ddim_oracle as O

# the oracle restates the same torch ops in the same order, so it reproduces the reference to the last bit on
# the machine that minted the fixtures; a different host CPU may pick other oneDNN kernels -> fp32 noise floor
TOL = dict(rtol=0, atol=2e-4)


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def test_state_dict_spec_matches_reference_layout(oracle_sd, golden_dir):
    spec = O.state_dict_spec()
    assert len(spec) == 389                                   # SURVEY.md §7 hard part 9
    assert sum(int(np.prod(s)) for k, s in spec.items() if k != "temb.te") == 47_155_266
    assert "up_modules.0.3.conv.weight" in spec and spec["up_modules.0.3.conv.weight"] == (256, 192, 4, 4)
    assert "down_modules.1.0.norm.2.bias" not in spec
    fp = _load(golden_dir, "weights_fingerprint")
    for k, s in zip(fp["keys"], fp["sums"]):
        assert float(oracle_sd[str(k)].double().sum()) == pytest.approx(float(s), rel=1e-12, abs=1e-12)


def test_schedule(golden_dir):
    assert np.array_equal(O.alphas_cumprod().numpy(), _load(golden_dir, "schedule")["alphas"])


@pytest.mark.parametrize("name", ["model_b2_t64", "model_b1_t32", "model_b3_t96"])
def test_model_forward(oracle_sd, golden_dir, name):
    g = _load(golden_dir, name)
    b, T, seed = (int(v) for v in g["meta"])
    x, _, t = O.synthetic_inputs(b, T, seed)
    assert np.array_equal(t.numpy(), g["t"])
    out = O.OracleModel(oracle_sd)(x, t)
    np.testing.assert_allclose(out.numpy(), g["out"], **TOL)


@pytest.mark.parametrize("name", ["ddim_b2_t32_eta0", "ddim_b1_t64_eta1"])
def test_generalized_steps(oracle_sd, golden_dir, name):
    g = _load(golden_dir, name)
    b, T, seed, nsteps = (int(v) for v in g["meta"])
    eta = float(g["eta"])
    sel = set(int(v) for v in g["sel"]) if bool(g["has_sel"]) else None
    x, _, _ = O.synthetic_inputs(b, T, seed)
    seq = O.uniform_seq(1000, nsteps)
    gen = torch.Generator().manual_seed(seed + 1000)
    noise = [torch.randn(x.shape, generator=gen) for _ in seq]
    xs, x0 = O.generalized_steps(x, seq, O.OracleModel(oracle_sd), O.alphas_cumprod(), sel, noise=noise, eta=eta)
    assert torch.equal(xs[0], x)
    assert len(xs) - 1 == g["xs"].shape[0] and len(x0) == g["x0"].shape[0]
    for got, want in zip(xs[1:], g["xs"]):
        np.testing.assert_allclose(got.numpy(), want, rtol=0, atol=5e-4)
    for got, want in zip(x0, g["x0"]):
        np.testing.assert_allclose(got.numpy(), want, rtol=0, atol=5e-4)


def test_loss(oracle_sd, golden_dir):
    g = _load(golden_dir, "loss_b4_t32")
    b, T, seed = (int(v) for v in g["meta"])
    x0, e, t = O.synthetic_inputs(b, T, seed)
    m = O.OracleModel(oracle_sd)
    a = O.alphas_cumprod()
    assert float(O.noise_estimation_loss(m, x0, t, e, a)) == pytest.approx(float(g["loss"]), rel=1e-5)
    np.testing.assert_allclose(O.noise_estimation_loss(m, x0, t, e, a, keepdim=True).numpy(), g["per_sample"], rtol=1e-5)


def test_ddim_coefficients_edge_cases():
    a = O.alphas_cumprod()
    co = O.ddim_coefficients(a, range(0, 1000, 20), eta=0.0)
    assert len(co) == 50 and co[0]["i"] == 980 and co[-1]["j"] == -1
    assert co[-1]["at_next"] == 1.0 and co[-1]["c2"] == 0.0          # functions/denoising.py:12: alpha[0] = 1.0
    assert all(c["c1"] == 0.0 for c in co)
    co1 = O.ddim_coefficients(a, range(0, 1000, 250), eta=1.0)
    assert all(c["c1"] > 0 for c in co1[:-1]) and co1[-1]["c1"] == 0.0
    assert O.select_index_for(1000, 10) == {999, 888, 777, 666, 555, 444, 333, 222, 111, 0}


def test_loss_gradients_match_reference_fingerprint(oracle_sd, golden_dir):
    """autograd through the oracle vs gradients of the UNMODIFIED reference (per-parameter norm / sum / seeded projection)."""
    g = _load(golden_dir, "grad_b2_t32")
    b, T, seed = (int(v) for v in g["meta"])
    x0, e, t = O.synthetic_inputs(b, T, seed)
    loss, grads = O.loss_and_grads(oracle_sd, x0, t, e, O.alphas_cumprod())
    assert float(loss) == pytest.approx(float(g["loss"]), rel=1e-5)
    keys = [str(k) for k in g["keys"]]
    assert keys == [k for k in oracle_sd.keys() if k != "temb.te"] and len(keys) == 388
    for k, fp in zip(keys, g["fp"]):
        norm, s, proj = O.grad_fingerprint(k, grads[k])
        assert norm == pytest.approx(fp[0], rel=2e-3, abs=1e-6), k
        assert abs(proj - fp[2]) <= 2e-3 * max(fp[0], 1e-6) + 1e-6, k      # projection on a unit-variance direction ~ norm


@pytest.mark.parametrize("name", ["denoise_b2_t48", "denoise_b1_t33_f128"])
def test_denoise_2d(golden_dir, name):
    """the post-filter restatement against outputs of the reference's own denoise_2d
    (External/UPU/signal/denoise.py:45-70), minted by oracle/make_golden_denoise.py."""
    g = _load(golden_dir, name)
    b, T, F_, seed = (int(v) for v in g["meta"])
    x = O.denoise_inputs(b, T, seed, f_size=F_)
    got = O.denoise_2d(x)
    np.testing.assert_allclose(got.numpy(), g["out"], rtol=0, atol=1e-5)


def test_lower_median_is_torch_median():
    g = torch.Generator().manual_seed(3)
    for n in (1, 2, 7, 8, 33):
        x = torch.randn(3, n, 5, generator=g)
        assert torch.equal(O.lower_median(x, 1), x.median(dim=1, keepdim=True)[0])


def test_ema_update(golden_dir):
    """the EMA restatement against the final shadow of the reference's own EMAHelper (models/ema.py:16-23)
    after three updates, minted by oracle/make_golden_ema.py."""
    g = _load(golden_dir, "ema_steps")
    mu = float(g["mu"])
    _, traj = O.ema_case(int(g["seed"]))
    shadow = {k: v.clone() for k, v in traj[0].items()}
    for step in traj[1:]:
        shadow = O.ema_update(shadow, step, mu)
    for k, v in shadow.items():
        assert np.array_equal(v.numpy(), g[k.replace(".", "__")]), k


@pytest.mark.parametrize("name", ["adabelief", "adamw"])
def test_optimizer_tail(golden_dir, name):
    """clip_grad_norm_ + AdaBelief / AdamW restatements against three steps of the reference's own
    optimizers (oracle/make_golden_optim.py).  fp32 rounding only (torch's kernels fuse some multiply-adds)."""
    g = _load(golden_dir, "optim_steps")
    cfg = {"adabelief": dict(lr=0.0003, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.00001, warmup=1000),
           "adamw": dict(lr=0.0005, betas=(0.9, 0.998), eps=1e-6, weight_decay=0.0001, warmup=10000)}[name]
    params, grads = O.optim_case(int(g["seed"]))
    m = {k: torch.zeros_like(v) for k, v in params.items()}
    v = {k: torch.zeros_like(v_) for k, v_ in params.items()}
    fn = O.adabelief_step if name == "adabelief" else O.adamw_step
    for i, gr in enumerate(grads):
        total, cg = O.clip_grad_norm(gr, 1.0)
        assert float(total) == pytest.approx(float(g[f"{name}__norms"][i]), rel=1e-6)
        lr = cfg["lr"] * min(((1 + i) / cfg["warmup"]) ** -0.5, (1 + i) / cfg["warmup"])
        for k in params:
            params[k], m[k], v[k] = fn(params[k], cg[k], m[k], v[k], i + 1, lr, cfg["betas"], cfg["eps"], cfg["weight_decay"])
    for k in params:
        kk = k.replace(".", "__")
        np.testing.assert_allclose(O.optim_probe(params[k]).numpy(), g[f"{name}__p__{kk}"], rtol=0, atol=1e-7)
        np.testing.assert_allclose(O.optim_probe(m[k]).numpy(), g[f"{name}__m__{kk}"], rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(O.optim_probe(v[k]).numpy(), g[f"{name}__v__{kk}"], rtol=1e-5, atol=1e-12)


def test_fullshape_goldens_forward_t8192(oracle_sd, golden_dir):
    """The oracle at the sampling geometry x[1,2,8192,256] against the reference-minted probe (oracle/make_golden_fullshape.py)."""
    g = _load(golden_dir, "model_b1_t8192")
    b, T, seed = (int(v) for v in g["meta"])
    x, _, t = O.synthetic_inputs(b, T, seed)
    out = O.OracleModel(oracle_sd)(x, t)
    np.testing.assert_allclose(out[..., ::61, ::7].numpy(), g["probe"], **TOL)
    assert float(out.double().norm()) == pytest.approx(float(g["summary"][0]), rel=1e-5)


def test_fullshape_goldens_ddim_configs0_geometry(oracle_sd, golden_dir):
    """The oracle's generalized_steps at BASELINE configs[0] geometry x[4,2,1024,256] (last 3 steps of range(0,1000,20))."""
    g = _load(golden_dir, "ddim_b4_t1024_s3")
    b, T, seed, nsteps = (int(v) for v in g["meta"])
    x, _, _ = O.synthetic_inputs(b, T, seed)
    xs, x0 = O.generalized_steps(x, [int(v) for v in g["seq"]], O.OracleModel(oracle_sd), O.alphas_cumprod(), None, eta=0.0)
    for got, want in zip(xs[1:] + x0, list(g["xs_probe"]) + list(g["x0_probe"])):
        np.testing.assert_allclose(got[..., ::61, ::7].numpy(), want, rtol=0, atol=5e-4)


@pytest.mark.parametrize("name", ["pfft2wav_b2_t64", "pfft2wav_b3_t97_loud", "pfft2wav_b1_t130_hpi"])
def test_pfft2wav_oracle_vs_reference_golden(golden_dir, name):
    """oracle/pfft2wav_oracle.py against int32 PCM minted from the reference's own pfft2wav (oracle/make_golden_pfft2wav.py): the
    restatement uses the same scipy pocketfft and the same in-place float32 operations, so it is bit-exact."""
    from oracle import pfft2wav_oracle as P
    g = _load(golden_dir, name)
    b, T, seed, hpi = (int(v) for v in g["meta"])
    x = P.synthetic_spectrogram(b, T, seed, scale=float(g["scale"]))
    for j in range(b):
        img = np.ascontiguousarray(np.transpose(x[j], (2, 1, 0)))
        with np.errstate(all="ignore"):
            got = P.pfft2wav(img, int(g["samplerate"]), dtype=np.int32, HPI=bool(hpi))
        # a different host CPU / scipy build may round the float32 FFT differently by an ulp (32 counts near the 3e8 peak)
        assert np.abs(got.astype(np.int64) - g["wav"][j].astype(np.int64)).max() <= 64
