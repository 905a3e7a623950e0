"""CPU-only checks of the host side: the C-ABI library loads and exports everything include/ddim_b200.h declares,
the drop-in Model owns the reference's parameters (names, shapes, default init), and the product path refuses to
run without CUDA instead of falling back."""
import os
import re

import numpy as np
import pytest
import torch

from oracle import ddim_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    from ddim_audio_b200 import _C
    return _C


def test_library_exports_every_declared_symbol(built):
    hdr = open(os.path.join(ROOT, "include", "ddim_b200.h")).read()
    declared = set(re.findall(r"\b(ddimb_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(built.lib, name), f"libddim_b200.so lacks {name}"
    assert declared == set(built.EXPORTS), declared ^ set(built.EXPORTS)
    assert built.version() >= 100


def test_model_state_dict_is_the_references(built):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    torch.manual_seed(0)
    m = Model(audio_config("torch.FloatTensor"))
    sd = m.state_dict()
    spec = O.state_dict_spec()
    assert list(sd.keys()) == list(spec.keys())
    assert all(tuple(sd[k].shape) == spec[k] for k in spec)
    # bit-identical default init to the reference under the same seed (fixture minted by oracle/make_golden.py)
    fp = np.load(os.path.join(ROOT, "tests", "golden", "init_fingerprint.npz"))
    assert [str(k) for k in fp["keys"]] == list(sd.keys())
    for k, s, a in zip(fp["keys"], fp["sums"], fp["abs_sums"]):
        v = sd[str(k)].double()
        assert float(v.sum()) == float(s) and float(v.abs().sum()) == float(a), str(k)
    assert float(sd["down_modules.1.0.norm.2.weight"].abs().sum()) == 0.0      # models/diffusion.py:25
    m.load_state_dict(O.make_state_dict(0), strict=True)


def test_no_cpu_fallback(built):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.functions.denoising import generalized_steps
    from ddim_audio_b200.functions.losses import noise_estimation_loss
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200 import ops
    m = Model(audio_config("torch.FloatTensor"))
    x = torch.zeros(1, 2, 32, 256)
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(x, torch.zeros(1, dtype=torch.long))
    with pytest.raises(RuntimeError, match="no CPU path"):
        noise_estimation_loss(m, x, torch.zeros(1, dtype=torch.long), x, torch.ones(1000))
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.ddim_step(x, x, None, 0.5, 0.6, 0.0, 0.1)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            generalized_steps(x, range(0, 1000, 500), m, torch.ones(1000), None)


def test_no_cpu_fallback_widened_rows(built):
    """the post-filter, the fused optimizers / clip and the EMA refuse CPU tensors instead of falling
    back to torch; the chunk-table builder (host logic) cuts tensors into <= CHUNK records with the right offsets."""
    import numpy as np
    import torch.nn as nn
    from ddim_audio_b200.functions import optim as fused
    from ddim_audio_b200.models.ema import EMAHelper
    from ddim_audio_b200.signal.denoise import denoise_2d
    with pytest.raises(RuntimeError, match="no CPU path"):
        denoise_2d(torch.zeros(1, 2, 8, 256))
    with pytest.raises(RuntimeError, match="no CPU"):
        denoise_2d(np.zeros((1, 2, 8, 256), np.float32))
    p = nn.Parameter(torch.zeros(8))
    p.grad = torch.ones(8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        fused.clip_grad_norm_([p], 1.0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        fused.AdamW([p]).step()
    with pytest.raises(RuntimeError, match="no CPU path"):
        fused.AdaBelief([p]).step()
    lin = nn.Linear(4, 4)
    h = EMAHelper(mu=0.9)
    h.register(lin)
    with pytest.raises(RuntimeError, match="no CPU path"):
        h.update(lin)
    a, b = torch.zeros(fused.CHUNK * 2 + 5), torch.zeros(7)
    table, total = fused.ChunkTable().get([None, [a, b], None, None], "cpu")
    assert total == a.numel() + b.numel() and table.shape == (4, 8)
    assert table[:, 4].tolist() == [fused.CHUNK, fused.CHUNK, 5, 7]
    assert table[:3, 1].tolist() == [a.data_ptr() + 4 * fused.CHUNK * i for i in range(3)] and int(table[3, 1]) == b.data_ptr()
    assert int(table[:, [0, 2, 3]].abs().sum()) == 0


def test_sampler_scalars_match_oracle():
    from ddim_audio_b200.functions.denoising import _ddim_scalars
    a = O.alphas_cumprod()
    al = [1.0] + a.numpy().tolist()
    for eta in (0.0, 0.5, 1.0):
        for c in O.ddim_coefficients(a, range(0, 1000, 100), eta):
            assert _ddim_scalars(al, c["i"], c["j"], eta) == (c["at"], c["at_next"], c["c1"], c["c2"])


def test_fourier_operands_reproduce_fft():
    # host-built DFT operands (bf16 hi+lo planes) vs torch.fft on the CPU, in float64 from the stored planes
    from ddim_audio_b200.models.diffusion import _fourier_operands
    S, H = 3, 512
    wh, a2, Sp = _fourier_operands(S, H, "cpu")
    whf = wh[:, :H].double() + wh[:, 2 * H:].double()
    cs = (a2[:, :Sp].double() + a2[:, 2 * Sp:3 * Sp].double())[:, :S]
    ss = (a2[:, Sp:2 * Sp].double() + a2[:, 3 * Sp:].double())[:, :S]
    x = torch.randn(S, H, dtype=torch.float64)
    u = whf @ x.T                                    # [2H, S]
    y = cs @ u[:H].T + ss @ u[H:].T                  # a2 stores -sin
    ref = torch.fft.fftn(x, dim=(0, 1)).real
    assert float((y - ref).abs().max()) < 1e-3


def test_operand_layouts_are_strided_gathers_of_the_masters():
    """models/diffusion.py::_gather_torch with the (dims, strides, offset) Model._pack records for ddimb_repack must equal the torch
    statements of the operand layouts (tap-major 3x3 / 4x4 kernels, flipped + transposed data-gradient twins, Linear transposes)."""
    import torch

    from ddim_audio_b200.models.diffusion import _gather_torch as G

    g = torch.Generator().manual_seed(11)
    Co, Ci = 5, 7
    w = torch.randn(Co, Ci, 3, 3, generator=g)
    assert torch.equal(G(w, (Co, 3, 3, Ci), (Ci * 9, 3, 1, 9), 0).reshape(Co, -1), w.permute(0, 2, 3, 1).reshape(Co, -1))
    assert torch.equal(G(w, (Ci, 3, 3, Co), (9, -3, -1, Ci * 9), 8).reshape(Ci, -1), w.flip(2, 3).permute(1, 2, 3, 0).reshape(Ci, -1))
    O, I = 6, 4
    w4 = torch.randn(O, I, 4, 4, generator=g)
    assert torch.equal(G(w4, (O, 4, 4, I), (I * 16, 4, 1, 16), 0).reshape(O, -1), w4.permute(0, 2, 3, 1).reshape(O, -1))
    assert torch.equal(G(w4, (4, 4, I, O), (4, 1, 16, I * 16), 0).reshape(16 * I, O), w4.permute(2, 3, 1, 0).reshape(-1, O))
    l = torch.randn(9, 4, generator=g)
    assert torch.equal(G(l, (1, 1, 9, 4), (0, 0, 4, 1), 0).reshape(9, 4), l)
    assert torch.equal(G(l, (1, 1, 4, 9), (0, 0, 1, 4), 0).reshape(4, 9), l.t())
    wo = torch.randn(2, 32, 3, 3, generator=g)
    assert torch.equal(G(wo, (32, 2, 3, 3), (9, 288, -3, -1), 8), wo.flip(2, 3).permute(1, 0, 2, 3))
