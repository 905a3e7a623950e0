"""-m gpu: the fused EMA update behind the reference's ``EMAHelper`` interface (models/ema.py), against
the golden shadow of the reference's own helper and against the oracle formula on the full denoiser.  Bar: BIT-EXACT (the
kernel keeps the reference's operation order and fp32 roundings)."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import ddim_oracle as O

pytestmark = pytest.mark.gpu


class Holder(nn.Module):
    def __init__(self, tensors):
        super().__init__()
        self.keys = list(tensors)
        for i, k in enumerate(self.keys):
            self.register_parameter(f"p{i}", nn.Parameter(tensors[k].clone()))

    def named_parameters(self, *a, **k):
        for i, key in enumerate(self.keys):
            yield key, getattr(self, f"p{i}")


def test_vs_reference_golden(golden_dir):
    from ddim_audio_b200 import _C
    from ddim_audio_b200.models.ema import EMAHelper

    g = np.load(os.path.join(golden_dir, "ema_steps.npz"))
    _, traj = O.ema_case(int(g["seed"]))
    m = Holder(traj[0]).cuda()
    h = EMAHelper(mu=float(g["mu"]))
    h.register(m)
    n0 = _C.launch_count()
    for step in traj[1:]:
        with torch.no_grad():
            for k, p in m.named_parameters():
                p.copy_(step[k])
        h.update(m)
    assert _C.launch_count() - n0 == len(traj) - 1            # one launch per update, whatever the number of tensors
    for k, v in h.state_dict().items():
        assert np.array_equal(v.cpu().numpy(), g[k.replace(".", "__")]), k


def test_full_model_bit_exact_and_interface(oracle_sd):
    """All 389-entry denoiser parameters: two updates equal the oracle formula bit for bit; ``ema`` / ``state_dict`` /
    ``load_state_dict`` / frozen parameters behave as models/ema.py."""
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.models.diffusion import Model
    from ddim_audio_b200.models.ema import EMAHelper

    model = Model(audio_config())
    model.load_state_dict(oracle_sd, strict=True)
    model.cuda()
    frozen = next(iter(model.parameters()))
    frozen.requires_grad_(False)
    h = EMAHelper(mu=0.9999)
    h.register(model)
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    assert set(h.shadow) == set(names) and len(names) == len(list(model.parameters())) - 1
    want = {k: v.clone() for k, v in h.shadow.items()}
    g = torch.Generator(device="cuda").manual_seed(3)
    for _ in range(2):
        with torch.no_grad():
            for p in model.parameters():
                p.add_(0.01 * torch.randn(p.shape, device="cuda", generator=g))
        h.update(model)
        want = O.ema_update(want, {n: p.data for n, p in model.named_parameters() if p.requires_grad}, 0.9999)
    for k in names:
        assert torch.equal(h.shadow[k], want[k]), k
    sd = h.state_dict()
    h2 = EMAHelper(mu=0.9999)
    h2.load_state_dict({k: v.clone() for k, v in sd.items()})
    h2.update(model)                                           # table rebuilt for the loaded shadow
    h.update(model)
    assert all(torch.equal(h.shadow[k], h2.shadow[k]) for k in names)
    h.ema(model)
    assert all(torch.equal(p.data, h.shadow[n]) for n, p in model.named_parameters() if p.requires_grad)


def test_errors():
    from ddim_audio_b200.models.ema import EMAHelper

    m = nn.Linear(4, 4)
    h = EMAHelper()
    h.register(m)
    with pytest.raises(RuntimeError):
        h.update(m)                                            # CPU parameters: no CPU path
