"""Load the UNMODIFIED reference (klae01/ddim-audio) read-only from /root/reference, on CPU.

TEST INFRASTRUCTURE ONLY -- used by oracle/make_golden.py (build container) to mint the
golden fixtures and to validate oracle/ddim_oracle.py.  /root/reference does not exist on
the GPU box; nothing under tests -m gpu, smoke() or bench.py imports this module.

Shims, none of which touch reference source:
  1. configs/audio.yml:26,42 hard-code ``torch.cuda.FloatTensor`` -> overridden to
     ``torch.FloatTensor`` in the loaded YAML dict.
  2. functions/denoising.py:18-19 call ``x.type("torch.cuda.FloatTensor")`` -> a scoped
     patch of ``torch.Tensor.type`` maps ``torch.cuda.*Tensor`` strings to their CPU types.
  3. on CPU ``xt.to("cpu")`` aliases ``xt`` (every snapshot would be the final tensor) ->
     a scoped patch of ``torch.Tensor.to`` clones when the result is the same object.
"""
from __future__ import annotations

import contextlib
import os
import sys

import torch
import yaml

REF = os.environ.get("DDIM_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF, "models", "diffusion.py"))


@contextlib.contextmanager
def _ref_cwd():
    """models/diffusion.py:7 does sys.path.append("External") relative to the cwd."""
    old_cwd, old_path = os.getcwd(), list(sys.path)
    sys.dont_write_bytecode = True
    os.chdir(REF)
    sys.path.insert(0, REF)
    sys.path.insert(1, os.path.join(REF, "External"))
    try:
        yield
    finally:
        os.chdir(old_cwd)
        sys.path[:] = old_path


def load_config():
    with _ref_cwd():
        from utils import dict2namespace  # reference utils.py:4-12

        cfg = yaml.safe_load(open(os.path.join(REF, "configs", "audio.yml")))
    cfg["model"]["dtype"] = "torch.FloatTensor"
    cfg["model"]["transformers"]["dtype"] = "torch.FloatTensor"
    return dict2namespace(cfg)


def load_model(state_dict=None, eval_mode=True):
    config = load_config()
    with _ref_cwd():
        from models.diffusion import Model

        model = Model(config)
    if state_dict is not None:
        model.load_state_dict(state_dict, strict=True)
    if eval_mode:
        model.eval()
    return model, config


def reference_functions():
    with _ref_cwd():
        from functions.denoising import generalized_steps
        from functions.losses import noise_estimation_loss
    return generalized_steps, noise_estimation_loss


@contextlib.contextmanager
def cpu_shims(noise=None):
    """Scoped patches 2 and 3 above; ``noise`` (iterable of tensors) replaces the draws of
    ``torch.randn_like`` inside the block so eta>0 runs can be replayed elsewhere."""
    orig_type, orig_to, orig_randn_like = torch.Tensor.type, torch.Tensor.to, torch.randn_like
    noise_it = iter(noise) if noise is not None else None

    def type_shim(self, dtype=None, *a, **k):
        if isinstance(dtype, str) and dtype.startswith("torch.cuda."):
            dtype = dtype.replace("torch.cuda.", "torch.")
        return orig_type(self, dtype, *a, **k)

    def to_shim(self, *a, **k):
        out = orig_to(self, *a, **k)
        if out is self and a and a[0] == "cpu":
            out = self.clone()
        return out

    def randn_like_shim(x, *a, **k):
        if noise_it is not None:
            return next(noise_it).to(x.dtype)
        return orig_randn_like(x, *a, **k)

    torch.Tensor.type, torch.Tensor.to, torch.randn_like = type_shim, to_shim, randn_like_shim
    try:
        yield
    finally:
        torch.Tensor.type, torch.Tensor.to, torch.randn_like = orig_type, orig_to, orig_randn_like
