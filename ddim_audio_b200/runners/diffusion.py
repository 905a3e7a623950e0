"""``Diffusion`` runner behind the reference's ``Diffusion(args, config).train() / .sample()`` (runners/diffusion.py:90-523),
reduced to the hot path: schedules (:32-62, :91-128), ``sample_image`` (:475-520), ``sample_sequence`` (:377-398),
``train_step`` (:130-183) and ``train`` (:201-287).  Dataset loading and the STFT/PNG/WAV codec are out of scope
(SURVEY.md section 2): ``train`` takes any iterable of ``(x, y)`` batches and ``sample_sequence`` RETURNS the
``(xs, x0_preds)`` host tensors -- the x0 list post-filtered by ``denoise_2d`` when ``config.sampling.denoise`` (:400-401) --
optionally handing them to ``self.writer`` (what the reference does at :402-422).

Rank-aware: under torchrun ``sample_sequence`` shards ``num_samples`` by batch with no collective and
``train_step`` all-reduces one flat gradient buffer.  There is no CPU path.
"""
from __future__ import annotations

import logging
import os

import numpy as np
import torch

from ..functions.denoising import generalized_steps
from ..functions import optim as fused_optim
from ..functions.losses import loss_registry
from ..models.diffusion import Model
from ..models.ema import EMAHelper
from ..signal.denoise import denoise_2d
from ..signal.pfft2wav import pfft2wav_batch
from . import dist as rdist


def get_beta_schedule(beta_schedule, *, beta_start, beta_end, num_diffusion_timesteps):
    """runners/diffusion.py:32-62 (float64 numpy on the host)."""
    n = num_diffusion_timesteps
    if beta_schedule == "quad":
        betas = np.linspace(beta_start ** 0.5, beta_end ** 0.5, n, dtype=np.float64) ** 2
    elif beta_schedule == "linear":
        betas = np.linspace(beta_start, beta_end, n, dtype=np.float64)
    elif beta_schedule == "const":
        betas = beta_end * np.ones(n, dtype=np.float64)
    elif beta_schedule == "jsd":
        betas = 1.0 / np.linspace(n, 1, n, dtype=np.float64)
    elif beta_schedule == "sigmoid":
        betas = 1 / (np.exp(-np.linspace(-6, 6, n)) + 1) * (beta_end - beta_start) + beta_start
    else:
        raise NotImplementedError(beta_schedule)
    assert betas.shape == (n,)
    return betas


class _Group:
    def __init__(self, config):
        self.config = config
        self.params = []


def classify_group(config, model):
    """runners/diffusion.py:71-87: parameters grouped by their top-level module name."""
    from ..config import dict2namespace

    top, groups = {}, {}
    for name, sub in vars(config).items():
        sub = dict(vars(sub))
        for n in sub.pop("top_level_name"):
            top[n] = name
        groups[name] = _Group(dict2namespace(sub))
    for name, p in model.named_parameters():
        groups[top.get(name.split(".")[0], "default")].params.append(p)
    return {k: v for k, v in groups.items() if v.params}


def get_optimizer(config, parameters):
    """functions/__init__.py:5-50.  AdamW and AdaBelief (the two the reference's audio.yml configures) are the fused library
    optimizers of ``functions/optim.py`` on CUDA parameters; the other names stay stock torch objects on the fp32 masters."""
    import torch.optim as optim

    from ..functions import optim as fused

    parameters = list(parameters)
    on_cuda = bool(parameters) and all(p.is_cuda for p in parameters)
    kw = dict(lr=config.lr, weight_decay=config.weight_decay, betas=tuple(config.beta), amsgrad=config.amsgrad, eps=config.eps)
    if config.optimizer == "Adam":
        return optim.Adam(parameters, **kw)
    if config.optimizer == "AdamW":
        return fused.AdamW(parameters, **kw) if on_cuda and not config.amsgrad else optim.AdamW(parameters, **kw)
    if config.optimizer == "AdaBelief":
        if not on_cuda:
            raise RuntimeError("ddim_audio_b200: AdaBelief needs CUDA parameters (there is no CPU path)")
        return fused.AdaBelief(parameters, weight_decouple=True, fixed_decay=False, rectify=False, clip_step=config.clip_step,
                               norm_ord=config.norm_ord, **kw)
    if config.optimizer == "RMSProp":
        return optim.RMSprop(parameters, lr=config.lr, weight_decay=config.weight_decay)
    if config.optimizer == "SGD":
        return optim.SGD(parameters, lr=config.lr, momentum=0.9)
    raise NotImplementedError(f"Optimizer {config.optimizer} not understood.")


def get_scheduler(config, optimizer):
    """functions/__init__.py:53-60."""
    if config.warmup:
        from torch.optim.lr_scheduler import LambdaLR
        return LambdaLR(optimizer, lambda step: min(((1 + step) / config.warmup) ** -0.5, (1 + step) / config.warmup))
    return None


def antithetic_timesteps(n, num_timesteps, generator=None):
    """runners/diffusion.py:141-142 (host int64)."""
    t = torch.randint(low=0, high=num_timesteps, size=((n + 1) // 2,), generator=generator)
    return torch.cat([t, num_timesteps - t - 1], dim=0)[:n]


class Diffusion(object):
    def __init__(self, args, config, device=None):
        self.args = args
        # extension: --train_graph replays one captured (re-pack + forward + backward) per train step (fixed batch shape)
        self.train_graph = bool(getattr(args, "train_graph", False))
        self.config = config
        if device is None:
            if not torch.cuda.is_available():
                raise RuntimeError("ddim_audio_b200.Diffusion needs a CUDA device: there is no CPU path")
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        self.model_var_type = config.model.var_type
        betas = get_beta_schedule(beta_schedule=config.diffusion.beta_schedule, beta_start=config.diffusion.beta_start,
                                  beta_end=config.diffusion.beta_end, num_diffusion_timesteps=config.diffusion.num_diffusion_timesteps)
        # :107-115 -- cast to fp32 FIRST, then cumprod in fp32 (the reference's dtype string is a FloatTensor type)
        alphas = torch.from_numpy(np.concatenate([[1], 1.0 - betas], axis=-1)).float()
        betas = torch.from_numpy(betas).float()
        self.num_timesteps = betas.shape[0]
        alphas = alphas.cumprod(dim=0)
        alphas_cumprod, alphas_cumprod_prev = alphas[1:], alphas[:-1]
        posterior_variance = betas * (1.0 - alphas_cumprod_prev) / (1.0 - alphas_cumprod)
        if self.model_var_type == "fixedlarge":
            self.logvar = betas.log().to(self.device)
        elif self.model_var_type == "fixedsmall":
            self.logvar = posterior_variance.clamp(min=1e-20).log().to(self.device)
        self.alphas = alphas_cumprod.to(self.device)
        self.betas = betas.to(self.device)
        self.writer = None          # callable(xs, x0_preds) for the out-of-scope PNG/WAV stage

    # ------------------------------------------------------------------ sampling
    def sample_image(self, x, model, select_index=None):
        """:475-520 -- builds ``seq`` and runs the DDIM loop; returns ``(xs, x0_preds)``."""
        if getattr(self.args, "sample_type", "generalized") != "generalized":
            raise NotImplementedError("ddim_audio_b200: only sample_type 'generalized' (DDIM) is on the hot path; ddpm_noisy is out of scope")
        skip_type = getattr(self.args, "skip_type", "uniform")
        if skip_type == "uniform":
            seq = range(0, self.num_timesteps, self.num_timesteps // self.args.timesteps)
        elif skip_type == "quad":
            seq = [int(s) for s in list(np.linspace(0, np.sqrt(self.num_timesteps * 0.8), self.args.timesteps) ** 2)]
        else:
            raise NotImplementedError
        return generalized_steps(x, seq, model, self.alphas, eta=getattr(self.args, "eta", 0.0), select_index=select_index)

    def select_index(self):
        """:388-394."""
        if self.args.sequence in [-1, 0]:
            return range(self.args.timesteps)
        index = np.linspace(1, self.args.timesteps, self.args.sequence, dtype=np.int32)
        return set((self.args.timesteps - index).tolist())

    def sample_sequence(self, model, generator=None):
        """:377-398.  The whole ``num_samples`` batch is drawn ONCE (so an N-rank run equals the 1-rank run sample for sample)
        and this rank denoises its contiguous slice; rank 0 gets the concatenated host lists, other ranks ``None``."""
        cfg = self.config
        rank, world = rdist.rank_world()
        shape = (cfg.sampling.num_samples, cfg.model.channels, cfg.sampling.t_size, cfg.model.f_size)
        lo, hi = rdist.shard_bounds(shape[0], rank, world)
        if generator is None and world == 1:
            x = torch.randn(*shape, device=self.device)                               # :380-386, torch's CUDA stream
        else:
            if generator is None:
                generator = torch.Generator().manual_seed(int(getattr(self.args, "seed", 1234)))
            x = torch.randn(*shape, generator=generator)[lo:hi].to(self.device)       # identical on every rank, then sliced
        if hi > lo:
            with torch.no_grad():
                xs, x0 = self.sample_image(x, model, select_index=self.select_index())
        else:
            xs, x0 = [], []
        # :400-420 per predicted x0: denoise_2d, then (AUDIO) permute(0,3,2,1) + pfft2wav -> int32 PCM; both per sample, so they
        # shard like the loop.  The PNG preview (pfft2img) and the file names are host-side (out of scope).
        want_wav = getattr(getattr(cfg, "data", None), "dataset", None) == "AUDIO" and getattr(self, "render_wav", True)
        wavs = []
        for k, y in enumerate(x0):
            yd = y.to(self.device)
            if getattr(cfg.sampling, "denoise", False):
                yd = denoise_2d(yd)
            if want_wav:
                wavs.append(pfft2wav_batch(yd, cfg.sampling.virtual_samplerate, HPI=cfg.sampling.HPI).cpu())
            x0[k] = yd.cpu()
            del yd
        if world > 1:
            xs = rdist.gather_host_lists([t.cpu() for t in xs])
            x0 = rdist.gather_host_lists(x0)
            wavs = rdist.gather_host_lists(wavs) if want_wav else wavs
        self.wavs = wavs                      # list (one entry per selected step) of int32 [num_samples, (F-1)*(T-1)] PCM, rank 0
        if xs is not None and want_wav and getattr(self.args, "image_folder", None):
            self.write_wavs(wavs, self.args.image_folder)
        if self.writer is not None and xs is not None:
            self.writer(xs, x0)
        return xs, x0

    def write_wavs(self, wavs, folder):
        """:403-420 -- ``{sample}_{step:0{digits}d}.wav`` at the dataset's virtual sample rate (scipy.io.wavfile, host-side file IO)."""
        from scipy.io.wavfile import write as WAV_write

        os.makedirs(folder, exist_ok=True)
        digits = int(np.ceil(np.log10(len(wavs) + 1)))
        rate = self.config.data.dataset_kwargs.virtual_samplerate
        for i, batch in enumerate(wavs):
            for j, wav in enumerate(batch):
                WAV_write(os.path.join(folder, f"{j}_{i:0{digits}d}.wav"), rate, wav.numpy())

    def load_model(self, path=None):
        """:289-316 -- ``Model(config)`` + ``ckpt.pth`` in the reference's list format [model_sd, opt_sd, epoch, step, ema_sd]."""
        model = Model(self.config).to(self.device)
        if path is None:
            ckpt_id = getattr(self.config.sampling, "ckpt_id", None)
            path = os.path.join(self.args.log_path, "ckpt.pth" if ckpt_id is None else f"ckpt_{ckpt_id}.pth")
        states = torch.load(path, map_location=self.device)
        model.load_state_dict(states[0], strict=True)
        if self.config.model.ema and len(states) > 4:
            with torch.no_grad():
                shadow = states[-1]
                for name, p in model.named_parameters():                             # models/ema.py:25-30
                    if p.requires_grad and name in shadow:
                        p.copy_(shadow[name])
        return model

    def sample(self, model=None):
        model = self.load_model() if model is None else model
        model.eval()
        if getattr(self.args, "sequence", None) is None:
            raise NotImplementedError("Sample procedeure not defined")
        return self.sample_sequence(model)

    # ------------------------------------------------------------------ training
    def build_training(self, model):
        """:218-234 -- optimizers, schedulers, clip groups (EMA is the caller's object; any with .update(model) works)."""
        optimizers, schedulers = {}, {}
        for name, g in classify_group(self.config.optimization.optimizer, model).items():
            optimizers[name] = get_optimizer(g.config, g.params)
            s = get_scheduler(g.config, optimizers[name])
            if s:
                schedulers[name] = s
        grad_group = dict(classify_group(self.config.optimization.grad_norm, model))
        return optimizers, schedulers, grad_group

    def loss_and_backward(self, model, x, t=None, e=None):
        """:137-154 -- the hot part of ``train_step``: noise, antithetic t, loss, backward, DP gradient all-reduce."""
        rank, world = rdist.rank_world()
        n = x.size(0)
        x = x.to(self.device)
        if e is None:
            e = torch.randn_like(x)
        if t is None:
            t = antithetic_timesteps(n, self.num_timesteps)
        t = t.to(self.device)
        loss_fn = loss_registry[self.config.model.type]
        if getattr(self, "train_graph", False) and model.training:
            # one captured (re-pack + forward + backward) per batch shape, replayed per step (models/training.py::GraphedLossBackward)
            from ..models.training import GraphedLossBackward
            cache = model.__dict__.setdefault("_train_graphs", {})
            key = (tuple(x.shape), str(x.device), world)
            gl = cache.get(key)
            if gl is None:
                gl = cache[key] = GraphedLossBackward(model, self.alphas, x, t, e, world)
            loss = gl(x, t, e)
        else:
            loss = loss_fn(model, x, t, e, self.alphas)
            (loss / world if world > 1 else loss).backward()
        if world > 1:
            rdist.allreduce_gradients(list(model.parameters()), getattr(model, "flat_grad", None))
        return loss

    def train_step(self, model, x, optimizers, schedulers, grad_group, ema_helper, step, epoch, t=None, e=None):
        """:130-199.  ``t`` / ``e`` (extension, default None = drawn as the reference draws them, :137-142) inject the timesteps
        and the noise for reproducible runs."""
        if not model.training:
            model.train()
        if not getattr(self, "train_graph", False):        # a captured step owns static gradient tensors and overwrites them
            for opt in optimizers.values():
                opt.zero_grad(set_to_none=True)
        loss = self.loss_and_backward(model, x, t=t, e=e)
        loggings = {"step": step, "loss": float(loss.item())}
        for g in grad_group.values():
            if g.config.grad_clip is not None:
                fused_optim.clip_grad_norm_(g.params, g.config.grad_clip)               # :155-160, two launches per group
        for name, opt in optimizers.items():
            out = opt.step()
            if type(out) is dict:
                loggings.update({f"{k}_{name}": v for k, v in out.items() if v is not None and k != "loss"})
        for s in schedulers.values():
            s.step()
        logging.info(", ".join((f"{k}: {v:.4f}" if type(v) is float else f"{k}: {v}") for k, v in loggings.items()))
        if ema_helper is not None:
            ema_helper.update(model)
        rank, _ = rdist.rank_world()
        log_path = getattr(self.args, "log_path", None)
        if rank == 0 and log_path and (step % self.config.training.snapshot_freq == 0 or step == 1):
            states = [model.state_dict(), opt.state_dict(), epoch, step]
            if ema_helper is not None:
                states.append(ema_helper.state_dict())
            torch.save(states, os.path.join(log_path, f"ckpt_{step}.pth"))
            torch.save(states, os.path.join(log_path, "ckpt.pth"))
        return loggings

    def train(self, loader=None, model=None, ema_helper=None, max_steps=None):
        """:201-287 with the dataset injected: ``loader`` yields ``(x[B,2,T,F], y)``; under torchrun every rank passes ITS shard."""
        if loader is None:
            raise NotImplementedError("ddim_audio_b200: dataset loading is out of scope -- pass an iterable of (x, y) batches")
        model = Model(self.config).to(self.device) if model is None else model
        optimizers, schedulers, grad_group = self.build_training(model)
        if ema_helper is None and getattr(self.config.model, "ema", False):          # :232-236
            ema_helper = EMAHelper(mu=self.config.model.ema_rate)
            ema_helper.register(model)
        self.ema_helper = ema_helper
        n_iters = self.config.training.n_iters if max_steps is None else max_steps
        step, epoch, log = 0, 0, None
        while step < n_iters:
            for x, _ in loader:
                step += 1
                log = self.train_step(model, x, optimizers, schedulers, grad_group, ema_helper, step, epoch)
                if step >= n_iters:
                    break
            epoch += 1
        return model, log

    def test(self):
        pass
