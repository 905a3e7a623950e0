"""The hot-path keys of the reference's configs/audio.yml as a nested ``argparse.Namespace`` (what the reference's
``utils.dict2namespace`` produces, utils.py:4-12), for callers that have no YAML at hand (tests, bench, smoke).
A namespace loaded from the reference's own audio.yml works the same way."""
from __future__ import annotations

import argparse
import copy

AUDIO = {
    "data": {"dataset": "AUDIO", "pfft_format": "CTF", "num_workers": 0,
             "dataset_kwargs": {"f_size": 256, "t_size": 1024, "virtual_samplerate": 48000, "axis": "CTF", "HPI": False}},
    "model": {
        "dtype": "torch.cuda.FloatTensor",                                  # audio.yml:26
        "type": "simple",
        "transformers": {
            "imports": "import transformers; from transformers.models.fnet.modeling_fnet import FNetEncoder",
            "module": "FNetEncoder",
            "config": "transformers.FNetConfig",
            "kwargs": {"hidden_size": 512, "num_hidden_layers": 12, "intermediate_size": 2048, "hidden_act": "gelu_new",
                       "hidden_dropout_prob": 0.1, "initializer_range": 0.02, "layer_norm_eps": 1e-6},
            "channels": 512,
            "dtype": "torch.cuda.FloatTensor",                              # audio.yml:42
        },
        "channels": 2, "t_size": 1024, "f_size": 256,
        "ch": [32, 64, 96, 128, 192, 256], "krn": [3, 3, 3, 3, 3, 3], "res": [2, 2, 3, 3, 3, 3],
        "var_type": "fixedlarge", "ema_rate": 0.9999, "ema": True,
    },
    "diffusion": {"beta_schedule": "linear", "beta_start": 0.0001, "beta_end": 0.02, "num_diffusion_timesteps": 1000},
    "training": {"batch_size": 14, "n_epochs": None, "n_iters": 5000000, "snapshot_freq": 5000, "validation_freq": 2000},
    "sampling": {"batch_size": 64, "last_only": True, "denoise": True, "HPI": False, "virtual_samplerate": 48000,
                 "num_samples": 2, "t_size": 8192},
    "optimization": {
        "optimizer": {
            "transformer": {"top_level_name": ["transformer"], "weight_decay": 0.0001, "optimizer": "AdamW", "warmup": 10000,
                            "lr": 0.0005, "beta": [0.9, 0.998], "amsgrad": False, "eps": 1e-6},
            "default": {"top_level_name": [], "weight_decay": 0.00001, "optimizer": "AdaBelief", "warmup": 1000, "lr": 0.0003,
                        "beta": [0.9, 0.999], "amsgrad": False, "eps": 1e-8, "clip_step": None, "norm_ord": 2},
        },
        "grad_norm": {"transformer": {"top_level_name": [], "grad_clip": 1}, "default": {"top_level_name": [], "grad_clip": 1}},
    },
}


def dict2namespace(d):
    ns = argparse.Namespace()
    for k, v in d.items():
        setattr(ns, k, dict2namespace(v) if isinstance(v, dict) else v)
    return ns


def audio_config(device_dtype: str = "torch.cuda.FloatTensor"):
    """``device_dtype='torch.FloatTensor'`` builds the parameters on the CPU (host-side tests); forward still needs CUDA."""
    d = copy.deepcopy(AUDIO)
    d["model"]["dtype"] = device_dtype
    d["model"]["transformers"]["dtype"] = device_dtype
    return dict2namespace(d)
