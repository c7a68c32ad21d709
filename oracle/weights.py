"""Oracle: fixed-seed Keras-layout weights, generated without touching the product library.

ImageNet weights (/root/reference/models.py:26, :51 ``weights='imagenet'``) need the network, so both the
engine and the oracle run on the same fixed-seed random initialisation in Keras layout.  This is the oracle's
own copy of the generator (same presets, same draws in the same order as
distributed-machine-learning_b200/keras_weights.py; tests/test_oracle_cpu.py compares the two), so that
bench.py's reference arm and the CPU tests never load libdml_b200.so.  Test infrastructure only.
"""
from __future__ import annotations

import numpy as np

from . import keras_nets

# the presets tests/golden/make_golden.py settled on (largest minimum top-1 margin over the bundled images)
DEFAULT_PRESET = {"ResNet50": ("keras_default", 0), "InceptionV3": ("he_uniform", 7)}


def generate(model: str, preset: str | None = None, seed: int | None = None):
    """List of fp32 arrays in blob order (layer creation order, get_weights() order per layer)."""
    dp, ds = DEFAULT_PRESET[model]
    preset = preset or dp
    seed = ds if seed is None else seed
    rng = np.random.default_rng(seed)
    out = []
    for name, shape in keras_nets.weight_spec(model):
        kind = name.rsplit("/", 1)[1]
        if kind == "kernel":
            if len(shape) == 4:
                kh, kw, cin, cout = shape
                fan_in, fan_out = kh * kw * cin, kh * kw * cout
            else:
                fan_in, fan_out = shape
            if preset == "he_uniform" and len(shape) == 4:
                lim = np.sqrt(6.0 / fan_in)
            else:
                lim = np.sqrt(6.0 / (fan_in + fan_out))  # glorot_uniform, Keras' default
            a = rng.uniform(-lim, lim, size=shape)
        elif preset == "folding_stress":
            if kind in ("bias", "beta", "moving_mean"):
                a = rng.normal(0.0, 0.1, size=shape)
            elif kind in ("gamma", "moving_variance"):
                a = rng.uniform(0.5, 1.5, size=shape)
            else:
                raise ValueError(name)
        else:
            a = np.ones(shape) if kind in ("gamma", "moving_variance") else np.zeros(shape)
        out.append(np.ascontiguousarray(a, dtype=np.float32))
    return out
