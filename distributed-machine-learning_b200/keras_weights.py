"""Keras-layout weight blobs for the two models.

The reference loads ImageNet weights through Keras (/root/reference/models.py:26,
:51), which needs the network.  Offline, the engine takes a flat fp32 blob in
Keras layout (layer creation order, each layer in get_weights() order; see
dml_model_weight_info) from a file, or a fixed-seed random initialisation:

  keras_default  - what ``weights=None`` gives: glorot-uniform kernels, zero
                   biases, identity BatchNorm.
  he_uniform     - U(-l, l), l = sqrt(6 / fan_in); keeps InceptionV3's
                   activations O(1) (no residuals there).
  folding_stress - glorot kernels with random biases and BN statistics, to
                   exercise BN/bias folding (fp32 check mode).
"""
from __future__ import annotations

import os
import re
import warnings

import numpy as np

from . import _lib

# Seeds chosen by tests/golden/make_golden.py (largest minimum top-1 margin over
# the bundled images among the seeds it tried).
DEFAULT_PRESET = {"ResNet50": ("keras_default", 0), "InceptionV3": ("he_uniform", 7)}


def _fans(shape):
    if len(shape) == 4:
        kh, kw, cin, cout = shape
        return kh * kw * cin, kh * kw * cout
    return shape[0], shape[1]


def generate(model_name: str, preset: str | None = None, seed: int | None = None):
    """List of fp32 arrays in blob order."""
    dp, ds = DEFAULT_PRESET[model_name]
    preset = preset or dp
    seed = ds if seed is None else seed
    rng = np.random.default_rng(seed)
    out = []
    for name, shape in _lib.weight_spec(model_name):
        kind = name.rsplit("/", 1)[1]
        if kind == "kernel":
            fan_in, fan_out = _fans(shape)
            if preset == "he_uniform" and len(shape) == 4:
                lim = np.sqrt(6.0 / fan_in)
            else:
                lim = np.sqrt(6.0 / (fan_in + fan_out))
            a = rng.uniform(-lim, lim, size=shape)
        elif preset == "folding_stress":
            if kind == "bias":
                a = rng.normal(0.0, 0.1, size=shape)
            elif kind == "gamma":
                a = rng.uniform(0.5, 1.5, size=shape)
            elif kind == "beta":
                a = rng.normal(0.0, 0.1, size=shape)
            elif kind == "moving_mean":
                a = rng.normal(0.0, 0.1, size=shape)
            elif kind == "moving_variance":
                a = rng.uniform(0.5, 1.5, size=shape)
            else:
                raise ValueError(name)
        else:
            a = np.ones(shape) if kind in ("gamma", "moving_variance") else np.zeros(shape)
        out.append(np.ascontiguousarray(a, dtype=np.float32))
    return out


def to_blob(arrays) -> np.ndarray:
    return np.concatenate([np.asarray(a, dtype=np.float32).ravel() for a in arrays])


def _natural_key(k: str):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", k)]


def arrays_from_npz(model_name: str, z) -> list:
    """The arrays of an .npz in blob order, validated one by one against dml_model_weight_info.

    Accepted key conventions: the weight_spec names themselves ("conv1_conv/kernel", ...; "/" may be written "__"),
    or positional keys in natural numeric order (numpy's arr_0, arr_1, ..., arr_319 -- NOT lexicographic, which would
    put arr_10 before arr_2).  A wrong count or any shape mismatch raises: a scrambled checkpoint must not load."""
    spec = _lib.weight_spec(model_name)
    files = list(z.files)
    by_name = {k.replace("__", "/"): k for k in files}
    if all(name in by_name for name, _ in spec):
        keys = [by_name[name] for name, _ in spec]
    else:
        keys = sorted(files, key=_natural_key)
    if len(keys) != len(spec):
        raise _lib.DmlError(f"{model_name}: checkpoint has {len(keys)} arrays, the model needs {len(spec)}")
    out = []
    for k, (name, shape) in zip(keys, spec):
        a = np.asarray(z[k])
        if tuple(a.shape) != tuple(shape):
            raise _lib.DmlError(f"{model_name}: array {k!r} has shape {tuple(a.shape)}, {name} needs {tuple(shape)}")
        out.append(np.ascontiguousarray(a, dtype=np.float32))
    return out


_warned = set()


def load_blob(model_name: str) -> np.ndarray:
    """Blob from $DML_B200_WEIGHTS_DIR/<model>.npy (flat fp32) or .npz (arrays, see arrays_from_npz) if present,
    else the fixed-seed default preset -- with a warning, because that is NOT what the reference's
    ``weights='imagenet'`` (/root/reference/models.py:26, :51) predicts with."""
    d = os.environ.get("DML_B200_WEIGHTS_DIR")
    if d:
        npy = os.path.join(d, f"{model_name}.npy")
        npz = os.path.join(d, f"{model_name}.npz")
        if os.path.exists(npy):
            blob = np.ascontiguousarray(np.load(npy), dtype=np.float32).ravel()
            want = sum(int(np.prod(sh)) for _, sh in _lib.weight_spec(model_name))
            if blob.size != want:
                raise _lib.DmlError(f"{npy}: {blob.size} floats, {model_name} needs {want}")
            return blob
        if os.path.exists(npz):
            return to_blob(arrays_from_npz(model_name, np.load(npz)))
    if model_name not in _warned:
        _warned.add(model_name)
        warnings.warn(f"dml_b200: no {model_name} checkpoint under DML_B200_WEIGHTS_DIR -- serving FIXED-SEED RANDOM weights "
                      f"(preset {DEFAULT_PRESET[model_name]}); predictions are not ImageNet predictions", RuntimeWarning,
                      stacklevel=2)
    return to_blob(generate(model_name))
