"""Oracle: the reference's per-file predict loop, restated on the CPU.

Mirrors /root/reference/models.py:23-46 (run_inference_on_InceptionV3) and
:48-71 (run_inference_on_ResNet50): for each file, load_img -> img_to_array ->
expand_dims -> preprocess_input -> predict on a batch of ONE -> top-5, keyed by
the file's basename.  This is also what bench.py times as the CPU baseline
("port": Keras itself is unavailable offline).  Test infrastructure only.
"""
from __future__ import annotations

import numpy as np

from . import decode, frontend, keras_nets


def run_inference(model: str, image_files, weights, class_index, batch: int = 1):
    """Reference loop (batch=1 is the reference's shape; >1 is the 'what if it batched' variant)."""
    results = {}
    files = list(image_files)
    for i in range(0, len(files), batch):
        chunk = files[i:i + batch]
        xs = []
        for f in chunk:
            with open(f, "rb") as fh:
                xs.append(frontend.load_and_preprocess(fh.read(), model))
        probs, _ = keras_nets.forward(model, weights, np.stack(xs))
        dec = decode.decode_predictions(probs, class_index, top=5)
        for f, d in zip(chunk, dec):
            results[f.split("/")[-1]] = [d]  # extra outer list = decode_predictions' batch dim
    return results


def predict_arrays(model: str, u8_images, weights, batch: int = 1):
    """Same loop on already-decoded uint8 HxWx3 images; returns (probs, logits)."""
    size, mode = frontend.MODEL_INPUT[model]
    ps, ls = [], []
    for i in range(0, len(u8_images), batch):
        xs = [frontend.preprocess(frontend.resize_nearest_tables(im, size), mode) for im in u8_images[i:i + batch]]
        p, l = keras_nets.forward(model, weights, np.stack(xs))
        ps.append(p)
        ls.append(l)
    return np.concatenate(ps), np.concatenate(ls)
