"""Oracle front end: load_img / img_to_array / preprocess_input restated.

Follows /root/reference/models.py:34-38 (InceptionV3, 299x299, tf mode) and
/root/reference/models.py:59-63 (ResNet50, 224x224, caffe mode).  Keras itself
is absent (see oracle/__init__.py); ``load_img`` is restated from its published
behaviour: PIL open -> convert("RGB") if needed -> resize((W, H), NEAREST) when
the size differs.  Test infrastructure only.
"""
from __future__ import annotations

import io

import numpy as np
from PIL import Image

# keras.applications.imagenet_utils: caffe-mode BGR means, as fp32.
CAFFE_MEAN_BGR = np.array([103.939, 116.779, 123.68], dtype=np.float32)

MODEL_INPUT = {"ResNet50": (224, "caffe"), "InceptionV3": (299, "tf")}


def pillow_nearest_table(n_in: int, n_out: int) -> np.ndarray:
    """Source index for each output index of Pillow's NEAREST resize.

    Pillow's ImagingScaleAffine keeps a running fp64 sum (xo = a/2; xo += a) and
    truncates; closed forms differ on exact-integer landings.
    """
    a = np.float64(n_in) / np.float64(n_out)
    xo = a * np.float64(0.5)
    tab = np.empty(n_out, dtype=np.int32)
    for x in range(n_out):
        tab[x] = int(xo)
        xo = xo + a
    return np.minimum(tab, n_in - 1)


def decode_rgb(data: bytes) -> np.ndarray:
    """JPEG bytes -> HxWx3 uint8 (PIL decode, convert to RGB if needed)."""
    img = Image.open(io.BytesIO(data))
    if img.mode != "RGB":
        img = img.convert("RGB")
    return np.asarray(img, dtype=np.uint8)


def load_img(data: bytes, size: int) -> np.ndarray:
    """tf.keras.utils.load_img(path, target_size=(size,size)) -> uint8 HxWx3.

    Uses Pillow's own resize, exactly as Keras does (models.py:34, :59).
    """
    img = Image.open(io.BytesIO(data))
    if img.mode != "RGB":
        img = img.convert("RGB")
    if img.size != (size, size):
        img = img.resize((size, size), Image.NEAREST)
    return np.asarray(img, dtype=np.uint8)


def resize_nearest_tables(rgb: np.ndarray, size: int) -> np.ndarray:
    """Same result as load_img's resize, computed from the index tables."""
    h, w = rgb.shape[:2]
    if (w, h) == (size, size):
        return rgb
    yt = pillow_nearest_table(h, size)
    xt = pillow_nearest_table(w, size)
    return rgb[yt][:, xt]


def preprocess(u8: np.ndarray, mode: str) -> np.ndarray:
    """img_to_array + preprocess_input: uint8 [...,H,W,3] RGB -> fp32 NHWC.

    caffe (models.py:63): RGB->BGR then subtract the fp32 mean per channel.
    tf    (models.py:38): x /= 127.5 ; x -= 1.0, two rounded fp32 operations.
    """
    x = u8.astype(np.float32)
    if mode == "caffe":
        x = x[..., ::-1] - CAFFE_MEAN_BGR
        return np.ascontiguousarray(x, dtype=np.float32)
    if mode == "tf":
        x = x / np.float32(127.5)
        x = x - np.float32(1.0)
        return x.astype(np.float32)
    raise ValueError(mode)


def load_and_preprocess(data: bytes, model: str) -> np.ndarray:
    size, mode = MODEL_INPUT[model]
    return preprocess(load_img(data, size), mode)
