"""Host side of load_img: JPEG decode with Pillow and the NEAREST index tables.

The reference decodes and resizes with Pillow through
``tf.keras.utils.load_img(path, target_size=(S, S))`` (/root/reference/models.py:34,
:59).  Here Pillow still decodes (same libjpeg), but the resize itself runs on
the GPU: the host only computes, per image, which source row/column each output
row/column reads.  Pillow's NEAREST walks a running fp64 sum (xo = a/2; xo += a)
and truncates, so the tables are built the same way.
"""
from __future__ import annotations

import functools
import io

import numpy as np
from PIL import Image


@functools.lru_cache(maxsize=4096)
def _table(n_in: int, n_out: int) -> np.ndarray:
    a = np.float64(n_in) / np.float64(n_out)
    # cumulative fp64 additions, exactly the order Pillow performs them
    steps = np.full(n_out, a, dtype=np.float64)
    steps[0] = a * np.float64(0.5)
    xo = np.add.accumulate(steps)  # sequential left-to-right accumulation
    tab = np.minimum(xo.astype(np.int64), n_in - 1).astype(np.int32)
    tab.setflags(write=False)
    return tab


def nearest_table(n_in: int, n_out: int) -> np.ndarray:
    """Source index for every output index of Pillow's NEAREST resize."""
    if n_in == n_out:
        return np.arange(n_out, dtype=np.int32)  # load_img skips the resize when sizes match
    return _table(int(n_in), int(n_out))


def decode_rgb(data: bytes) -> np.ndarray:
    """JPEG/PNG bytes -> HxWx3 uint8 RGB (convert('RGB') if needed, as load_img does)."""
    img = Image.open(io.BytesIO(data))
    if img.mode != "RGB":
        img = img.convert("RGB")
    return np.asarray(img, dtype=np.uint8)


def decode_file(path: str) -> np.ndarray:
    with open(path, "rb") as fh:
        return decode_rgb(fh.read())


def shrink_oversized(im: np.ndarray, size: int, max_bytes: int) -> np.ndarray:
    """An image whose pixels would not fit the engine's staging buffer is resized on the host through the very
    index tables the GPU kernel would use (rows first, then columns: a pure gather, so the result is the image
    load_img(target_size=(size, size)) returns).  Anything smaller goes to the GPU untouched."""
    if im.nbytes <= max_bytes:
        return im
    h, w = im.shape[:2]
    return np.ascontiguousarray(im[nearest_table(h, size)][:, nearest_table(w, size)])


class HostBatch:
    """Concatenated uint8 pixels + per-image offsets, sizes and index tables.

    ``into``: a dict of preallocated (ideally pinned) numpy arrays ``pixels`` / ``img_off`` / ``hw`` / ``xtab`` /
    ``ytab`` to fill instead of allocating (the engine's reusable host slots)."""

    def __init__(self, images, size: int, into: dict | None = None):
        n = len(images)
        self.n = n
        self.size = size
        for im in images:
            if im.ndim != 3 or im.shape[2] != 3 or im.dtype != np.uint8:
                raise ValueError("images must be HxWx3 uint8")
        sizes = np.fromiter((im.size for im in images), dtype=np.int64, count=n)
        total = int(sizes.sum())
        if into is None:
            offs = np.zeros(n, dtype=np.int64)
            hw = np.zeros((n, 2), dtype=np.int32)
            xt = np.zeros((n, size), dtype=np.int32)
            yt = np.zeros((n, size), dtype=np.int32)
            pix = np.empty(total + 16, dtype=np.uint8)
        else:
            if n > into["img_off"].shape[0] or total + 16 > into["pixels"].size:
                raise ValueError("batch does not fit the preallocated host buffers")
            offs, hw, xt, yt = into["img_off"][:n], into["hw"][:n], into["xtab"][:n], into["ytab"][:n]
            pix = into["pixels"][:total + 16]
        if n:
            offs[0] = 0
            np.cumsum(sizes[:-1], out=offs[1:])
            # one vectorised copy of all pixels (no per-image Python loop over the bytes)
            np.concatenate([im.reshape(-1) for im in images], out=pix[:total])
        ident = np.arange(size, dtype=np.int32)
        for i, im in enumerate(images):
            h, w = im.shape[:2]
            hw[i] = (h, w)
            # load_img resizes only when (W, H) != (S, S); both tables are identity otherwise
            xt[i] = _table(w, size) if w != size else ident
            yt[i] = _table(h, size) if h != size else ident
        self.pixels, self.img_off, self.hw, self.xtab, self.ytab = pix, offs, hw, xt, yt
