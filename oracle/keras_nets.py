"""Oracle: Keras ResNet50 / InceptionV3 graphs restated with torch CPU fp32 ops.

What the reference runs at /root/reference/models.py:40 and :65
(``modelObj.predict(input_img)``) is keras.applications' ResNet50 (v1: stride on
the first 1x1, conv biases, BN eps 1.001e-5, explicit ZeroPadding) and
InceptionV3 (bias-free convs, BN scale=False eps 1e-3, avg-pool 'same' dividing
by the number of in-bounds taps).  TensorFlow/Keras 2.11 (un-vendored, absent
here) is restated from its published application definitions; parity unpinned,
see oracle/__init__.py.  Test infrastructure only.

Weights are a list of fp32 numpy arrays in layer CREATION order, each layer in
Keras ``get_weights()`` order: Conv2D [kernel HWIO (, bias)], BatchNormalization
[(gamma,) beta, moving_mean, moving_variance], Dense [kernel (in,out), bias].
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

RESNET_BN_EPS = 1.001e-5
INCEPTION_BN_EPS = 1e-3


class _Weights:
    """Hands out weight arrays in creation order; in spec mode records shapes."""

    def __init__(self, arrays=None):
        self.arrays = arrays
        self.i = 0
        self.spec = []  # (name, shape)

    def take(self, name, shape):
        shape = tuple(int(s) for s in shape)
        if self.arrays is None:
            self.spec.append((name, shape))
            return torch.empty(shape, device="meta")
        a = self.arrays[self.i]
        if tuple(a.shape) != shape:
            raise ValueError(f"weight {self.i} ({name}): got {a.shape}, want {shape}")
        self.i += 1
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))


class _Ctx:
    def __init__(self, weights: _Weights, taps):
        self.w = weights
        self.taps = taps  # dict or None
        self.nconv = 0
        self.nbn = 0

    def tap(self, name, x):
        if self.taps is not None and x.device.type != "meta":
            self.taps[name] = x.permute(0, 2, 3, 1).contiguous().numpy()


def _conv(ctx, x, cout, kh, kw, stride, padding, bias, name):
    cin = x.shape[1]
    k = ctx.w.take(name + "/kernel", (kh, kw, cin, cout))  # HWIO
    b = ctx.w.take(name + "/bias", (cout,)) if bias else None
    if padding == "same":
        assert stride == 1
        pad = ((kh - 1) // 2, (kw - 1) // 2)
    else:
        pad = (0, 0)
    # TF conv is cross-correlation; torch wants OIHW.
    return F.conv2d(x, k.permute(3, 2, 0, 1), b, stride=stride, padding=pad)


def _bn(ctx, x, scale, eps, name):
    c = x.shape[1]
    gamma = ctx.w.take(name + "/gamma", (c,)) if scale else None
    beta = ctx.w.take(name + "/beta", (c,))
    mean = ctx.w.take(name + "/moving_mean", (c,))
    var = ctx.w.take(name + "/moving_variance", (c,))
    return F.batch_norm(x, mean, var, gamma, beta, training=False, eps=eps)


def _head(ctx, x, name_pool="avg_pool"):
    x = x.mean(dim=(2, 3))  # GlobalAveragePooling2D
    if ctx.taps is not None and x.device.type != "meta":
        ctx.taps[name_pool] = x.numpy()
    k = ctx.w.take("predictions/kernel", (x.shape[1], 1000))
    b = ctx.w.take("predictions/bias", (1000,))
    logits = x @ k + b
    return logits


# ----------------------------------------------------------------------------
# ResNet50 (keras.applications.resnet.ResNet50)
# ----------------------------------------------------------------------------
def _resnet_block(ctx, x, f, stride, conv_shortcut, name):
    eps = RESNET_BN_EPS
    if conv_shortcut:
        sc = _conv(ctx, x, 4 * f, 1, 1, stride, "valid", True, name + "_0_conv")
        sc = _bn(ctx, sc, True, eps, name + "_0_bn")
    else:
        sc = x
    y = _conv(ctx, x, f, 1, 1, stride, "valid", True, name + "_1_conv")
    y = F.relu(_bn(ctx, y, True, eps, name + "_1_bn"))
    y = _conv(ctx, y, f, 3, 3, 1, "same", True, name + "_2_conv")
    y = F.relu(_bn(ctx, y, True, eps, name + "_2_bn"))
    y = _conv(ctx, y, 4 * f, 1, 1, 1, "valid", True, name + "_3_conv")
    y = _bn(ctx, y, True, eps, name + "_3_bn")
    out = F.relu(sc + y)
    ctx.tap(name + "_out", out)
    return out


def _resnet50(ctx, x):
    x = F.pad(x, (3, 3, 3, 3))  # conv1_pad
    x = _conv(ctx, x, 64, 7, 7, 2, "valid", True, "conv1_conv")
    x = F.relu(_bn(ctx, x, True, RESNET_BN_EPS, "conv1_bn"))
    ctx.tap("conv1_relu", x)
    x = F.pad(x, (1, 1, 1, 1))  # pool1_pad (zeros; inputs are >= 0)
    x = F.max_pool2d(x, 3, 2)
    ctx.tap("pool1_pool", x)
    for stage, (f, blocks, stride1) in enumerate([(64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)], start=2):
        x = _resnet_block(ctx, x, f, stride1, True, f"conv{stage}_block1")
        for b in range(2, blocks + 1):
            x = _resnet_block(ctx, x, f, 1, False, f"conv{stage}_block{b}")
    return _head(ctx, x)


# ----------------------------------------------------------------------------
# InceptionV3 (keras.applications.inception_v3.InceptionV3)
# ----------------------------------------------------------------------------
def _cb(ctx, x, f, kh, kw, padding="same", stride=1):
    """conv2d_bn: Conv2D(no bias) -> BN(scale=False, eps=1e-3) -> ReLU."""
    i = ctx.nconv
    ctx.nconv += 1
    cname = "conv2d" if i == 0 else f"conv2d_{i}"
    bname = "batch_normalization" if i == 0 else f"batch_normalization_{i}"
    x = _conv(ctx, x, f, kh, kw, stride, padding, False, cname)
    x = _bn(ctx, x, False, INCEPTION_BN_EPS, bname)
    return F.relu(x)


def _avgpool_same(x):
    # TF AveragePooling2D(3, strides 1, 'same') divides by the in-bounds tap count.
    return F.avg_pool2d(x, 3, 1, padding=1, count_include_pad=False)


def _inception_v3(ctx, x):
    x = _cb(ctx, x, 32, 3, 3, "valid", 2)
    ctx.tap("stem_conv0", x)
    x = _cb(ctx, x, 32, 3, 3, "valid")
    x = _cb(ctx, x, 64, 3, 3)
    x = F.max_pool2d(x, 3, 2)
    x = _cb(ctx, x, 80, 1, 1, "valid")
    x = _cb(ctx, x, 192, 3, 3, "valid")
    x = F.max_pool2d(x, 3, 2)
    ctx.tap("stem_pool2", x)

    # mixed0..2 (35x35)
    for i, pool_f in enumerate([32, 64, 64]):
        b1 = _cb(ctx, x, 64, 1, 1)
        b5 = _cb(ctx, x, 48, 1, 1)
        b5 = _cb(ctx, b5, 64, 5, 5)
        b3 = _cb(ctx, x, 64, 1, 1)
        b3 = _cb(ctx, b3, 96, 3, 3)
        b3 = _cb(ctx, b3, 96, 3, 3)
        bp = _avgpool_same(x)
        bp = _cb(ctx, bp, pool_f, 1, 1)
        x = torch.cat([b1, b5, b3, bp], dim=1)
        ctx.tap(f"mixed{i}", x)

    # mixed3 (17x17)
    b3 = _cb(ctx, x, 384, 3, 3, "valid", 2)
    bd = _cb(ctx, x, 64, 1, 1)
    bd = _cb(ctx, bd, 96, 3, 3)
    bd = _cb(ctx, bd, 96, 3, 3, "valid", 2)
    bp = F.max_pool2d(x, 3, 2)
    x = torch.cat([b3, bd, bp], dim=1)
    ctx.tap("mixed3", x)

    # mixed4..7 (17x17)
    for i, c in enumerate([128, 160, 160, 192], start=4):
        b1 = _cb(ctx, x, 192, 1, 1)
        b7 = _cb(ctx, x, c, 1, 1)
        b7 = _cb(ctx, b7, c, 1, 7)
        b7 = _cb(ctx, b7, 192, 7, 1)
        bd = _cb(ctx, x, c, 1, 1)
        bd = _cb(ctx, bd, c, 7, 1)
        bd = _cb(ctx, bd, c, 1, 7)
        bd = _cb(ctx, bd, c, 7, 1)
        bd = _cb(ctx, bd, 192, 1, 7)
        bp = _avgpool_same(x)
        bp = _cb(ctx, bp, 192, 1, 1)
        x = torch.cat([b1, b7, bd, bp], dim=1)
        ctx.tap(f"mixed{i}", x)

    # mixed8 (8x8)
    b3 = _cb(ctx, x, 192, 1, 1)
    b3 = _cb(ctx, b3, 320, 3, 3, "valid", 2)
    b7 = _cb(ctx, x, 192, 1, 1)
    b7 = _cb(ctx, b7, 192, 1, 7)
    b7 = _cb(ctx, b7, 192, 7, 1)
    b7 = _cb(ctx, b7, 192, 3, 3, "valid", 2)
    bp = F.max_pool2d(x, 3, 2)
    x = torch.cat([b3, b7, bp], dim=1)
    ctx.tap("mixed8", x)

    # mixed9, mixed10 (8x8)
    for i in (9, 10):
        b1 = _cb(ctx, x, 320, 1, 1)
        b3 = _cb(ctx, x, 384, 1, 1)
        b3a = _cb(ctx, b3, 384, 1, 3)
        b3b = _cb(ctx, b3, 384, 3, 1)
        b3 = torch.cat([b3a, b3b], dim=1)
        bd = _cb(ctx, x, 448, 1, 1)
        bd = _cb(ctx, bd, 384, 3, 3)
        bda = _cb(ctx, bd, 384, 1, 3)
        bdb = _cb(ctx, bd, 384, 3, 1)
        bd = torch.cat([bda, bdb], dim=1)
        bp = _avgpool_same(x)
        bp = _cb(ctx, bp, 192, 1, 1)
        x = torch.cat([b1, b3, bd, bp], dim=1)
        ctx.tap(f"mixed{i}", x)
    return _head(ctx, x)


_NETS = {"ResNet50": (_resnet50, 224), "InceptionV3": (_inception_v3, 299)}


def weight_spec(model: str):
    """[(name, shape)] in blob order (layer creation order)."""
    fn, size = _NETS[model]
    w = _Weights(None)
    fn(_Ctx(w, None), torch.empty((1, 3, size, size), device="meta"))
    return w.spec


def forward(model: str, weights, x_nhwc: np.ndarray, taps: dict | None = None):
    """fp32 NHWC preprocessed batch -> (softmax probs [N,1000], logits [N,1000])."""
    fn, _ = _NETS[model]
    w = _Weights(list(weights))
    x = torch.from_numpy(np.ascontiguousarray(x_nhwc, dtype=np.float32)).permute(0, 3, 1, 2)
    with torch.no_grad():
        logits = fn(_Ctx(w, taps), x)
        assert w.i == len(w.arrays), "unused weights"
        probs = torch.softmax(logits, dim=1)
    return probs.numpy(), logits.numpy()


def blob_to_arrays(model: str, blob: np.ndarray):
    """Split a flat fp32 Keras-layout blob into the per-variable arrays."""
    out, off = [], 0
    for _, shape in weight_spec(model):
        n = int(np.prod(shape))
        out.append(blob[off:off + n].reshape(shape))
        off += n
    if off != blob.size:
        raise ValueError(f"blob has {blob.size} floats, {model} needs {off}")
    return out
