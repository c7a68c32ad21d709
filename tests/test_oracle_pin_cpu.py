"""Independent cross-check of the oracle's network arithmetic (CPU).

TensorFlow/Keras cannot be installed offline, so the oracle's Keras-graph restatement (oracle/keras_nets.py,
following /root/reference/models.py:26 and :51) has no reference run to be pinned against.  The one independent
implementation of the same two architectures on this box is torchvision 0.26:

* ``inception_v3(aux_logits=False, transform_input=False)`` has Keras InceptionV3's topology and BN eps; it differs in
  two places that are bridged explicitly here: its BatchNorm has a gamma (set to 1) and its branch average pools count
  the padding (patched to ``count_include_pad=False``; that divisor rule is covered by its own closed-form test in
  test_oracle_cpu.py).
* ``resnet50`` is v1.5 (stride on the 3x3) and bias-free; moving the stride of the first block of layers 2-4 to the
  first 1x1 and folding each conv bias into the following BatchNorm's running mean gives Keras' v1.

The same Keras-layout weight list goes into both; logits must agree to fp32 round-off.  What this pins: the layer
creation order <-> graph wiring (concat order, branch order, which array belongs to which conv), kernel layout
(HWIO, (rows, cols) of the 1x7 / 7x1 filters), paddings, stride placement, eps, the BN formula and the dense layer.
"""
import numpy as np
import pytest
import torch

from oracle import frontend as ofe, keras_nets, weights as oweights

tv = pytest.importorskip("torchvision")


def _randomise_bn(model, W, seed):
    """Non-trivial BN statistics / biases on top of a kernel preset, so that a swapped array cannot go unnoticed."""
    rng = np.random.default_rng(seed)
    out = []
    for (name, shape), a in zip(keras_nets.weight_spec(model), W):
        kind = name.rsplit("/", 1)[1]
        if kind in ("bias", "beta", "moving_mean"):
            a = rng.normal(0.0, 0.1, size=shape).astype(np.float32)
        elif kind in ("gamma", "moving_variance"):
            a = rng.uniform(0.5, 1.5, size=shape).astype(np.float32)
        out.append(a)
    return out


def _set_conv(conv, kernel_hwio):
    w = torch.from_numpy(np.ascontiguousarray(kernel_hwio.transpose(3, 2, 0, 1)))
    assert tuple(conv.weight.shape) == tuple(w.shape), (conv, w.shape)
    conv.weight.data.copy_(w)


def _set_bn(bn, gamma, beta, mean, var, eps, conv_bias=None):
    bn.eps = eps
    bn.weight.data.copy_(torch.from_numpy(gamma) if gamma is not None else torch.ones_like(bn.weight))
    bn.bias.data.copy_(torch.from_numpy(beta))
    m = mean if conv_bias is None else mean - conv_bias  # BN(conv + b) == BN'(conv) with mean' = mean - b
    bn.running_mean.data.copy_(torch.from_numpy(m))
    bn.running_var.data.copy_(torch.from_numpy(var))


def _inputs(model, images):
    size, mode = ofe.MODEL_INPUT[model]
    return np.stack([ofe.preprocess(ofe.resize_nearest_tables(im, size), mode) for im in images])


def test_inception_v3_matches_torchvision(images, monkeypatch):
    model = "InceptionV3"
    W = _randomise_bn(model, oweights.generate(model, "he_uniform", 3), 11)
    net = tv.models.inception_v3(weights=None, aux_logits=False, transform_input=False, init_weights=False).eval()
    # Keras layer creation order (conv2d, conv2d_1, ... conv2d_93) -> torchvision BasicConv2d modules
    order = ["Conv2d_1a_3x3", "Conv2d_2a_3x3", "Conv2d_2b_3x3", "Conv2d_3b_1x1", "Conv2d_4a_3x3"]
    for blk in ("Mixed_5b", "Mixed_5c", "Mixed_5d"):
        order += [f"{blk}.{b}" for b in ("branch1x1", "branch5x5_1", "branch5x5_2", "branch3x3dbl_1", "branch3x3dbl_2",
                                         "branch3x3dbl_3", "branch_pool")]
    order += [f"Mixed_6a.{b}" for b in ("branch3x3", "branch3x3dbl_1", "branch3x3dbl_2", "branch3x3dbl_3")]
    for blk in ("Mixed_6b", "Mixed_6c", "Mixed_6d", "Mixed_6e"):
        order += [f"{blk}.{b}" for b in ("branch1x1", "branch7x7_1", "branch7x7_2", "branch7x7_3", "branch7x7dbl_1",
                                         "branch7x7dbl_2", "branch7x7dbl_3", "branch7x7dbl_4", "branch7x7dbl_5", "branch_pool")]
    order += [f"Mixed_7a.{b}" for b in ("branch3x3_1", "branch3x3_2", "branch7x7x3_1", "branch7x7x3_2", "branch7x7x3_3",
                                        "branch7x7x3_4")]
    for blk in ("Mixed_7b", "Mixed_7c"):
        order += [f"{blk}.{b}" for b in ("branch1x1", "branch3x3_1", "branch3x3_2a", "branch3x3_2b", "branch3x3dbl_1",
                                         "branch3x3dbl_2", "branch3x3dbl_3a", "branch3x3dbl_3b", "branch_pool")]
    assert len(order) == 94
    it = iter(W)
    for path in order:
        mod = net.get_submodule(path)
        _set_conv(mod.conv, next(it))
        _set_bn(mod.bn, None, next(it), next(it), next(it), keras_nets.INCEPTION_BN_EPS)
    net.fc.weight.data.copy_(torch.from_numpy(np.ascontiguousarray(next(it).T)))
    net.fc.bias.data.copy_(torch.from_numpy(next(it)))
    assert next(it, None) is None
    # torchvision's branch pools are F.avg_pool2d(x, 3, 1, 1) with the padding counted; Keras divides by the in-bounds taps
    real = torch.nn.functional.avg_pool2d

    def avg_pool_excluding_pad(x, *a, **k):
        k.setdefault("count_include_pad", False)
        return real(x, *a, **k)

    monkeypatch.setattr(torch.nn.functional, "avg_pool2d", avg_pool_excluding_pad)
    x = _inputs(model, [images[101], images[150], images[199]])
    _, ref = keras_nets.forward(model, W, x)
    with torch.no_grad():
        got = net(torch.from_numpy(x).permute(0, 3, 1, 2)).numpy()
    rel = np.abs(got - ref).max() / np.abs(ref).max()
    print(f"\nInceptionV3 oracle vs torchvision: logits max|d|/max|ref| = {rel:.2e} (max|ref| {np.abs(ref).max():.3f})")
    assert np.abs(ref).max() > 1e-2  # not a vacuous comparison
    assert rel <= 2e-5


def test_resnet50_v1_matches_torchvision_blocks(images):
    model = "ResNet50"
    W = oweights.generate(model, "folding_stress", 5)
    arr = {name: a for (name, _), a in zip(keras_nets.weight_spec(model), W)}
    eps = keras_nets.RESNET_BN_EPS
    net = tv.models.resnet50(weights=None).eval()
    for layer in (net.layer2, net.layer3, net.layer4):  # v1.5 -> v1: the stride sits on the first 1x1 of the stage
        assert layer[0].conv2.stride == (2, 2)
        layer[0].conv1.stride, layer[0].conv2.stride = (2, 2), (1, 1)

    def load(conv, bn, cname, bname):
        _set_conv(conv, arr[cname + "/kernel"])
        _set_bn(bn, arr[bname + "/gamma"], arr[bname + "/beta"], arr[bname + "/moving_mean"], arr[bname + "/moving_variance"],
                eps, conv_bias=arr[cname + "/bias"])

    load(net.conv1, net.bn1, "conv1_conv", "conv1_bn")
    used = 6
    for s, nblocks in zip((2, 3, 4, 5), (3, 4, 6, 3)):
        layer = getattr(net, f"layer{s - 1}")
        assert len(layer) == nblocks
        for b in range(1, nblocks + 1):
            n, blk = f"conv{s}_block{b}", layer[b - 1]
            if b == 1:
                load(blk.downsample[0], blk.downsample[1], n + "_0_conv", n + "_0_bn")
                used += 6
            load(blk.conv1, blk.bn1, n + "_1_conv", n + "_1_bn")
            load(blk.conv2, blk.bn2, n + "_2_conv", n + "_2_bn")
            load(blk.conv3, blk.bn3, n + "_3_conv", n + "_3_bn")
            used += 18
    net.fc.weight.data.copy_(torch.from_numpy(np.ascontiguousarray(arr["predictions/kernel"].T)))
    net.fc.bias.data.copy_(torch.from_numpy(arr["predictions/bias"]))
    assert used + 2 == len(W) == 320
    x = _inputs(model, [images[0], images[57], images[99]])
    _, ref = keras_nets.forward(model, W, x)
    with torch.no_grad():
        got = net(torch.from_numpy(x).permute(0, 3, 1, 2)).numpy()
    rel = np.abs(got - ref).max() / np.abs(ref).max()
    print(f"\nResNet50-v1 oracle vs torchvision blocks: logits max|d|/max|ref| = {rel:.2e} (max|ref| {np.abs(ref).max():.3f})")
    assert np.abs(ref).max() > 1e-2
    assert rel <= 2e-5


def test_oracle_weight_generator_equals_the_engines():
    """bench.py's reference arm draws its weights from the oracle's own generator (it must not load the product
    library); both generators give the same arrays."""
    import importlib
    kw = importlib.import_module("distributed-machine-learning_b200.keras_weights")
    for model, preset, seed in (("ResNet50", None, None), ("InceptionV3", None, None), ("ResNet50", "folding_stress", 5),
                                ("InceptionV3", "keras_default", 1)):
        a, b = oweights.generate(model, preset, seed), kw.generate(model, preset, seed)
        assert len(a) == len(b) and all(np.array_equal(x, y) for x, y in zip(a, b))
