"""The fused convolution + 3x3 / stride-2 max-pool kernel (csrc/conv_pool.inc) against the convolution kernel followed by the pool.

InceptionV3's conv2d_2 (3x3 same, 32 -> 64) + max_pooling2d and ResNet50's conv1 + pool1_pad + pool1_pool are separate Keras layers
in /root/reference/models.py:40 / :65; the fused kernel keeps the convolution's rows in shared memory and writes only the pooled
rows.  max() of bf16 values is exact, so not one bit may differ from the separate kernels.
"""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def _case(n, hi, wi, pad, seed):
    import torch
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda")
    x = torch.from_numpy(rng.standard_normal((n, hi, wi, 32)).astype(np.float32)).to(dev).bfloat16()
    w = torch.from_numpy((rng.standard_normal((64, 9, 32)) / np.sqrt(288)).astype(np.float32)).to(dev).bfloat16()
    wp = torch.zeros((64, 9, 64), dtype=torch.bfloat16, device=dev)
    wp[:, :, :32] = w
    bias = torch.from_numpy(rng.standard_normal(64).astype(np.float32)).to(dev)
    return dict(n=n, hi=hi, wi=wi, pad=pad, x=x, wp=wp, bias=bias)


def _separate(c, pool_pad):
    import torch
    import torch.nn.functional as F
    lib = pkg("_lib")
    n, hi, wi, pad = c["n"], c["hi"], c["wi"], c["pad"]
    ho, wo = hi + 2 * pad - 2, wi + 2 * pad - 2
    conv = torch.empty((n, ho, wo, 64), dtype=torch.bfloat16, device="cuda")
    lib.check(lib.lib().dml_conv2d_bf16(7, c["x"].data_ptr(), c["wp"].data_ptr(), c["bias"].data_ptr(), None, conv.data_ptr(), n, hi, wi, 32, 32,
                                        64, 64, 0, 3, 3, 1, pad, pad, 64, 1, 0, 0, None))
    torch.cuda.synchronize()
    y = conv.float().permute(0, 3, 1, 2)
    if pool_pad:
        y = F.pad(y, (1, 1, 1, 1))  # ZeroPadding2D(1): zeros, not -inf
    return F.max_pool2d(y, 3, 2).permute(0, 2, 3, 1).bfloat16()  # exact: a maximum of bf16 values


def _fused(c, pool_pad):
    import torch
    lib = pkg("_lib")
    n, hi, wi, pad = c["n"], c["hi"], c["wi"], c["pad"]
    ho, wo = hi + 2 * pad - 2, wi + 2 * pad - 2
    hp, wq = (ho + 2 * pool_pad - 3) // 2 + 1, (wo + 2 * pool_pad - 3) // 2 + 1
    out = torch.full((n, hp, wq, 64), 7.0, dtype=torch.bfloat16, device="cuda")
    lib.check(lib.lib().dml_conv_pool_bf16(c["x"].data_ptr(), c["wp"].data_ptr(), c["bias"].data_ptr(), out.data_ptr(), n, hi, wi, 32, 64, 3, 3,
                                           pad, 1, pool_pad, None))
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("n,hi,wi,pad,pool_pad", [(1, 147, 147, 1, 0), (3, 147, 147, 1, 0), (130, 31, 31, 1, 0), (2, 112, 112, 1, 1),
                                                  (5, 40, 56, 0, 1), (1, 9, 9, 1, 0), (160, 24, 24, 1, 1)])
def test_conv_pool_matches_conv_then_pool_bit_for_bit(n, hi, wi, pad, pool_pad):
    import torch
    c = _case(n, hi, wi, pad, seed=n + hi)
    got = _fused(c, pool_pad)
    ref = _separate(c, pool_pad)
    assert got.shape == ref.shape
    assert torch.equal(got, ref)


def test_conv_pool_refuses_what_it_cannot_do():
    import torch
    lib = pkg("_lib")
    c = _case(1, 20, 20, 1, seed=1)
    out = torch.empty((1, 9, 9, 64), dtype=torch.bfloat16, device="cuda")
    # the pool's zero padding is only neutral behind a ReLU
    rc = lib.lib().dml_conv_pool_bf16(c["x"].data_ptr(), c["wp"].data_ptr(), c["bias"].data_ptr(), out.data_ptr(), 1, 20, 20, 32, 64, 3, 3, 1, 0, 1, None)
    assert rc == -1 and b"ReLU" in lib.lib().dml_last_error()


@pytest.mark.parametrize("model,first", [("ResNet50", 0), ("InceptionV3", 100)])
def test_engine_with_and_without_the_fused_pool_gives_the_same_bits(images, monkeypatch, model, first):
    """Whole networks through the engine: the stem + pool1 (ResNet50, space-to-depth stem with the pool's zero padding) and
    conv2d_2 + max_pooling2d (InceptionV3) fused (forced, whatever the timing would pick) against DML_NO_CONV_POOL."""
    kw = pkg("keras_weights")
    W = kw.generate(model, "folding_stress" if model == "ResNet50" else "he_uniform", 9)
    sel = images[first:first + 37]
    outs = []
    for fused in (True, False):
        monkeypatch.delenv("DML_NO_CONV_POOL", raising=False)
        monkeypatch.delenv("DML_FORCE_CONV_POOL", raising=False)
        monkeypatch.setenv("DML_FORCE_CONV_POOL" if fused else "DML_NO_CONV_POOL", "1")
        eng = pkg("engine").Engine(model, max_batch=40, weights=W)
        try:
            pools = [o for o in eng.ops() if o[1] == 1 and o[4] == 0]  # max-pools that report no bytes of their own: done by their conv
            assert (len(pools) == 1) == fused, pools
            idx, prob, probs = eng.predict_images(sel, want_probs=True)
            outs.append((idx, prob, probs, eng.logits(len(sel)).cpu().numpy()))
        finally:
            eng.close()
    for a, b in zip(outs[0], outs[1]):
        assert np.array_equal(a, b)
