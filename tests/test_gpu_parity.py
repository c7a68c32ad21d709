"""End-to-end parity on the 200 bundled images, through the C ABI, against the
committed oracle vectors (tests/golden/*_oracle.npz):

  fp32 check mode : logits within 1e-3 relative (max|d| / max|ref|)
  bf16 modes      : softmax within 2e-2 max-abs (BASELINE.json tolerance); logits
                    reported relative; identical top-1 on all 200 images
and per-layer taps against the oracle run live on a few images."""
import numpy as np
import pytest

from conftest import pkg
from oracle import keras_nets, frontend as ofe

pytestmark = pytest.mark.gpu

MODES = {
    "fp32_check": dict(fp32_check=True),
    "bf16_cuda_core": dict(tensor_cores=False),
    "bf16_tcgen05": dict(),
}


def softmax(l):
    e = np.exp(l - l.max(1, keepdims=True))
    return e / e.sum(1, keepdims=True)


@pytest.mark.parametrize("mode", list(MODES))
@pytest.mark.parametrize("model", ["ResNet50", "InceptionV3"])
def test_all_200_images(model, mode, images, oracle_golden):
    g = oracle_golden[model]
    kw = pkg("keras_weights")
    W = kw.generate(model, str(g["preset"]), int(g["seed"]))
    eng = pkg("engine").Engine(model, max_batch=200, micro_batch=32, weights=W, **MODES[mode])
    try:
        idx, prob, probs = eng.predict_images(images, want_probs=True)
        logits = eng.logits(len(images)).cpu().numpy()
    finally:
        eng.close()
    ref_logits = g["logits"]
    ref_probs = softmax(ref_logits)
    rel = np.abs(logits - ref_logits).max() / np.abs(ref_logits).max()
    dsm = np.abs(probs - ref_probs).max()
    top1_same = int((idx[:, 0] == g["top5_idx"][:, 0]).sum())
    top5_same = int(sum(set(a) == set(b) for a, b in zip(idx, g["top5_idx"])))
    print(f"\n{model} {mode}: logits rel {rel:.3e}  softmax max-abs {dsm:.3e}  top1 {top1_same}/200  top5-set {top5_same}/200")
    assert np.isfinite(logits).all()
    if mode == "fp32_check":
        assert rel <= 1e-3
        assert dsm <= 1e-3
    else:
        assert dsm <= 2e-2
        assert rel <= 3e-2
    assert top1_same == 200
    # the engine's top-5 is consistent with its own softmax, ties -> larger index
    for r in range(0, 200, 17):
        order = np.argsort(probs[r], kind="stable")[-5:][::-1]
        assert list(order) == list(idx[r])
        assert np.array_equal(prob[r], probs[r][order])


@pytest.mark.parametrize("mode", ["fp32_check", "bf16_tcgen05"])
@pytest.mark.parametrize("model", ["ResNet50", "InceptionV3"])
def test_layer_taps(model, mode, images):
    """Intermediate tensors against the oracle, relative to each tensor's own scale."""
    kw = pkg("keras_weights")
    W = kw.generate(model)
    sel = [images[3], images[150], images[77]]
    size, pmode = ofe.MODEL_INPUT[model]
    x = np.stack([ofe.preprocess(ofe.resize_nearest_tables(im, size), pmode) for im in sel])
    taps = {}
    keras_nets.forward(model, W, x, taps)
    eng = pkg("engine").Engine(model, max_batch=4, micro_batch=4, weights=W, keep_all=True, **MODES[mode])
    try:
        eng.predict_images(sel)
        have = eng.tensors()
        checked = 0
        worst = 0.0
        for name, ref in taps.items():
            if name not in have or ref.ndim != 4:
                continue
            got = eng.read_tensor(name, len(sel))
            assert got.shape == ref.shape, name
            rel = np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-20)
            worst = max(worst, rel)
            assert rel <= (1e-3 if mode == "fp32_check" else 4e-2), (name, rel)
            checked += 1
        print(f"\n{model} {mode}: {checked} taps, worst rel {worst:.3e}")
        assert checked >= 10
    finally:
        eng.close()


def test_results_do_not_depend_on_micro_batch(images):
    kw = pkg("keras_weights")
    W = kw.generate("InceptionV3")
    outs = []
    for mb in (1, 5, 16):
        eng = pkg("engine").Engine("InceptionV3", max_batch=16, micro_batch=mb, weights=W)
        try:
            outs.append(eng.predict_images(images[100:113], want_probs=True)[2])
        finally:
            eng.close()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])


def test_folded_projection_shortcut_matches_separate_convs(images, monkeypatch):
    """ResNet50's projection shortcuts run as a second GEMM source of the block's last conv (graph.h, Op::in2).
    Against the same engine with the shortcut as its own conv + residual add (DML_NO_FUSE_SHORTCUT), on weights with
    non-trivial biases / BN statistics: the only difference is that the shortcut is no longer rounded to bf16."""
    from oracle import models_ref
    kw = pkg("keras_weights")
    W = kw.generate("ResNet50", "folding_stress", 5)
    sel = images[20:45]   # 25 images: partial tiles in conv4 / conv5
    outs = []
    for fuse in (True, False):
        if fuse:
            monkeypatch.delenv("DML_NO_FUSE_SHORTCUT", raising=False)
        else:
            monkeypatch.setenv("DML_NO_FUSE_SHORTCUT", "1")
        eng = pkg("engine").Engine("ResNet50", max_batch=32, weights=W)
        try:
            names = [o[0] for o in eng.ops()]
            assert ("conv2_block1_0_conv" in names) == (not fuse)
            eng.predict_images(sel)
            outs.append(eng.logits(len(sel)).cpu().numpy().copy())
        finally:
            eng.close()
    _, ref = models_ref.predict_arrays("ResNet50", sel, W, batch=25)
    scale = np.abs(ref).max()
    err_f, err_s = np.abs(outs[0] - ref).max() / scale, np.abs(outs[1] - ref).max() / scale
    print(f"\nlogits vs fp32 oracle: folded {err_f:.3e}, separate {err_s:.3e}; folded vs separate {np.abs(outs[0] - outs[1]).max() / scale:.3e}")
    assert np.abs(outs[0] - outs[1]).max() / scale <= 2e-2
    assert err_f <= max(2e-2, 1.5 * err_s)
    assert np.array_equal(outs[0].argmax(1), outs[1].argmax(1))


def test_merged_sibling_reductions_are_bit_identical(images, monkeypatch):
    """InceptionV3's two sibling 1x1 branch reductions of a block run as one GEMM (graph.h, Op::merge_with): every output
    channel sees the same K loop, so the logits must not change by a single bit against the separate convolutions."""
    kw = pkg("keras_weights")
    W = kw.generate("InceptionV3", "folding_stress", 11)
    sel = images[50:69]
    outs = []
    monkeypatch.setenv("DML_NO_POOL_COMMUTE", "1")  # the pool-branch commute (next test) changes the rounding order, this merge does not
    for fuse in (True, False):
        if fuse:
            monkeypatch.delenv("DML_NO_FUSE_SHORTCUT", raising=False)
        else:
            monkeypatch.setenv("DML_NO_FUSE_SHORTCUT", "1")
        eng = pkg("engine").Engine("InceptionV3", max_batch=24, weights=W)
        try:
            followers = [o[0] for o in eng.ops() if o[1] == 0 and o[3] == 0]
            assert (len(followers) == 19) == fuse, followers   # 9 block heads x 2 reductions + mixed8's pair
            eng.predict_images(sel)
            outs.append(eng.logits(len(sel)).cpu().numpy().copy())
        finally:
            eng.close()
    assert np.array_equal(outs[0], outs[1])


def test_pool_branch_conv_before_pool_matches_pool_before_conv(images, monkeypatch):
    """InceptionV3's pool branches are AveragePooling2D -> 1x1 conv -> BN -> ReLU (Keras conv2d_bn after the pool).  The engine
    runs the 1x1 conv first, as one more member of the block's merged head GEMM, and averages its few output channels
    afterwards (bias and ReLU after the average): the same linear map, with the bf16 rounding at a different place.  Against
    the pool-first order on weights with non-trivial BN statistics, and against the fp32 oracle."""
    from oracle import models_ref
    kw = pkg("keras_weights")
    W = kw.generate("InceptionV3", "he_uniform", 7)
    sel = images[100:124]
    outs = {}
    for commute in (True, False):
        if commute:
            monkeypatch.delenv("DML_NO_POOL_COMMUTE", raising=False)
        else:
            monkeypatch.setenv("DML_NO_POOL_COMMUTE", "1")
        eng = pkg("engine").Engine("InceptionV3", max_batch=24, weights=W)
        try:
            followers = [o[0] for o in eng.ops() if o[1] == 0 and o[3] == 0]
            assert len(followers) == (28 if commute else 19), followers  # + the 9 pool-branch convs
            _, _, probs = eng.predict_images(sel, want_probs=True)
            outs[commute] = (probs, eng.logits(len(sel)).cpu().numpy().copy())
        finally:
            eng.close()
    ref_probs, ref_logits = models_ref.predict_arrays("InceptionV3", sel, W, batch=8)
    scale = np.abs(ref_logits).max()
    d_orders = np.abs(outs[True][1] - outs[False][1]).max() / scale
    d_ref = {c: np.abs(outs[c][1] - ref_logits).max() / scale for c in (True, False)}
    print(f"\npool-branch order: conv-first vs pool-first logits rel {d_orders:.2e}; vs oracle: conv-first {d_ref[True]:.2e}, pool-first {d_ref[False]:.2e}")
    assert d_orders <= 2e-2
    assert d_ref[True] <= 3e-2 and np.abs(outs[True][0] - ref_probs).max() <= 2e-2
    assert (outs[True][0].argmax(1) == ref_probs.argmax(1)).all()


def test_folding_stress_fp32(images):
    """Non-trivial biases and BN statistics: folding checked in fp32 mode."""
    from oracle import models_ref
    kw = pkg("keras_weights")
    for model in ("ResNet50", "InceptionV3"):
        W = kw.generate(model, "folding_stress", 3)
        sel = images[:6]
        _, ref_logits = models_ref.predict_arrays(model, sel, W, batch=6)
        eng = pkg("engine").Engine(model, max_batch=8, weights=W, fp32_check=True)
        try:
            eng.predict_images(sel)
            logits = eng.logits(len(sel)).cpu().numpy()
        finally:
            eng.close()
        rel = np.abs(logits - ref_logits).max() / np.abs(ref_logits).max()
        print(f"\n{model} folding_stress fp32: logits rel {rel:.3e}")
        assert rel <= 1e-3


@pytest.mark.parametrize("model", ["ResNet50", "InceptionV3"])
def test_deterministic_and_batch_invariant(model, images):
    """Same images -> bit-identical logits, run to run and whatever else shares the batch
    (stale shared memory / buffer contents must never leak into results)."""
    eng = pkg("engine").Engine(model, max_batch=64)
    try:
        sel = images[:48]
        eng.predict_images(sel)
        a = eng.logits(48).cpu().numpy().copy()
        eng.predict_images(images[60:70])
        eng.predict_images(sel)
        b = eng.logits(48).cpu().numpy().copy()
        eng.predict_images(sel[:7])
        c = eng.logits(7).cpu().numpy().copy()
    finally:
        eng.close()
    assert np.array_equal(a, b)
    assert np.array_equal(a[:7], c)
