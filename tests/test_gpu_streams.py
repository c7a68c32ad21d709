"""Branch streams (csrc/graph.cc plan_streams, csrc/dml_api.cu run_layers): InceptionV3's independent branches on several CUDA
streams must give exactly the bits of the serial order -- the same kernels run on the same data, only their order in time
changes (/root/reference/models.py:40 modelObj.predict: one result whatever the executor's order).  Repeated to give a missing
dependency the chance to show."""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def _run(images, mode, monkeypatch, micro_batch=0, reps=3):
    kw = pkg("keras_weights")
    W = kw.generate("InceptionV3", "he_uniform", 7)
    monkeypatch.setenv("DML_BRANCH_STREAMS", str(mode))
    eng = pkg("engine").Engine("InceptionV3", max_batch=48, micro_batch=micro_batch, weights=W)
    try:
        sched, ns = eng.schedule()
        outs = []
        for _ in range(reps):
            idx, prob, probs = eng.predict_images(images, want_probs=True)
            outs.append((idx.copy(), prob.copy(), probs.copy(), eng.logits(len(images)).cpu().numpy()))
    finally:
        eng.close()
    return sched, ns, outs


def test_streams_give_the_serial_bits(images, monkeypatch):
    sel = images[100:141]  # 41 images
    _, ns0, ref = _run(sel, 0, monkeypatch, reps=1)
    assert ns0 == 1
    sched, ns, outs = _run(sel, 1, monkeypatch, reps=4)
    assert ns >= 3
    assert len({s for _, s, _ in sched}) == ns
    assert sum(1 for _, _, w in sched if w > 0) >= 30  # every block forks and joins
    for o in outs:
        for a, b in zip(o, ref[0]):
            assert np.array_equal(a, b)


def test_streams_with_micro_batches(images, monkeypatch):
    """run_layers is called once per micro-batch: the side streams must have joined before the next one reuses the arena."""
    sel = images[100:137]
    _, _, ref = _run(sel, 0, monkeypatch, micro_batch=8, reps=1)
    _, ns, outs = _run(sel, 1, monkeypatch, micro_batch=8, reps=3)
    assert ns >= 3
    for o in outs:
        for a, b in zip(o, ref[0]):
            assert np.array_equal(a, b)


def test_resnet50_has_no_side_streams(monkeypatch):
    kw = pkg("keras_weights")
    eng = pkg("engine").Engine("ResNet50", max_batch=8, weights=kw.generate("ResNet50", "keras_default", 0))
    try:
        sched, ns = eng.schedule()
        assert ns == 1 and all(s == 0 and w == 0 for _, s, w in sched)
    finally:
        eng.close()
