"""Parity at the plans bench.py times (BASELINE.json configs[2] / configs[3]) and at the task chunkings of
configs[0] / configs[1].

bench.py builds Engine(model, max_batch = 256 | 128, micro_batch = 0): whole-batch layers, autotuned on those
shapes.  The other GPU tests use smaller engines, so here the very construction bench.py uses is fed bench.py's
own synthetic images and compared with the CPU oracle (reference loop: /root/reference/models.py:30-44, :55-69):
softmax within 2e-2 max-abs (BASELINE.json's bf16 tolerance), logits relative, top-1 wherever the oracle's own
margin is larger than the measured error.
"""
import asyncio
import os

import numpy as np
import pytest

from conftest import pkg
from oracle import models_ref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model,batch", [("ResNet50", 256), ("InceptionV3", 128)])
def test_benchmarked_plan_against_oracle(model, batch):
    import bench
    kw = pkg("keras_weights")
    fe = pkg("frontend")
    W = kw.generate(model)
    imgs = bench.synthetic_images(model, batch, seed=0)
    eng = pkg("engine").Engine(model, max_batch=batch, micro_batch=0, load=False)  # bench.py's construction
    try:
        eng.load_weights(W)
        hb = fe.HostBatch(imgs, eng.size)
        eng.upload(hb)
        eng.infer_device(batch, want_probs=True)
        eng.sync()
        probs = eng.d_probs[:batch].cpu().numpy()
        idx = eng.d_idx[:batch].cpu().numpy()
        logits = eng.logits(batch).cpu().numpy()
        # the pipelined host-buffer path bench.py's e2e leg uses gives the same top-5
        t = eng.submit(eng.pin(hb))
        idx2, prob2 = eng.collect(t)
        assert np.array_equal(idx2, idx)
        assert np.array_equal(prob2, np.take_along_axis(probs, idx, axis=1))
    finally:
        eng.close()
    ref_probs, ref_logits = models_ref.predict_arrays(model, imgs, W, batch=32)
    dsm = float(np.abs(probs - ref_probs).max())
    dl = float(np.abs(logits - ref_logits).max())
    rel = dl / float(np.abs(ref_logits).max())
    srt = np.sort(ref_logits, axis=1)
    margin = srt[:, -1] - srt[:, -2]
    decided = margin > 4.0 * dl  # rounding alone cannot flip these
    same = idx[:, 0] == ref_probs.argmax(1)
    print(f"\n{model} batch {batch} (bench plan): softmax max-abs {dsm:.3e}  logits rel {rel:.3e}  "
          f"top-1 {int(same.sum())}/{batch} ({int(decided.sum())} with margin > 4x max|dlogit|)")
    assert np.isfinite(logits).all()
    assert dsm <= 2e-2
    assert rel <= 3e-2
    assert same[decided].all()
    assert same.mean() >= 0.9


def _strip(res):
    return {k: [[(t[0], t[1], float(t[2])) for t in v[0]]] if not isinstance(v, str) else v for k, v in res.items()}


@pytest.mark.parametrize("model,lo,chunks", [("ResNet50", 1, (10, 11)), ("InceptionV3", 101, (32, 33))])
def test_baseline_chunkings_match_single_batch(model, lo, chunks, image_dir):
    """configs[0]: ResNet50, testfiles/1-100.jpeg in tasks of 10 (and 11: the leader's chunker hands out
    batch_size + 1 files, /root/reference/worker.py:216-227); configs[1]: InceptionV3, 101-200 in tasks of 32.
    Every task goes through perform_inference like a WORKER_TASK_REQUEST does (worker.py:1379); the merged result
    equals the one-task run bit for bit."""
    m = pkg("models")
    files = [os.path.join(image_dir, f"{i}.jpeg") for i in range(lo, lo + 100)]
    whole = _strip(asyncio.run(m.perform_inference(model, files)))
    assert len(whole) == 100
    for c in chunks:
        merged = {}
        for i in range(0, 100, c):
            part = asyncio.run(m.perform_inference(model, files[i:i + c]))
            assert list(part) == [os.path.basename(f) for f in files[i:i + c]]
            merged.update(_strip(part))
        assert merged == whole
