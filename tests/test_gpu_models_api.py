"""The drop-in models module on the GPU: same call shapes and result structure as
/root/reference/models.py (worker.py:1358, :1379, :1452 call sites)."""
import asyncio
import json
import os

import numpy as np
import pytest

from conftest import pkg
from oracle import decode

pytestmark = pytest.mark.gpu


def test_perform_inference_matches_oracle_golden(image_dir, oracle_golden):
    m = pkg("models")
    for model, lo in (("ResNet50", 1), ("InceptionV3", 101)):
        files = [os.path.join(image_dir, f"{i}.jpeg") for i in range(lo, lo + 100)]
        res = asyncio.run(m.perform_inference(model, files))
        assert list(res) == [f"{i}.jpeg" for i in range(lo, lo + 100)]
        g = oracle_golden[model]
        ci = m.class_index()
        for k, name in enumerate(res):
            row = lo - 1 + k
            v = res[name]
            assert isinstance(v, list) and len(v) == 1 and len(v[0]) == 5  # extra outer list, top-5
            wnid, cname, p = v[0][0]
            assert (wnid, cname) == ci[int(g["top5_idx"][row][0])]
            assert isinstance(p, np.float32) and abs(float(p) - float(g["top5_prob"][row][0])) <= 2e-2
            assert [t[2] for t in v[0]] == sorted((t[2] for t in v[0]), reverse=True)
        res2 = m.perform_inference_without_async(model, files[:10])
        assert {k: [[t[:2] for t in v[0]]] for k, v in res2.items()} == \
               {k: [[t[:2] for t in res[k][0]]] for k in list(res)[:10]}


def test_output_json_shape_and_bad_files(image_dir, tmp_path):
    m = pkg("models")
    bad = tmp_path / "broken.jpeg"
    bad.write_bytes(b"not a jpeg")
    files = [os.path.join(image_dir, "117.jpeg"), str(bad), os.path.join(image_dir, "missing.jpeg")]
    res = m.run_inference_on_InceptionV3(files)
    assert list(res) == ["117.jpeg", "broken.jpeg", "missing.jpeg"]
    assert isinstance(res["broken.jpeg"], str) and isinstance(res["missing.jpeg"], str)
    out = tmp_path / "output_1_127.json"
    m.dump_to_file(res, str(out))
    text = out.read_text()
    assert text.endswith("\n") and text == decode.dumps(res)
    back = json.loads(text)
    top = back["117.jpeg"][0][0]
    assert len(back["117.jpeg"]) == 1 and len(back["117.jpeg"][0]) == 5
    assert isinstance(top[0], str) and isinstance(top[1], str) and isinstance(top[2], float)
    assert m.perform_inference_without_async("VGG16", files) == "Invalid model"
    assert m.run_inference_on_ResNet50([]) == {}


def test_pipelined_submit_collect_matches_serial(images):
    """Two batches in flight (H2D overlapping kernels) give the same answers as the serial call."""
    fe = pkg("frontend")
    eng = pkg("engine").Engine("ResNet50", max_batch=16)
    try:
        chunks = [images[i:i + 16] for i in range(0, 80, 16)]
        want = [eng.predict_images(c) for c in chunks]
        hbs = [eng.pin(fe.HostBatch(c, eng.size)) for c in chunks]
        got, pending = [], None
        for hb in hbs:
            t = eng.submit(hb)
            if pending is not None:
                i, p = eng.collect(pending)
                got.append((i.copy(), p.copy()))
            pending = t
        i, p = eng.collect(pending)
        got.append((i.copy(), p.copy()))
        for (wi, wp), (gi, gp) in zip(want, got):
            assert np.array_equal(wi, gi) and np.array_equal(wp, gp)
        with pytest.raises(pkg("_lib").DmlError):
            eng.submit(hbs[0]); eng.submit(hbs[1]); eng.submit(hbs[2])
    finally:
        eng.sync()
        eng.close()
