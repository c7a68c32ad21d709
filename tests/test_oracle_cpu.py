"""CPU tests: the oracle against its golden vectors, Pillow, and the reference's
sample output; the host-side front end against the oracle."""
import hashlib
import io
import json
import os

import numpy as np
import pytest
from PIL import Image

from conftest import GOLDEN, ROOT, pkg
from oracle import decode, frontend as ofe, keras_nets, models_ref


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_param_counts_match_keras():
    # keras.applications' published parameter counts
    assert sum(int(np.prod(s)) for _, s in keras_nets.weight_spec("ResNet50")) == 25_636_712
    assert sum(int(np.prod(s)) for _, s in keras_nets.weight_spec("InceptionV3")) == 23_851_784
    assert len(keras_nets.weight_spec("ResNet50")) == 320
    assert len(keras_nets.weight_spec("InceptionV3")) == 378


def test_nearest_table_matches_pillow():
    rng = np.random.default_rng(0)
    sizes = list(range(100, 301, 7)) + [128, 154, 170, 182, 192, 204, 212, 218, 220, 222, 224, 234, 240, 254, 256,
                                        262, 270, 278, 284, 286, 292, 299, 300, 1, 2, 3]
    for n_in in sizes:
        col = np.arange(n_in, dtype=np.uint8 if n_in <= 256 else np.int32)
        img = Image.fromarray((np.arange(n_in) % 251).astype(np.uint8)[None, :].repeat(2, 0))
        for n_out in (224, 299):
            want = np.asarray(img.resize((n_out, 2), Image.NEAREST))[0]
            tab = ofe.pillow_nearest_table(n_in, n_out)
            assert np.array_equal((tab % 251).astype(np.uint8), want), (n_in, n_out)
            assert np.array_equal(pkg("frontend").nearest_table(n_in, n_out), tab) or n_in == n_out


def test_frontend_golden_hashes(jpeg_bytes, frontend_golden):
    g = frontend_golden
    fe = pkg("frontend")
    for i, name in enumerate(g["names"]):
        data = jpeg_bytes[str(name)]
        rgb = ofe.decode_rgb(data)
        assert sha(rgb) == g["sha_rgb"][i]
        assert tuple(rgb.shape[:2]) == tuple(g["hw"][i])
        assert np.array_equal(fe.decode_rgb(data), rgb)
        for size, mode in ((224, "caffe"), (299, "tf")):
            r = ofe.load_img(data, size)
            assert sha(r) == g[f"sha_resized_{size}"][i]
            assert np.array_equal(ofe.resize_nearest_tables(rgb, size), r)
            if i % 10 == 0:
                assert sha(ofe.preprocess(r, mode)) == g[f"sha_pre_{size}_{mode}"][i]
    # host batch tables == oracle tables
    imgs = [ofe.decode_rgb(jpeg_bytes[str(n)]) for n in g["names"][:8]]
    hb = fe.HostBatch(imgs, 299)
    for i, im in enumerate(imgs):
        assert np.array_equal(hb.ytab[i], ofe.pillow_nearest_table(im.shape[0], 299))
        assert np.array_equal(hb.xtab[i], ofe.pillow_nearest_table(im.shape[1], 299))
        assert np.array_equal(hb.pixels[hb.img_off[i]:hb.img_off[i] + im.size], im.ravel())


def test_preprocess_modes():
    v = np.arange(256, dtype=np.uint8)
    rgb = np.stack([v, v[::-1], (v.astype(np.int32) * 7) % 256], -1).astype(np.uint8)[None]
    c = ofe.preprocess(rgb, "caffe")
    assert c.dtype == np.float32
    assert np.array_equal(c[0, :, 0], rgb[0, :, 2].astype(np.float32) - np.float32(103.939))
    assert np.array_equal(c[0, :, 2], rgb[0, :, 0].astype(np.float32) - np.float32(123.68))
    t = ofe.preprocess(rgb, "tf")
    assert t.min() == -1.0 and t.max() == 1.0
    assert np.array_equal(t[0, :, 0], (v.astype(np.float32) / np.float32(127.5)) - np.float32(1))


def test_oracle_reproduces_golden(images, oracle_golden):
    kw = pkg("keras_weights")
    for model in ("ResNet50", "InceptionV3"):
        g = oracle_golden[model]
        W = kw.generate(model, str(g["preset"]), int(g["seed"]))
        sel = [0, 57, 123, 199]
        probs, logits = models_ref.predict_arrays(model, [images[i] for i in sel], W, batch=1)
        assert np.abs(logits - g["logits"][sel]).max() <= 1e-3 * np.abs(g["logits"]).max()
        for k, i in enumerate(sel):
            assert list(decode.top5_indices(probs[k])) == list(g["top5_idx"][i])


def test_avgpool_and_layer_semantics():
    import torch
    x = torch.arange(16, dtype=torch.float32).reshape(1, 1, 4, 4)
    y = keras_nets._avgpool_same(x)
    assert y[0, 0, 0, 0].item() == pytest.approx((0 + 1 + 4 + 5) / 4)  # corner divides by 4 in-bounds taps
    assert y[0, 0, 0, 1].item() == pytest.approx((0 + 1 + 2 + 4 + 5 + 6) / 6)
    assert y[0, 0, 1, 1].item() == pytest.approx(sum([0, 1, 2, 4, 5, 6, 8, 9, 10]) / 9)


def test_decode_predictions_and_json(tmp_path):
    ci = pkg("models").class_index()
    assert len(ci) == 1000 and ci[153] == ("n02085936", "Maltese_dog")
    p = np.full((1, 1000), 1e-4, dtype=np.float32)
    p[0, [5, 7, 900]] = [0.3, 0.3, 0.2]
    d = decode.decode_predictions(p, ci)
    assert [r[1] for r in d[0][:3]] == [ci[7][1], ci[5][1], ci[900][1]]  # tie: larger index first
    assert d[0][3][1] == ci[999][1] and d[0][4][1] == ci[998][1]
    with pytest.raises(ValueError):
        decode.decode_predictions(np.zeros((1000,), np.float32), ci)
    # the product's dump_to_file writes exactly what the reference's would
    res = {"1.jpeg": [[(w, n, pr) for (w, n, pr) in d[0]]]}
    out = tmp_path / "o.json"
    pkg("models").dump_to_file(res, str(out))
    assert out.read_text() == decode.dumps(res)
    back = json.loads(out.read_text())
    assert back["1.jpeg"][0][0] == [ci[7][0], ci[7][1], float(np.float32(0.3))]


def test_reference_sample_output_pairs():
    """(wnid, name) pairs of the reference's own sample outputs agree with our class index."""
    ci = pkg("models").class_index()
    by_name = {v[1]: v[0] for v in ci.values()}
    sample = os.path.join(GOLDEN, "reference_sample_pairs.json")
    pairs = json.load(open(sample))
    assert len(pairs) >= 30
    for wnid, name in pairs:
        assert by_name[name] == wnid


def test_model_parameters_and_merge():
    m = pkg("models")
    assert m.ModelParameters(1, 5.6, 2, 0.325, 10).execution_time_per_vm() == pytest.approx(10 + 5.6 + 2 + 0.325 * 9)
    d2 = {"a": 1}
    assert m.Merge({"b": 2}, d2) is None and d2 == {"a": 1, "b": 2}
    assert m.perform_inference_without_async("VGG16", []) == "Invalid model"
    import asyncio
    assert asyncio.run(m.perform_inference("nope", [])) == "Invalid model"
