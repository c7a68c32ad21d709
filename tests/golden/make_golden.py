"""Generates the committed fixtures under tests/golden/ (run in the BUILD container,
where /root/reference exists; the GPU box only reads the outputs).

  python tests/golden/make_golden.py [--search]

  testfiles.tar        the reference's 200 bundled test JPEGs (testfiles/1-100.jpeg,
                       testfiles_more/101-200.jpeg), raw bytes, packed
  frontend.npz         per image: (H, W), sha256 of the decoded RGB bytes and of
                       load_img's output at 224 and 299 (Pillow's own resize), and of
                       the fp32 preprocess_input result in caffe / tf mode
  <model>_oracle.npz   seed/preset, oracle logits [200,1000] fp32, top-5 idx/prob

--search additionally tries seeds 0..15 per model and prints the minimum top-1
logit margin over the 200 images for each (the committed default seeds in
keras_weights.DEFAULT_PRESET were picked from that table).
Keras/TensorFlow are not installed, so the network outputs come from the CPU
restatement in oracle/ (parity unpinned); the front-end hashes come from Pillow,
the library the reference itself calls.
"""
import glob
import hashlib
import importlib
import io
import os
import sys
import tarfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from oracle import decode, frontend, keras_nets, models_ref  # noqa: E402

kw = importlib.import_module("distributed-machine-learning_b200.keras_weights")


def ref_files():
    num = lambda p: int(os.path.basename(p).split(".")[0])
    a = sorted(glob.glob(os.path.join(REF, "testfiles", "*.jpeg")), key=num)
    b = sorted(glob.glob(os.path.join(REF, "testfiles_more", "*.jpeg")), key=num)
    assert len(a) == 100 and len(b) == 100
    return a + b


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    files = ref_files()
    tar_path = os.path.join(HERE, "testfiles.tar")
    with tarfile.open(tar_path, "w") as tf:
        for f in files:
            ti = tarfile.TarInfo(os.path.basename(f))
            data = open(f, "rb").read()
            ti.size = len(data)
            ti.mtime = 0
            tf.addfile(ti, io.BytesIO(data))
    raw = [open(f, "rb").read() for f in files]
    imgs = [frontend.decode_rgb(d) for d in raw]
    names = [os.path.basename(f) for f in files]
    rec = {"names": np.array(names), "hw": np.array([im.shape[:2] for im in imgs], dtype=np.int32),
           "sha_rgb": np.array([sha(im) for im in imgs])}
    for size, mode in ((224, "caffe"), (299, "tf")):
        res = [frontend.load_img(d, size) for d in raw]  # Pillow's own resize
        for im, r in zip(imgs, res):  # the index-table restatement must agree with Pillow
            assert np.array_equal(frontend.resize_nearest_tables(im, size), r)
        rec[f"sha_resized_{size}"] = np.array([sha(r) for r in res])
        rec[f"sha_pre_{size}_{mode}"] = np.array([sha(frontend.preprocess(r, mode)) for r in res])
    np.savez_compressed(os.path.join(HERE, "frontend.npz"), **rec)

    if "--search" in sys.argv:
        for model in ("ResNet50", "InceptionV3"):
            preset = kw.DEFAULT_PRESET[model][0]
            for seed in range(16):
                W = kw.generate(model, preset, seed)
                _, lg = models_ref.predict_arrays(model, imgs, W, batch=25)
                s = np.sort(lg, 1)
                mg = s[:, -1] - s[:, -2]
                print(f"{model} {preset} seed {seed}: min margin {mg.min():.4f} median {np.median(mg):.4f} "
                      f"distinct top1 {len(set(lg.argmax(1)))} logit std {lg.std():.3f}", flush=True)

    for model in ("ResNet50", "InceptionV3"):
        preset, seed = kw.DEFAULT_PRESET[model]
        W = kw.generate(model, preset, seed)
        probs, logits = models_ref.predict_arrays(model, imgs, W, batch=25)
        # batch-1 (the reference's loop shape) must agree with the batched oracle
        p1, _ = models_ref.predict_arrays(model, imgs[:4], W, batch=1)
        d1 = float(np.abs(p1 - probs[:4]).max())
        print(model, 'batch-1 vs batch-25 oracle max|dprob|', d1)
        assert d1 < 1e-4
        top = np.stack([decode.top5_indices(r) for r in probs]).astype(np.int32)
        np.savez_compressed(os.path.join(HERE, f"{model}_oracle.npz"), preset=preset, seed=seed, names=np.array(names),
                            logits=logits.astype(np.float32), top5_idx=top,
                            top5_prob=np.take_along_axis(probs, top, 1).astype(np.float32))
        s = np.sort(logits, 1)
        print(model, preset, seed, "min top-1 logit margin", float((s[:, -1] - s[:, -2]).min()))


if __name__ == "__main__":
    main()
