"""Finds the first tensor that differs between two runs / two batch sizes (GPU box)."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
E = importlib.import_module("distributed-machine-learning_b200.engine")
model = sys.argv[1] if len(sys.argv) > 1 else "InceptionV3"
rng = np.random.default_rng(5)
imgs = [rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8) for h, w in [(225, 300), (300, 300), (224, 224), (114, 169)] * 6]
eng = E.Engine(model, max_batch=24, keep_all=True)
names = list(eng.tensors())
def run(n):
    eng.predict_images(imgs[:n])
    out = {}
    for nm in names:
        t = eng.tensors()[nm]
        if t[1] * t[2] == 1:
            continue
        out[nm] = eng.read_tensor(nm, n)
    out["logits"] = eng.logits(n).cpu().numpy()
    return out
a = run(24); b = run(24); c = run(5); d = run(24)
for label, x, y, n in (("run1 vs run2 (n=24)", a, b, 24), ("n=24 vs n=5", a, c, 5), ("run1 vs run4 (n=24)", a, d, 24)):
    bad = [(nm, float(np.abs(x[nm][:n] - y[nm][:n]).max()), int((x[nm][:n] != y[nm][:n]).sum())) for nm in x if not np.array_equal(x[nm][:n], y[nm][:n])]
    print(label, "differing tensors:", len(bad))
    for nm, mx, cnt in bad[:6]:
        print("   ", nm, "max|d|", mx, "count", cnt)
