import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
E = importlib.import_module("distributed-machine-learning_b200.engine")
rng = np.random.default_rng(5)
imgs = [rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8) for h, w in [(225, 300), (300, 300), (224, 224), (114, 169)] * 6]
for tc in (True, False):
    eng = E.Engine("ResNet50", max_batch=24, keep_all=True, tensor_cores=tc)
    outs, ins = [], []
    for i in range(4):
        eng.predict_images(imgs)
        ins.append(eng.read_tensor("input", 24)); outs.append(eng.read_tensor("conv1_relu", 24))
    print("tensor_cores", tc)
    for i in range(1, 4):
        bad = np.argwhere(outs[i] != outs[0])
        print(" run", i, "input equal", np.array_equal(ins[i], ins[0]), "conv1 diff count", len(bad))
        if len(bad):
            print("   imgs", np.unique(bad[:, 0]), "rows", np.unique(bad[:, 1])[:20], "cols", np.unique(bad[:, 2])[:20], "ch", np.unique(bad[:, 3])[:10])
    eng.close()
