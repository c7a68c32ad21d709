"""Builder probe: does splitting one batch over several concurrent engines (own arena, own stream, shared GPU) beat one
whole-batch engine?  Two pipelines fill each other's kernel tails.  Not a bench record.

    python tools/lanes_probe.py ResNet50 256 1 2 4
"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "distributed-machine-learning_b200"


def main():
    import torch
    import bench
    engine_mod = importlib.import_module(PKG + ".engine")
    fe = importlib.import_module(PKG + ".frontend")
    kw = importlib.import_module(PKG + ".keras_weights")
    model, B = sys.argv[1], int(sys.argv[2])
    W = kw.generate(model)
    dev = torch.device("cuda", 0)
    for lanes in [int(v) for v in sys.argv[3:]]:
        n = B // lanes
        engs, data = [], []
        for l in range(lanes):
            eng = engine_mod.Engine(model, device=0, max_batch=n, weights=W)
            imgs = bench.synthetic_images(model, n, seed=l)
            hb = fe.HostBatch(imgs, eng.size)
            data.append(tuple(torch.from_numpy(getattr(hb, k)).to(dev) for k in ("pixels", "img_off", "hw", "xtab", "ytab")))
            engs.append(eng)
        for _ in range(5):
            for eng, d in zip(engs, data):
                eng.infer_ptrs(*d, n)
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            t0 = time.perf_counter()
            for _ in range(30):
                for eng, d in zip(engs, data):
                    eng.infer_ptrs(*d, n)
            torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t0) / 30)
        print(f"{model} batch {B} as {lanes} x {n}: {best * 1e3:.3f} ms/step  {B / best:.0f} images/s", flush=True)
        for eng in engs:
            eng.close()
        del engs, data
        torch.cuda.empty_cache()
    return 0


if __name__ == "__main__":
    sys.exit(main())
