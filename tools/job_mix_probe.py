"""Probe: 8 worker processes on ONE GPU, C3 = 1, leader loop under cProfile (where does a chunk's time go?)."""
import cProfile, importlib, os, pstats, sys, tarfile, tempfile, time, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import job_mix
if __name__ == "__main__":
    PKG = "distributed-machine-learning_b200"
    workers_mod = importlib.import_module(PKG + ".workers")
    tmp = tempfile.mkdtemp(prefix="dml_mix_")
    with tarfile.open(os.path.join(ROOT, "tests", "golden", "testfiles.tar")) as tf:
        tf.extractall(tmp)
    sdfs = sorted(f for f in os.listdir(tmp) if f.endswith(".jpeg"))[:100]
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    workers = [f"GPU{i}" for i in range(n)]
    pool = workers_mod.GpuWorkerPool([0] * n, max_batch=16, warm_files=[os.path.join(tmp, f) for f in sdfs[:8]])
    per_worker = collections.Counter()
    orig = pool.get_ack
    def get_ack(timeout=None):
        r = orig(timeout); per_worker[r[0]] += 1; return r
    pool.get_ack = get_ack
    pr = cProfile.Profile(); pr.enable()
    rep, acked, chunks = job_mix.run_mix(pool, workers, sdfs, tmp, 1, 2, 200)
    pr.disable()
    print(n, "workers:", round(rep["images_per_s"]), "img/s", rep["ResNet50"], dict(per_worker))
    pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
    pool.close()
