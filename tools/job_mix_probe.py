"""Probe of the C3 = 1 job mix: where does a chunk's time go between leader and workers?

    python tools/job_mix_probe.py N [same]     N workers on GPUs 0..N-1 (or all on GPU 0 with `same`)

Per chunk: leader submit -> worker receipt (task queue), receipt -> completion (decode + GPU), completion -> leader ack (ack queue),
and the fraction of the wall time each worker was busy.
"""
import collections, importlib, os, sys, tarfile, tempfile, time
import numpy as np
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
    devs = [0] * n if len(sys.argv) > 2 else list(range(n))
    workers = [f"GPU{i}" for i in range(n)]
    pool = workers_mod.GpuWorkerPool(devs, max_batch=16, warm_files=[os.path.join(tmp, f) for f in sdfs[:8]], timestamps=True)
    t_submit, rec = {}, []
    orig_submit, orig_ack = pool.submit, pool.get_ack
    def submit(w, tid, model, files):
        t_submit[(w, tid)] = time.time(); orig_submit(w, tid, model, files)
    def get_ack(timeout=None):
        r = orig_ack(timeout); t = time.time()
        if r[2] == "ok" and isinstance(r[5], tuple):
            rec.append((r[0], t_submit.get((r[0], r[1]), t), r[5][1], r[5][2], t))
        return r
    pool.submit, pool.get_ack = submit, get_ack
    for c3 in (1, 8):
        rec.clear()
        t0 = time.time()
        rep, acked, chunks = job_mix.run_mix(pool, workers, sdfs, tmp, c3, 2, 200)
        wall = time.time() - t0
        a = np.array([(r[2] - r[1], r[3] - r[2], r[4] - r[3]) for r in rec]) * 1e3
        busy = collections.defaultdict(float)
        for r in rec: busy[r[0]] += r[3] - r[2]
        print(f"C3={c3} {n} workers on {'one GPU' if len(sys.argv) > 2 else str(n) + ' GPUs'}: {rep['images_per_s']:.0f} img/s, wall {wall:.2f} s, {len(rec)} chunks | "
              f"ms p50/p99: to worker {np.percentile(a[:,0],50):.2f}/{np.percentile(a[:,0],99):.2f}, in worker {np.percentile(a[:,1],50):.2f}/{np.percentile(a[:,1],99):.2f}, "
              f"to leader {np.percentile(a[:,2],50):.2f}/{np.percentile(a[:,2],99):.2f} | worker busy fraction {min(busy.values())/wall:.2f}..{max(busy.values())/wall:.2f}")
    pool.close()
