"""BASELINE.json configs[4]: concurrent ResNet50 + InceptionV3 jobs under the restated fair-share
scheduler, C3 batch sizes 1 / 8 / 64, per-chunk latency p50 / p99.

    python tools/job_mix.py [--gpus N] [--jobs-per-model J] [--images-per-job I] [--fix-off-by-one]

One worker thread per GPU (the reference's worker VMs H3..H10 -> GPU 0..N-1), each with a resident
ResNet50 and InceptionV3 engine.  The main thread is the leader: it submits the jobs, calls
FairShareScheduler.schedule() whenever a chunk is acknowledged (as worker.py:1025-1026 does) and
pushes assignments into per-GPU queues.  Latency of a chunk = acknowledgement time - time the worker
received the task (worker.py:959, :1001), i.e. file read + JPEG decode + H2D + kernels + D2H.
Files are the 200 bundled test JPEGs (tests/golden/testfiles.tar).  Prints one JSON line.
"""
import argparse
import importlib
import json
import os
import queue
import sys
import tarfile
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "distributed-machine-learning_b200"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--jobs-per-model", type=int, default=4)
    ap.add_argument("--images-per-job", type=int, default=200)
    ap.add_argument("--c3", default="1,8,64")
    ap.add_argument("--fix-off-by-one", action="store_true")
    args = ap.parse_args()

    import torch
    engine_mod = importlib.import_module(PKG + ".engine")
    sched_mod = importlib.import_module(PKG + ".scheduler")
    models = importlib.import_module(PKG + ".models")
    ngpu = min(args.gpus, torch.cuda.device_count())
    tmp = tempfile.mkdtemp(prefix="dml_mix_")
    with tarfile.open(os.path.join(ROOT, "tests", "golden", "testfiles.tar")) as tf:
        tf.extractall(tmp)
    sdfs = sorted(f for f in os.listdir(tmp) if f.endswith(".jpeg"))[:100]  # the 100-file list of worker.py:198-206

    workers = [f"GPU{i}" for i in range(ngpu)]
    engines = {w: {m: engine_mod.Engine(m, device=i, max_batch=72) for m in ("ResNet50", "InceptionV3")}
               for i, w in enumerate(workers)}
    inbox = {w: queue.Queue() for w in workers}
    acks = queue.Queue()

    def worker_loop(w):
        while True:
            task = inbox[w].get()
            if task is None:
                return
            t_recv = time.perf_counter()
            files = [os.path.join(tmp, f) for f in task.images]
            res = models._run(task.model, files, engines[w][task.model])
            acks.put((w, task, time.perf_counter() - t_recv, len(res)))

    threads = [threading.Thread(target=worker_loop, args=(w,), daemon=True) for w in workers]
    for t in threads:
        t.start()
    # warm-up: one chunk per engine
    for w in workers:
        for m in ("ResNet50", "InceptionV3"):
            models._run(m, [os.path.join(tmp, f) for f in sdfs[:8]], engines[w][m])

    report = {}
    for c3 in [int(x) for x in args.c3.split(",")]:
        sched = sched_mod.FairShareScheduler(workers, batch_size=c3, fix_off_by_one=args.fix_off_by_one)
        for m in sched_mod.MODELS:
            sched.set_batch_size(m, c3)
        jid = 0
        for j in range(args.jobs_per_model):
            for m in ("ResNet50", "InceptionV3"):
                jid += 1
                sched.submit_job(m, args.images_per_job, jid, sdfs)
        lat = {"ResNet50": [], "InceptionV3": []}
        cancelled = set()
        images_done = 0
        t0 = time.perf_counter()

        def dispatch():
            for a in sched.schedule():
                if a.preempted is not None:
                    cancelled.add((a.worker, a.preempted["job_id"], a.preempted["batch_id"]))
                inbox[a.worker].put(a)

        dispatch()
        while not sched.idle():
            w, task, dt, n = acks.get(timeout=120)  # a stuck schedule fails loudly instead of hanging
            key = (w, task.job_id, task.batch_id)
            if key in cancelled:          # preempted while running: result dropped, chunk already requeued
                cancelled.discard(key)
                continue
            cur = sched.workers_tasks_dict.get(w)
            if cur is None or (cur["job_id"], cur["batch_id"]) != (task.job_id, task.batch_id):
                continue
            sched.task_done(w)
            lat[task.model].append(dt)
            images_done += n
            dispatch()
        wall = time.perf_counter() - t0
        report[f"c3={c3}"] = {
            "chunk_images": c3 + (0 if args.fix_off_by_one else 1), "wall_s": wall, "images_per_s": images_done / wall,
            **{m: {"chunks": len(v), "p50_ms": float(np.percentile(v, 50) * 1e3), "p99_ms": float(np.percentile(v, 99) * 1e3),
                   "mean_ms": float(np.mean(v) * 1e3)} for m, v in lat.items() if v}}
    for w in workers:
        inbox[w].put(None)
    print(json.dumps({"metric": "per_chunk_latency", "n_gpus": ngpu, "jobs_per_model": args.jobs_per_model,
                      "images_per_job": args.images_per_job, "files": "200 bundled JPEGs, first 100 names (sorted)",
                      "off_by_one": not args.fix_off_by_one, "results": report}))


if __name__ == "__main__":
    main()
