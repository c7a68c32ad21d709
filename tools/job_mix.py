"""BASELINE.json configs[4]: concurrent ResNet50 + InceptionV3 jobs under the restated fair-share
scheduler, C3 batch sizes 1 / 8 / 64, per-chunk latency p50 / p99.

    python tools/job_mix.py [--gpus N] [--jobs-per-model J] [--images-per-job I] [--fix-off-by-one] [--out FILE]

One worker PROCESS per GPU (workers.GpuWorkerPool: the reference's worker VMs H3..H10 -> GPU 0..N-1), each with
resident ResNet50 and InceptionV3 engines and its own JPEG decode pool of cores / N threads.  This process is the
leader: it submits the jobs, calls FairShareScheduler.schedule() whenever a chunk is acknowledged (as
worker.py:1025-1026 does) and hands assignments to the per-GPU task queues; an assignment to a busy worker preempts
its task (worker.py:944-953).  Latency of a chunk = acknowledgement - receipt of the task in the worker
(worker.py:959, :1001), i.e. file read + JPEG decode + H2D + kernels + D2H.  Files are the 200 bundled test JPEGs
(tests/golden/testfiles.tar).  Prints one JSON line.
"""
import argparse
import importlib
import json
import os
import sys
import tarfile
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "distributed-machine-learning_b200"


def run_mix(pool, workers, sdfs, tmp, c3, jobs_per_model, images_per_job, fix_off_by_one=False, wire_cap=None):
    """Drives one job mix to completion; returns the report dict and the per-chunk acknowledgement counts."""
    sched_mod = importlib.import_module(PKG + ".scheduler")
    sched = sched_mod.FairShareScheduler(workers, batch_size=c3, fix_off_by_one=fix_off_by_one)
    for m in sched_mod.MODELS:
        sched.set_batch_size(m, c3, wire_cap=wire_cap)
    jid = 0
    all_chunks = []
    for _ in range(jobs_per_model):
        for m in ("ResNet50", "InceptionV3"):
            jid += 1
            all_chunks += [(m, c["job_id"], c["batch_id"]) for c in sched.submit_job(m, images_per_job, jid, sdfs)]
    lat = {"ResNet50": [], "InceptionV3": []}
    acked = {}
    preemptions = 0
    images_done = 0
    widx = {w: i for i, w in enumerate(workers)}
    t0 = time.perf_counter()

    def dispatch():
        nonlocal preemptions
        for a in sched.schedule():
            preemptions += a.preempted is not None
            assert a.fits_datagram(), "a task request beyond the reference's 32 KiB datagram"
            pool.submit(widx[a.worker], (a.model, a.job_id, a.batch_id), a.model, [os.path.join(tmp, f) for f in a.images])

    dispatch()
    while not sched.idle():
        wi, task_id, status, dt, n, info = pool.get_ack(timeout=120)  # a stuck schedule fails loudly instead of hanging
        if status == "error":
            raise RuntimeError(f"worker {wi}: {info}")
        w = workers[wi]
        model, job_id, batch_id = task_id
        cur = sched.workers_tasks_dict.get(w)
        if status != "ok" or cur is None or (cur["job_id"], cur["batch_id"]) != (job_id, batch_id):
            continue  # preempted (or finished after it was preempted): the chunk is back in its queue, this result is dropped
        sched.task_done(w)
        acked[task_id] = acked.get(task_id, 0) + 1
        lat[model].append(dt)
        images_done += n
        dispatch()
    wall = time.perf_counter() - t0
    rep = {"chunk_images": c3 + (0 if fix_off_by_one else 1), "wall_s": wall, "images_per_s": images_done / wall, "preemptions": preemptions,
           **{m: {"chunks": len(v), "p50_ms": float(np.percentile(v, 50) * 1e3), "p99_ms": float(np.percentile(v, 99) * 1e3),
                  "mean_ms": float(np.mean(v) * 1e3)} for m, v in lat.items() if v}}
    return rep, acked, all_chunks


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--jobs-per-model", type=int, default=4)
    ap.add_argument("--images-per-job", type=int, default=200)
    ap.add_argument("--c3", default="1,8,64")
    ap.add_argument("--fix-off-by-one", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    import torch
    workers_mod = importlib.import_module(PKG + ".workers")
    sched_mod = importlib.import_module(PKG + ".scheduler")
    ngpu = min(args.gpus, torch.cuda.device_count())
    tmp = tempfile.mkdtemp(prefix="dml_mix_")
    with tarfile.open(os.path.join(ROOT, "tests", "golden", "testfiles.tar")) as tf:
        tf.extractall(tmp)
    sdfs = sorted(f for f in os.listdir(tmp) if f.endswith(".jpeg"))[:100]  # the 100-file list of worker.py:198-206
    workers = [f"GPU{i}" for i in range(ngpu)]
    pool = workers_mod.GpuWorkerPool(list(range(ngpu)), max_batch=72, warm_files=[os.path.join(tmp, f) for f in sdfs[:8]])
    cap = sched_mod.max_images_per_task([f"{i}.jpeg" for i in range(1, 401)])
    report = {}
    try:
        for c3 in [int(x) for x in args.c3.split(",")]:
            rep, acked, chunks = run_mix(pool, workers, sdfs, tmp, c3, args.jobs_per_model, args.images_per_job, args.fix_off_by_one, wire_cap=cap)
            assert sorted(acked) == sorted(chunks) and set(acked.values()) == {1}, "every chunk is acknowledged exactly once"
            report[f"c3={c3}"] = rep
    finally:
        pool.close()
    line = json.dumps({"metric": "per_chunk_latency", "n_gpus": ngpu, "workers": "one process per GPU", "decode_threads_per_worker": pool.decode_threads,
                       "host_cores": os.cpu_count(), "jobs_per_model": args.jobs_per_model, "images_per_job": args.images_per_job,
                       "files": "200 bundled JPEGs, first 100 names (sorted)", "off_by_one": not args.fix_off_by_one,
                       "task_datagram_cap_images": cap, "results": report})
    print(line)
    if args.out:
        with open(args.out, "w") as fh:
            fh.write(line + "\n")


if __name__ == "__main__":
    main()
