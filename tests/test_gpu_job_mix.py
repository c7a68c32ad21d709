"""The per-GPU worker processes under the restated leader (BASELINE.json configs[4]) on one GPU.

Two worker processes share GPU 0 (the reference's worker VMs are separate machines; two of them are enough for the
fair-share split and for a preemption to happen).  2 x 40 images per model, C3 = 4: every chunk must be acknowledged
exactly once even though a worker's running task is cancelled when the other model's job arrives
(/root/reference/worker.py:389-408 leader side, :944-953 worker side), and the results of a task equal the in-process call.
"""
import importlib
import os
import sys

import pytest

from conftest import ROOT, pkg

pytestmark = pytest.mark.gpu


def test_job_mix_every_chunk_acknowledged_once_under_preemption(image_dir):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    job_mix = importlib.import_module("job_mix")
    workers_mod = pkg("workers")
    sched_mod = pkg("scheduler")
    sdfs = sorted(f for f in os.listdir(image_dir) if f.endswith(".jpeg"))[:100]
    workers = ["GPU0a", "GPU0b"]
    pool = workers_mod.GpuWorkerPool([0, 0], max_batch=16, decode_threads=4, warm_files=[os.path.join(image_dir, f) for f in sdfs[:4]],
                                     return_results=True)
    try:
        # a worker's answer equals the in-process call (same engine code, another process)
        files = [os.path.join(image_dir, f) for f in sdfs[:5]]
        pool.submit(0, "probe", "ResNet50", files)
        wi, tid, status, dt, n, res = pool.get_ack(timeout=120)
        assert (wi, tid, status, n) == (0, "probe", "ok", 5) and dt > 0
        mine = pkg("models").run_inference_on_ResNet50(files)
        assert list(res) == list(mine)
        assert all([(a[0], a[1], float(a[2])) for a in res[k][0]] == [(b[0], b[1], float(b[2])) for b in mine[k][0]] for k in mine)
        # preemption by hand: a long task, then another one on the same worker before it can finish
        pool.submit(1, "long", "InceptionV3", [os.path.join(image_dir, f) for f in sdfs] * 3)
        pool.submit(1, "short", "ResNet50", files)
        got = {}
        while len(got) < 2:
            wi, tid, status, dt, n, _ = pool.get_ack(timeout=120)
            got[tid] = status
        assert got["short"] == "ok" and got["long"] in ("cancelled", "ok")
        # the leader's loop: ResNet50 jobs first so that the InceptionV3 job must take a worker away
        sched = sched_mod.FairShareScheduler(workers, batch_size=4)
        rep, acked, chunks = job_mix.run_mix(pool, workers, sdfs, image_dir + "/", 4, jobs_per_model=2, images_per_job=40)
        assert sorted(acked) == sorted(chunks) and set(acked.values()) == {1}
        assert rep["ResNet50"]["chunks"] == rep["InceptionV3"]["chunks"] == 16  # 2 jobs x 40 images in chunks of 5
        assert rep["images_per_s"] > 0
        del sched
    finally:
        pool.close()
