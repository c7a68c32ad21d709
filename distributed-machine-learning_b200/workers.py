"""One worker PROCESS per GPU: the reference's worker VMs H3..H10 (/root/reference/worker.py:52) mapped onto the GPUs of one box.

In the reference every worker is its own machine: it receives a WORKER_TASK_REQUEST, cancels whatever task it is running
(worker.py:944-953), downloads the images, runs perform_inference and acknowledges (worker.py:518-537).  Here each GPU
gets its own process with its own interpreter, resident engines and JPEG decode pool (cores / GPUs threads), so that
decoding for one GPU never waits on another GPU's Python work (threads of one interpreter share the GIL: with 8 worker
threads in one process a 2-image chunk took 6.0 ms against 1.5 ms on one GPU).  The leader talks to the workers through
per-worker task queues -- the per-VM task packets -- and one shared acknowledgement queue.

Inside a worker the reference's structure is kept: an asyncio loop, one task at a time, a new task request preempts the
running one (``task.cancel()`` then ``await``), results acknowledged with the task's latency measured from receipt
(worker.py:959 ``start_time``).  Nothing here touches the control plane (UDP, SDFS): the queues stand in for it.
"""
from __future__ import annotations

import asyncio
import importlib
import multiprocessing as mp
import os
import queue as _queue
import time

PKG = __name__.rsplit(".", 1)[0]


def _worker_main(index, device, inbox, acks, cfg):
    os.environ["DML_B200_DEVICE"] = str(device)
    os.environ["DML_B200_MAX_BATCH"] = str(cfg["max_batch"])
    os.environ["DML_B200_DECODE_THREADS"] = str(cfg["decode_threads"])
    try:
        models = importlib.import_module(PKG + ".models")
        import warnings
        warnings.simplefilter("ignore", RuntimeWarning)  # the fixed-seed-weights notice: once per process is enough, in the leader
        for m in cfg["models"]:
            eng = models.get_engine(m, device)
            if cfg.get("warm_files"):
                models._run(m, list(cfg["warm_files"]), eng)
    except BaseException as exc:  # report instead of dying silently: the leader would wait forever
        acks.put((index, None, "failed", 0.0, 0, f"{type(exc).__name__}: {exc}"))
        return
    acks.put((index, None, "ready", 0.0, 0, None))

    async def run(task_id, model, files, t_recv):
        try:
            res = await models.perform_inference(model, files)
            n = len(res) if isinstance(res, dict) else 0
            info = res if cfg["return_results"] else (("wall", cfg["_t_recv_wall"], time.time()) if cfg.get("timestamps") else None)
            acks.put((index, task_id, "ok", time.perf_counter() - t_recv, n, info))
        except asyncio.CancelledError:
            acks.put((index, task_id, "cancelled", time.perf_counter() - t_recv, 0, None))
            raise
        except BaseException as exc:
            acks.put((index, task_id, "error", time.perf_counter() - t_recv, 0, f"{type(exc).__name__}: {exc}"))

    async def serve():
        loop = asyncio.get_running_loop()
        current = None
        while True:
            msg = await loop.run_in_executor(None, inbox.get)
            if msg is None:
                break
            task_id, model, files = msg
            t_recv = time.perf_counter()
            cfg["_t_recv_wall"] = time.time()
            if current is not None and not current.done():
                current.cancel()  # worker.py:944-953: a new task request preempts the running task
                try:
                    await current
                except asyncio.CancelledError:
                    pass
            current = asyncio.create_task(run(task_id, model, files, t_recv))
        if current is not None:
            try:
                await current
            except BaseException:
                pass

    asyncio.run(serve())


class GpuWorkerPool:
    """``devices``: one entry per worker (a device may appear more than once: several workers sharing a GPU)."""

    def __init__(self, devices, models=("ResNet50", "InceptionV3"), max_batch=72, decode_threads=None, warm_files=None,
                 return_results=False, start_timeout=600.0, timestamps=False):
        self.devices = list(devices)
        n = len(self.devices)
        if decode_threads is None:
            decode_threads = max(1, (os.cpu_count() or 4) // n)
        ctx = mp.get_context("spawn")  # CUDA does not survive a fork
        self.acks = ctx.Queue()
        self.inboxes = [ctx.Queue() for _ in range(n)]
        cfg = {"models": tuple(models), "max_batch": int(max_batch), "decode_threads": int(decode_threads),
               "warm_files": list(warm_files) if warm_files else None, "return_results": bool(return_results),
               "timestamps": bool(timestamps)}  # timestamps: acknowledgements carry (wall clock at receipt, at completion) for latency studies
        self.procs = [ctx.Process(target=_worker_main, args=(i, d, self.inboxes[i], self.acks, cfg), daemon=True)
                      for i, d in enumerate(self.devices)]
        for p in self.procs:
            p.start()
        ready, deadline = 0, time.monotonic() + start_timeout
        while ready < n:
            try:
                idx, _, status, _, _, info = self.acks.get(timeout=max(0.1, deadline - time.monotonic()))
            except _queue.Empty:
                self.close()
                raise RuntimeError("GPU workers did not start in time")
            if status == "failed":
                self.close()
                raise RuntimeError(f"GPU worker {idx} failed to start: {info}")
            ready += status == "ready"
        self.decode_threads = decode_threads

    def submit(self, worker: int, task_id, model: str, files):
        """Hands a task to a worker; a task still running there is preempted (acknowledged as 'cancelled')."""
        self.inboxes[worker].put((task_id, model, list(files)))

    def get_ack(self, timeout=None):
        """(worker, task_id, status in {'ok', 'cancelled', 'error'}, seconds since the worker received the task, results count, results | message)."""
        return self.acks.get(timeout=timeout)

    def close(self):
        for q in self.inboxes:
            try:
                q.put(None)
            except Exception:
                pass
        for p in self.procs:
            p.join(timeout=20)
            if p.is_alive():
                p.terminate()
