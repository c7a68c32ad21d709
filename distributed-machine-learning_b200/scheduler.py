"""The leader's job chunker and fair-share scheduler, restated as a pure state machine.

This is the caller side of the predict path: in the reference the
leader splits a job into chunks and hands them to its 8 worker VMs
(/root/reference/worker.py:188-245 chunking, :255-495 schedule_job, :989-1026 ACK
bookkeeping, :1028-1037 SET_BATCH_SIZE, :1279-1306 requeue on failure).  On one B200 box the
8 workers are the 8 GPUs, so the same decisions map chunks to per-GPU queues.  Nothing here
touches the network or a GPU: methods take the current state and return assignments, which is
what lets the benchmark harness (and the tests) drive it deterministically.

Kept quirks, because they are the reference's behaviour:
  * a chunk holds batch_size + 1 images (worker.py:216-227); ``fix_off_by_one=True`` removes it;
  * SET_BATCH_SIZE recomputes the cost model with the InceptionV3 constants for either model
    (worker.py:1035);
  * free workers are taken from the END of the free list and running workers of the other model
    are preempted from the end of their list (worker.py:331-372);
  * with both queues non-empty the split considered is (w-1, 1), (w-2, 2), ... (1, w-1) for w
    online workers and the one minimising |rate_I - rate_R| / max is chosen (worker.py:303-324).
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field

from .models import ModelParameters

MODELS = ("InceptionV3", "ResNet50")

# worker.py:61, :74 -- the CPU-era cost model; replace through set_cost_model() with measured values
REFERENCE_COST = {
    "InceptionV3": dict(download_time=1, model_load_time=5.6, first_image_predict_time=2, each_image_predict_time=0.325),
    "ResNet50": dict(download_time=1, model_load_time=3.5, first_image_predict_time=1, each_image_predict_time=0.250),
}


# The reference ships every task to its worker as ONE fixed-size datagram whose JSON payload field is
# 2048 * 16 = 32768 bytes (struct "i255s6si32768s", /root/reference/packets.py:70-73).  pack() truncates a longer
# payload silently, unpack() then fails in json.loads and returns None (:79-92): the task is dropped, never ACKed,
# and its job never completes.  Each image costs its name plus its replica map {machine: [versions]}
# (worker.py:289-297) -- about 265 bytes with 4 replicas -- so a WORKER_TASK_REQUEST holds ~120 images at most and
# C3 batch sizes above that cannot travel through the real control plane, whatever the GPU could take.
PACKET_JSON_BYTES = 2048 * 16


def task_request_bytes(job_id, batch_id, model, images, replicas=None):
    """Bytes of the JSON payload the leader would put into the WORKER_TASK_REQUEST for this chunk
    (worker.py:297: {"jobid", "batchid", "model", "images": {name: {machine: [versions]}}}).
    replicas: {image: {machine: [versions]}}; default = 4 replicas with one version each, named like the reference's
    unique node names and versioned files."""
    def default(image):
        return {f"fa22-cs425-69{m:02d}.cs.illinois.edu:8000": [f"{image}_version1"] for m in (3, 4, 5, 6)}
    payload = {"jobid": job_id, "batchid": batch_id, "model": model,
               "images": {im: (replicas[im] if replicas is not None else default(im)) for im in images}}
    return len(json.dumps(payload).encode())


def max_images_per_task(sample_images, model="InceptionV3", replicas=None):
    """Largest chunk of these image names whose task request still fits one datagram."""
    n = 0
    while n < len(sample_images) and task_request_bytes(10 ** 6, 10 ** 6, model, sample_images[:n + 1], replicas) <= PACKET_JSON_BYTES:
        n += 1
    return n


@dataclass
class Assignment:
    worker: str
    model: str
    job_id: int
    batch_id: int
    images: list
    preempted: dict | None = None  # the chunk that was taken off this worker, if any

    def wire_bytes(self, replicas=None) -> int:
        return task_request_bytes(self.job_id, self.batch_id, self.model, self.images, replicas)

    def fits_datagram(self, replicas=None) -> bool:
        """False: the reference's Packet.pack would truncate this task request (packets.py:70-73) and the worker
        would never see it."""
        return self.wire_bytes(replicas) <= PACKET_JSON_BYTES


@dataclass
class _ModelState:
    batch_size: int
    time: float
    queue: list = field(default_factory=list)
    inprogress_queue: list = field(default_factory=list)


class FairShareScheduler:
    def __init__(self, workers, batch_size: int = 10, fix_off_by_one: bool = False):
        self.worker_nodes = list(workers)            # worker.py:52 (H3..H10) -> GPU 0..7
        self.fix_off_by_one = fix_off_by_one
        self.cost = {m: dict(REFERENCE_COST[m]) for m in MODELS}
        self.model_dict = {m: _ModelState(batch_size, self._time(m, batch_size)) for m in MODELS}
        self.workers_tasks_dict = {}                 # worker -> {'model', 'job_id', 'batch_id'}
        self.job_requester_dict = {}                 # job_id -> batches pending

    # -- cost model ---------------------------------------------------------------
    def _time(self, model, batch_size):
        return ModelParameters(batch_size=batch_size, **self.cost[model]).execution_time_per_vm()

    def set_cost_model(self, model, **kw):
        """Replace the CPU-era constants (e.g. with measured B200 values)."""
        self.cost[model].update(kw)
        self.model_dict[model].time = self._time(model, self.model_dict[model].batch_size)

    def set_batch_size(self, model, batch_size, reference_quirk: bool = True, wire_cap: int | None = None):
        """C3 / SET_BATCH_SIZE (worker.py:1028-1037).  The reference accepts any value; ``wire_cap`` (see
        max_images_per_task) refuses one whose chunks (batch_size + 1 images with the off-by-one) would not fit a
        WORKER_TASK_REQUEST datagram, instead of losing those tasks silently."""
        if wire_cap is not None and int(batch_size) + (0 if self.fix_off_by_one else 1) > wire_cap:
            raise ValueError(f"batch size {batch_size}: a task of {int(batch_size) + (0 if self.fix_off_by_one else 1)} images "
                             f"does not fit the {PACKET_JSON_BYTES}-byte task datagram (at most {wire_cap} images)")
        st = self.model_dict[model]
        st.batch_size = int(batch_size)
        st.time = ModelParameters(batch_size=st.batch_size, **REFERENCE_COST["InceptionV3"]).execution_time_per_vm() \
            if reference_quirk else self._time(model, st.batch_size)

    # -- job intake (worker.py:188-245) ---------------------------------------------
    def submit_job(self, model, number_of_images, job_id, sdfs_images):
        sdfs_images = sorted(sdfs_images)
        if not sdfs_images:
            return []
        images, index = [], 0
        while len(images) < number_of_images:      # round-robin over the sorted file list
            if index >= len(sdfs_images):
                index = 0
            images.append(sdfs_images[index])
            index += 1
        batch_size = self.model_dict[model].batch_size
        limit = batch_size - 1 if self.fix_off_by_one else batch_size
        all_batches, batch = [], []
        for image in images:
            batch.append(image)
            if len(batch) > limit:
                all_batches.append({"job_id": job_id, "batch_id": len(all_batches) + 1, "images": batch})
                batch = []
        if batch:
            all_batches.append({"job_id": job_id, "batch_id": len(all_batches) + 1, "images": batch})
        self.model_dict[model].queue.extend(all_batches)
        self.job_requester_dict[job_id] = len(all_batches)
        return all_batches

    # -- helpers ----------------------------------------------------------------------
    def get_running_nodes(self, model):
        return [w for w, t in self.workers_tasks_dict.items() if t["model"] == model]

    def _free_workers(self, online):
        return [w for w in self.worker_nodes if w in online and w not in self.workers_tasks_dict]

    def _assign(self, worker, model):
        st = self.model_dict[model]
        chunk = st.queue[0]
        preempted = None
        if worker in self.workers_tasks_dict:        # preempt what it is running (worker.py:389-408)
            old = self.workers_tasks_dict[worker]
            old_st = self.model_dict[old["model"]]
            for i, item in enumerate(old_st.inprogress_queue):
                if item["job_id"] == old["job_id"] and item["batch_id"] == old["batch_id"]:
                    preempted = old_st.inprogress_queue.pop(i)
                    old_st.queue.insert(0, preempted)
                    break
        self.workers_tasks_dict[worker] = {"model": model, "job_id": chunk["job_id"], "batch_id": chunk["batch_id"]}
        st.inprogress_queue.append(chunk)
        st.queue.pop(0)
        return Assignment(worker, model, chunk["job_id"], chunk["batch_id"], chunk["images"], preempted)

    def predicted_split(self, online_worker_count):
        """(inception_workers, resnet_workers) minimising the relative query-rate difference."""
        splits, w, t = [], online_worker_count, 1
        while w > 1:
            splits.append((w - 1, t))
            w -= 1
            t += 1
        if not splits:
            return None
        inc, res = self.model_dict["InceptionV3"], self.model_dict["ResNet50"]
        diffs = []
        for vi, vr in splits:
            ri, rr = vi * inc.batch_size / inc.time, vr * res.batch_size / res.time
            diffs.append(abs(ri - rr) / max(ri, rr) * 100)
        return splits[diffs.index(min(diffs))]

    # -- schedule_job (worker.py:255-495) -----------------------------------------------
    def schedule(self, online_workers=None):
        """Returns the list of Assignments the leader would send as WORKER_TASK_REQUESTs now."""
        online = set(self.worker_nodes if online_workers is None else online_workers)
        qi, qr = self.model_dict["InceptionV3"].queue, self.model_dict["ResNet50"].queue
        out = []
        if bool(qi) != bool(qr):                      # exactly one model has work: every free worker takes a chunk
            model = "InceptionV3" if qi else "ResNet50"
            for worker in self._free_workers(online):
                if not self.model_dict[model].queue:
                    break
                out.append(self._assign(worker, model))
            return out
        if not qi:
            return out
        split = self.predicted_split(len(online & set(self.worker_nodes)))
        if split is None:
            # one worker online: the reference takes min() of an empty list here and raises (worker.py:323);
            # a single GPU simply alternates, serving the longer queue first
            model = "InceptionV3" if len(qi) >= len(qr) else "ResNet50"
            for worker in self._free_workers(online):
                if self.model_dict[model].queue:
                    out.append(self._assign(worker, model))
            return out
        want = {"InceptionV3": split[0], "ResNet50": split[1]}
        running = {m: self.get_running_nodes(m) for m in MODELS}
        free = self._free_workers(online)
        new_nodes = {}
        for model, other in (("InceptionV3", "ResNet50"), ("ResNet50", "InceptionV3")):
            picked = []
            if len(running[model]) < want[model]:
                while free and len(picked) + len(running[model]) < want[model]:
                    picked.append(free.pop())          # from the end of the free list
                while running[other] and len(picked) + len(running[model]) < want[model]:
                    picked.append(running[other].pop())  # preempt the other model's last worker
            new_nodes[model] = picked
        for model in ("ResNet50", "InceptionV3"):     # the reference dispatches ResNet50 first
            for worker in new_nodes[model]:
                if not self.model_dict[model].queue:
                    break
                out.append(self._assign(worker, model))
        return out

    # -- WORKER_TASK_REQUEST_ACK (worker.py:989-1026) --------------------------------------
    def task_done(self, worker):
        """Returns True when this ACK completed its job."""
        t = self.workers_tasks_dict.pop(worker)
        st = self.model_dict[t["model"]]
        for i, item in enumerate(st.inprogress_queue):
            if item["job_id"] == t["job_id"] and item["batch_id"] == t["batch_id"]:
                st.inprogress_queue.pop(i)
                break
        self.job_requester_dict[t["job_id"]] -= 1
        return self.job_requester_dict[t["job_id"]] == 0

    # -- worker failure (worker.py:1279-1306) ------------------------------------------------
    def worker_failed(self, worker):
        """The dead worker's chunk goes back to the FRONT of its model's queue."""
        if worker not in self.workers_tasks_dict:
            return None
        t = self.workers_tasks_dict.pop(worker)
        st = self.model_dict[t["model"]]
        for i, item in enumerate(st.inprogress_queue):
            if item["job_id"] == t["job_id"] and item["batch_id"] == t["batch_id"]:
                chunk = st.inprogress_queue.pop(i)
                st.queue.insert(0, chunk)
                return chunk
        return None

    def idle(self):
        return not self.workers_tasks_dict and all(not self.model_dict[m].queue for m in MODELS)
