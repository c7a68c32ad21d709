"""Drop-in for the reference's ``models`` module (/root/reference/models.py).

Same names, signatures and result shapes as what /root/reference/worker.py:24
imports and calls (perform_inference at worker.py:1358/:1379,
perform_inference_without_async at :1452, dump_to_file at :1532/:1581/:1611,
ModelParameters at :61/:74/:1035), but predict runs on the B200 engine: the
model stays resident per (process, GPU) instead of being rebuilt per task, and
all files of a task are batched on the GPU.

Result contract (models.py:44, :69): ``{basename(path): [[(wnid, name, prob) x 5]]}``
in the order of ``files``; a file that cannot be read or decoded gets a string
entry instead of raising.
"""
from __future__ import annotations

import asyncio
import json
import os
import threading
from concurrent.futures import ThreadPoolExecutor
from functools import partial

import numpy as np

from . import frontend
from ._lib import DmlCancelled
from .engine import Engine

_HERE = os.path.dirname(os.path.abspath(__file__))
_engines = {}
_lock = threading.Lock()
_decode_pool = None
_class_index = None

# models.py:18-21 keeps these module globals (always None in the reference).
model1 = None
model2 = None

MAX_BATCH = int(os.environ.get("DML_B200_MAX_BATCH", "256"))


def class_index():
    """{int: (wnid, name)} -- Keras' imagenet_class_index.json (decode_predictions, models.py:42)."""
    global _class_index
    if _class_index is None:
        path = os.environ.get("DML_B200_CLASS_INDEX") or os.path.join(_HERE, "imagenet_class_index.json")
        with open(path) as fh:
            raw = json.load(fh)
        _class_index = {int(k): (v[0], v[1]) for k, v in raw.items()}
        missing = sum(1 for v in _class_index.values() if "?" in v[0])
        if missing:
            import warnings
            warnings.warn(f"dml_b200: {missing} of {len(_class_index)} wnids in {path} are placeholders (Keras downloads "
                          "imagenet_class_index.json from the network); set DML_B200_CLASS_INDEX to the real file",
                          RuntimeWarning, stacklevel=2)
    return _class_index


def get_engine(model_name: str, device: int | None = None) -> Engine:
    """The resident engine for (model, device); built on first use."""
    if device is None:
        device = int(os.environ.get("DML_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    key = (model_name, device)
    with _lock:
        if key not in _engines:
            _engines[key] = Engine(model_name, device=device, max_batch=MAX_BATCH)
        return _engines[key]


def _pool():
    global _decode_pool
    if _decode_pool is None:
        n = int(os.environ.get("DML_B200_DECODE_THREADS", "0")) or min(32, os.cpu_count() or 4)
        _decode_pool = ThreadPoolExecutor(max_workers=n)  # Pillow releases the GIL while it decodes
    return _decode_pool


def _decode_one(path):
    try:
        return frontend.decode_file(path)
    except Exception as exc:  # unreadable / corrupt file: reported per image, never raised
        return f"Failed to decode image: {type(exc).__name__}: {exc}"


def decode_top5(idx: np.ndarray, prob: np.ndarray):
    """Engine top-5 -> what decode_predictions(pred, top=5) returns for one row."""
    ci = class_index()
    return [(ci[int(i)][0], ci[int(i)][1], np.float32(p)) for i, p in zip(idx, prob)]


class _Cancel:
    """Hand-shake between a cancelled perform_inference coroutine and its worker thread."""

    def __init__(self):
        self.event = threading.Event()
        self.lock = threading.Lock()
        self.engine = None  # set while the thread owns the engine

    def request(self):
        with self.lock:
            self.event.set()
            if self.engine is not None:
                self.engine.cancel()  # dml_cancel: the dml_infer in progress stops before its next micro-batch


def _run(model_name: str, image_files: list, modelObj=None, cancel: _Cancel | None = None):
    eng = modelObj if isinstance(modelObj, Engine) else get_engine(model_name)
    files = list(image_files)
    failed = {}
    good = []

    def decoded_images():
        # the pool decodes ahead while the GPU works on the batches already handed over
        for i, d in enumerate(_pool().map(_decode_one, files)):
            if isinstance(d, str):
                failed[i] = d
            else:
                good.append(i)
                yield d

    with eng.lock:  # one engine is used by one thread at a time
        if cancel is not None:
            with cancel.lock:
                if cancel.event.is_set():
                    raise DmlCancelled("cancelled before the first batch")
                cancel.engine = eng
        try:
            idx, prob = eng.predict_pipelined(decoded_images(), cancel.event if cancel is not None else None)
        finally:
            if cancel is not None:
                with cancel.lock:
                    cancel.engine = None
                    if cancel.event.is_set():
                        eng.l.dml_cancel_clear(eng.h)  # a flag raised after the last micro-batch must not hit the next task
    results = {}
    pos = {g: k for k, g in enumerate(good)}
    for i, f in enumerate(files):
        key = f.split("/")[-1]
        if i in pos:
            k = pos[i]
            results[key] = [decode_top5(idx[k], prob[k])]  # extra outer list: decode_predictions' batch dim
        else:
            results[key] = failed[i]
    return results


def run_inference_on_InceptionV3(image_files: list, modelObj=None, *, cancel=None):
    """models.py:23-46."""
    return _run("InceptionV3", image_files, modelObj, cancel)


def run_inference_on_ResNet50(image_files: list, modelObj=None, *, cancel=None):
    """models.py:48-71."""
    return _run("ResNet50", image_files, modelObj, cancel)


async def perform_inference(model_name, files):
    """models.py:74-91.  Stays a coroutine; the blocking decode + GPU call runs in
    a worker thread (ctypes releases the GIL) instead of a fresh process pool.
    Cancelling the awaiting task (the worker does that when a new WORKER_TASK_REQUEST
    preempts the running one, worker.py:944-953) stops the thread at its next batch
    boundary -- the reference's pool, by contrast, runs the cancelled task to the end."""
    if model_name == "InceptionV3":
        function = run_inference_on_InceptionV3
    elif model_name == "ResNet50":
        function = run_inference_on_ResNet50
    else:
        return "Invalid model"
    loop = asyncio.get_running_loop()
    cancel = _Cancel()
    fut = loop.run_in_executor(None, partial(function, files, cancel=cancel))
    try:
        return await asyncio.shield(fut)
    except asyncio.CancelledError:
        cancel.request()
        try:
            await fut  # the thread leaves at its next batch boundary; nothing of this task is left running
        except (DmlCancelled, asyncio.CancelledError):
            pass
        raise


def perform_inference_without_async(model_name, files):
    """models.py:93-106."""
    if model_name == "InceptionV3":
        return run_inference_on_InceptionV3(files, modelObj=model1)
    elif model_name == "ResNet50":
        return run_inference_on_ResNet50(files, modelObj=model2)
    else:
        return "Invalid model"


class NpEncoder(json.JSONEncoder):
    """models.py:109-118."""

    def default(self, obj):
        if isinstance(obj, np.integer):
            return int(obj)
        if isinstance(obj, np.floating):
            return float(obj)
        if isinstance(obj, np.ndarray):
            return obj.tolist()
        return json.JSONEncoder.default(self, obj)


def Merge(dict1, dict2):
    """models.py:120-121 (returns None, like dict.update)."""
    return dict2.update(dict1)


def dump_to_file(d, filename):
    """models.py:123-126: indent=4 JSON plus print's trailing newline."""
    with open(filename, "w") as fout:
        json_dumps_str = json.dumps(d, indent=4, cls=NpEncoder)
        print(json_dumps_str, file=fout)


def merge_all_json_files(all_files, local_dir, output_file):
    """The client's merge of a job's per-task result files into ``final_{jobid}.json``
    (/root/reference/worker.py:1513-1534, called from get_output_cli :1617-1627): files that are missing from
    ``local_dir`` are skipped, the dicts are merged with ``dict.update`` in list order (a later file wins on a
    duplicate image name) and the result is written with dump_to_file.  ``local_dir`` is used as a plain string prefix,
    as the reference does.  Returns the merged dict, or None when no file was there (nothing is written then)."""
    available_files = [f for f in all_files if os.path.exists(local_dir + f)]
    if len(available_files) == 0:
        print("requested files are not available in provided directory")
        return None
    final_output: dict = {}
    for f in available_files:
        with open(local_dir + f, encoding="utf-8") as f1:
            final_output.update(json.loads(f1.read()))
    dump_to_file(final_output, local_dir + output_file)
    return final_output


class ModelParameters:
    """models.py:128-139: the leader's per-VM task cost model."""

    def __init__(self, download_time, model_load_time, first_image_predict_time, each_image_predict_time,
                 batch_size) -> None:
        self.download_time = download_time
        self.model_load_time = model_load_time
        self.first_image_predict_time = first_image_predict_time
        self.each_image_predict_time = each_image_predict_time
        self.batch_size = batch_size

    def execution_time_per_vm(self):
        return (self.download_time * self.batch_size) + self.model_load_time + self.first_image_predict_time + \
            (self.each_image_predict_time * (self.batch_size - 1))
