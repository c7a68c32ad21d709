"""CPU tests of the host side around the predict path: result files (f2), the task datagram cap (f4), batch packing,
oversized images, checkpoint loading, and the coroutine's cancellation hand-shake."""
import asyncio
import importlib
import json
import os
import threading
import types

import numpy as np
import pytest
from PIL import Image

from conftest import GOLDEN, pkg

REF_OUTPUTS = ["reference_output_1_127.json", "reference_output_2_127.json"]  # /root/reference/download/output_{1,2}_127.json


def test_dump_to_file_reproduces_the_reference_sample_outputs_byte_for_byte(tmp_path):
    """/root/reference/models.py:123-126 wrote these two files; loading them and dumping them again through the
    drop-in's dump_to_file must give the same bytes (indent 4, float repr of the fp32 probability, trailing newline)."""
    m = pkg("models")
    for name in REF_OUTPUTS:
        raw = open(os.path.join(GOLDEN, name), "rb").read()
        d = json.loads(raw)
        # what predict returns: np.float32 probabilities (NpEncoder's path), tuples from decode_predictions
        d32 = {k: [[(t[0], t[1], np.float32(t[2])) for t in v[0]]] for k, v in d.items()}
        out = tmp_path / name
        m.dump_to_file(d32, str(out))
        assert out.read_bytes() == raw
        m.dump_to_file(d, str(out))  # and with plain floats
        assert out.read_bytes() == raw


def test_merge_all_json_files_like_the_client(tmp_path):
    """worker.py:1513-1534: skip missing files, dict.update in list order, dump_to_file into the same directory."""
    m = pkg("models")
    local = str(tmp_path) + "/"
    parts = []
    for i, name in enumerate(REF_OUTPUTS):
        d = json.load(open(os.path.join(GOLDEN, name)))
        parts.append(d)
        json.dump(d, open(local + f"output_7_{i}_fa22-cs425-6903.json", "w"))
    files = ["output_7_0_fa22-cs425-6903.json", "output_7_missing.json", "output_7_1_fa22-cs425-6903.json"]
    merged = m.merge_all_json_files(files, local, "final_7.json")
    want = {}
    want.update(parts[0])
    want.update(parts[1])
    assert merged == want and list(merged) == list(want)
    text = open(local + "final_7.json").read()
    assert text == json.dumps(want, indent=4) + "\n"
    # a later file wins on a duplicate image name
    dup = {next(iter(parts[0])): "Failed to download file from SDFS"}
    json.dump(dup, open(local + "output_7_2_x.json", "w"))
    merged = m.merge_all_json_files(files + ["output_7_2_x.json"], local, "final_7.json")
    assert merged[next(iter(parts[0]))] == "Failed to download file from SDFS"
    assert m.merge_all_json_files(["nope.json"], local, "final_8.json") is None and not os.path.exists(local + "final_8.json")


def test_task_datagram_cap():
    """packets.py:70-73: 32768 bytes of JSON per WORKER_TASK_REQUEST; ~120 images with 4 replicas each."""
    S = pkg("scheduler")
    names = [f"{i}.jpeg" for i in range(1, 401)]
    cap = S.max_images_per_task(names)
    assert 100 <= cap <= 140
    assert S.task_request_bytes(1, 1, "ResNet50", names[:cap]) <= S.PACKET_JSON_BYTES < S.task_request_bytes(1, 1, "ResNet50", names[:cap + 1])
    # the reference's own framing agrees: struct "i255s6si32768s"
    import struct
    assert struct.calcsize(f"i{255}s{6}si{2048 * 16}s") == 33040
    s = S.FairShareScheduler([f"H{i}" for i in range(3, 11)], batch_size=10)
    s.set_batch_size("ResNet50", 64, wire_cap=cap)       # C3 = 64 travels (65 images per chunk)
    with pytest.raises(ValueError):
        s.set_batch_size("ResNet50", 256, wire_cap=cap)  # 257 images per task would be truncated by Packet.pack
    s.set_batch_size("ResNet50", 256)                    # the reference itself accepts anything
    s.submit_job("ResNet50", 300, 1, names)  # 400 distinct names: the request's image map is keyed by name (worker.py:289-291)
    a = s.schedule()
    assert a and len(a[0].images) == 257 and not a[0].fits_datagram() and a[-1].wire_bytes() > 0
    s2 = S.FairShareScheduler([f"H{i}" for i in range(3, 11)], batch_size=64)
    s2.submit_job("InceptionV3", 300, 2, names[:100])
    assert all(x.fits_datagram() for x in s2.schedule())


def test_host_batch_into_preallocated_buffers(images):
    fe = pkg("frontend")
    sel = images[:7]
    a = fe.HostBatch(sel, 299)
    total = sum(im.size for im in sel)
    into = {"pixels": np.zeros(total + 64, np.uint8), "img_off": np.zeros(9, np.int64), "hw": np.zeros((9, 2), np.int32),
            "xtab": np.zeros((9, 299), np.int32), "ytab": np.zeros((9, 299), np.int32)}
    b = fe.HostBatch(sel, 299, into=into)
    assert b.n == a.n == 7
    for k in ("img_off", "hw", "xtab", "ytab"):
        assert np.array_equal(getattr(a, k), getattr(b, k))
    assert np.array_equal(a.pixels[:total], b.pixels[:total])
    assert b.pixels.base is into["pixels"] or np.shares_memory(b.pixels, into["pixels"])
    for i, im in enumerate(sel):
        assert np.array_equal(a.pixels[a.img_off[i]:a.img_off[i] + im.size], im.reshape(-1))
    with pytest.raises(ValueError):
        fe.HostBatch(images[:10], 299, into=into)
    assert fe.HostBatch([], 224).n == 0


def test_oversized_images_are_shrunk_exactly_like_load_img_and_batches_respect_the_byte_bound():
    fe = pkg("frontend")
    Engine = pkg("engine").Engine
    rng = np.random.default_rng(3)
    big = rng.integers(0, 256, size=(1080, 1920, 3), dtype=np.uint8)
    small = fe.shrink_oversized(big, 224, 300 * 300 * 3)
    want = np.asarray(Image.fromarray(big).resize((224, 224), Image.NEAREST))
    assert small.shape == (224, 224, 3) and np.array_equal(small, want)
    assert fe.shrink_oversized(big, 224, big.nbytes) is big
    # Engine.chunks needs no GPU: drive it on a stand-in with the fields it reads
    eng = types.SimpleNamespace(size=224, max_batch=8, max_pixels=8 * 300 * 300 * 3, max_image_bytes=2_000_000)
    imgs = [rng.integers(0, 256, size=(375, 500, 3), dtype=np.uint8) for _ in range(9)] + [big] + \
           [rng.integers(0, 256, size=(100, 120, 3), dtype=np.uint8) for _ in range(20)]
    runs = list(Engine.chunks(eng, imgs))
    assert sum(len(r) for r in runs) == len(imgs)
    for r in runs:
        assert len(r) <= 8 and sum(im.nbytes for im in r) <= eng.max_pixels
    flat = [im for r in runs for im in r]
    assert flat[9].shape == (224, 224, 3)            # the 1080p frame went through the host-side NEAREST tables
    assert all(a is b for a, b in zip(flat[10:], imgs[10:]))
    # 500x375 images are 562,500 bytes: only 3 fit a 8 x 270,000-byte staging buffer
    assert [len(r) for r in runs[:2]] == [3, 3]


def test_npz_checkpoints_load_in_natural_order_and_are_shape_checked(tmp_path):
    kw = pkg("keras_weights")
    lib = pkg("_lib")
    W = kw.generate("ResNet50", "folding_stress", 2)
    blob = kw.to_blob(W)
    np.savez(tmp_path / "pos.npz", *W)  # arr_0 ... arr_319: lexicographic order would scramble them
    assert np.array_equal(kw.to_blob(kw.arrays_from_npz("ResNet50", np.load(tmp_path / "pos.npz"))), blob)
    names = [n for n, _ in lib.weight_spec("ResNet50")]
    order = np.random.default_rng(0).permutation(len(W))
    np.savez(tmp_path / "named.npz", **{names[i].replace("/", "__"): W[i] for i in order})
    assert np.array_equal(kw.to_blob(kw.arrays_from_npz("ResNet50", np.load(tmp_path / "named.npz"))), blob)
    bad = list(W)
    bad[4], bad[5] = bad[5], bad[4].reshape(8, 8)
    np.savez(tmp_path / "bad.npz", *bad)
    with pytest.raises(lib.DmlError):
        kw.arrays_from_npz("ResNet50", np.load(tmp_path / "bad.npz"))
    np.savez(tmp_path / "short.npz", *W[:-1])
    with pytest.raises(lib.DmlError):
        kw.arrays_from_npz("ResNet50", np.load(tmp_path / "short.npz"))


def test_default_weights_and_placeholder_labels_warn(monkeypatch):
    kw = pkg("keras_weights")
    monkeypatch.delenv("DML_B200_WEIGHTS_DIR", raising=False)
    kw._warned.discard("InceptionV3")
    with pytest.warns(RuntimeWarning, match="RANDOM"):
        kw.load_blob("InceptionV3")


def test_cancelled_coroutine_stops_its_thread_at_a_batch_boundary(monkeypatch):
    """worker.py:944-953 cancels the running task when a new WORKER_TASK_REQUEST arrives.  The coroutine hands the
    cancellation to its worker thread (which stops at the next batch boundary) and waits for it to leave."""
    m = pkg("models")
    lib = pkg("_lib")
    seen = {}

    def fake_run(files, modelObj=None, *, cancel=None):
        seen["cancel"] = cancel
        seen["started"].set()
        for _ in range(500):  # "batches"
            if cancel.event.is_set():
                seen["stopped_early"] = True
                raise lib.DmlCancelled("cancelled between two batches")
            threading.Event().wait(0.01)
        return {"never": 1}

    monkeypatch.setattr(m, "run_inference_on_ResNet50", fake_run)

    async def scenario():
        seen["started"] = threading.Event()
        task = asyncio.create_task(m.perform_inference("ResNet50", ["a.jpeg"]))
        while not seen["started"].is_set():
            await asyncio.sleep(0.005)
        task.cancel()
        with pytest.raises(asyncio.CancelledError):
            await task

    asyncio.run(scenario())
    assert seen.get("stopped_early") is True
    assert asyncio.run(m.perform_inference("VGG16", [])) == "Invalid model"
