"""CPU tests of the restated chunker + fair-share scheduler against the
behaviour written down in /root/reference/worker.py and the operator notes in test.py."""
import pytest

from conftest import pkg

W = [f"H{i}" for i in range(3, 11)]  # the reference's 8 workers (worker.py:52) = GPU 0..7


def files(n):
    return [f"{i}.jpeg" for i in range(1, n + 1)]


def test_chunking_keeps_the_reference_off_by_one():
    S = pkg("scheduler").FairShareScheduler
    s = S(W, batch_size=10)
    chunks = s.submit_job("ResNet50", 25, 1, files(100))
    assert [len(c["images"]) for c in chunks] == [11, 11, 3]          # worker.py:216-227: batch_size + 1 per chunk
    assert [c["batch_id"] for c in chunks] == [1, 2, 3]
    # round-robin over the SORTED names (string order, as sdfs_images.sort() gives)
    assert chunks[0]["images"][:3] == ["1.jpeg", "10.jpeg", "100.jpeg"]
    s2 = S(W, batch_size=10, fix_off_by_one=True)
    assert [len(c["images"]) for c in s2.submit_job("ResNet50", 25, 1, files(100))] == [10, 10, 5]
    s3 = S(W, batch_size=64)
    assert sum(len(c["images"]) for c in s3.submit_job("InceptionV3", 300, 7, files(100))) == 300  # wraps around the list
    assert s3.submit_job("InceptionV3", 5, 8, []) == []


def test_fair_share_split_from_the_reference_cost_model():
    """worker.py:303-324 with the constants of worker.py:61/:74 (batch 10): rates 10*v/t, closest pair wins.
    (The 5/3 split in the operator log at test.py:133-134 was taken with other, unrecorded constants.)"""
    s = pkg("scheduler").FairShareScheduler(W, batch_size=10)
    assert s.model_dict["InceptionV3"].time == pytest.approx(10 + 5.6 + 2 + 0.325 * 9)
    assert s.model_dict["ResNet50"].time == pytest.approx(10 + 3.5 + 1 + 0.25 * 9)
    assert s.predicted_split(8) == (4, 4)
    assert s.predicted_split(3) == (1, 1) or s.predicted_split(3) == (2, 1)
    assert s.predicted_split(1) is None
    # a much faster ResNet50 shifts workers to InceptionV3
    s.set_cost_model("ResNet50", each_image_predict_time=0.01, model_load_time=0, first_image_predict_time=0)
    assert s.predicted_split(8)[0] > 4


def test_single_model_fills_every_free_worker_then_acks_free_them():
    s = pkg("scheduler").FairShareScheduler(W, batch_size=8)
    s.submit_job("ResNet50", 100, 1, files(100))       # 12 chunks of 9 (the last one 1)
    a = s.schedule()
    assert [x.worker for x in a] == W and all(x.model == "ResNet50" and x.preempted is None for x in a)
    assert [x.batch_id for x in a] == list(range(1, 9))
    assert s.schedule() == []                            # everybody busy
    assert s.task_done("H5") is False
    b = s.schedule()
    assert len(b) == 1 and b[0].worker == "H5" and b[0].batch_id == 9
    done = [s.task_done(x) for x in list(s.workers_tasks_dict)]
    while not s.idle():
        for x in s.schedule():
            done.append(s.task_done(x.worker))
    assert done.count(True) == 1 and s.job_requester_dict[1] == 0


def test_second_model_preempts_down_to_the_fair_share():
    s = pkg("scheduler").FairShareScheduler(W, batch_size=10)
    s.submit_job("ResNet50", 200, 1, files(100))
    assert len(s.schedule()) == 8                         # ResNet50 alone takes all 8 workers
    s.submit_job("InceptionV3", 200, 2, files(100))
    a = s.schedule()                                      # fair share 4 / 4: four ResNet50 chunks are preempted
    assert len(a) == 4 and all(x.model == "InceptionV3" and x.preempted is not None for x in a)
    assert [x.worker for x in a] == W[::-1][:4]           # taken from the end of the running list (worker.py:344-349)
    assert len(s.get_running_nodes("InceptionV3")) == 4 and len(s.get_running_nodes("ResNet50")) == 4
    # preempted chunks went back to the FRONT of the ResNet50 queue, last preempted first
    assert [c["batch_id"] for c in s.model_dict["ResNet50"].queue[:4]] == [5, 6, 7, 8]
    assert all(c["job_id"] == 1 for c in s.model_dict["ResNet50"].queue)


def test_failure_requeues_at_the_front_and_c3_quirk():
    s = pkg("scheduler").FairShareScheduler(W, batch_size=10)
    s.submit_job("InceptionV3", 50, 3, files(100))
    a = s.schedule()
    lost = s.worker_failed(a[1].worker)
    assert lost["batch_id"] == a[1].batch_id and s.model_dict["InceptionV3"].queue[0] is lost
    assert s.worker_failed("H99") is None
    b = s.schedule(online_workers=[w for w in W if w != a[1].worker])
    assert all(x.worker != a[1].worker for x in b) and b[0].batch_id == lost["batch_id"]
    # C3 always recomputes the cost with the InceptionV3 constants (worker.py:1035)
    s.set_batch_size("ResNet50", 64)
    assert s.model_dict["ResNet50"].time == pytest.approx(64 + 5.6 + 2 + 0.325 * 63)
    s.set_batch_size("ResNet50", 64, reference_quirk=False)
    assert s.model_dict["ResNet50"].time == pytest.approx(64 + 3.5 + 1 + 0.25 * 63)
    s.set_cost_model("ResNet50", download_time=0, model_load_time=0, first_image_predict_time=0, each_image_predict_time=2e-5)
    assert s.model_dict["ResNet50"].time == pytest.approx(2e-5 * 63)


def test_mixed_queues_always_drain(tmp_path):
    """Whatever the worker count, a mixed job set runs to completion (one worker: alternate models)."""
    S = pkg("scheduler").FairShareScheduler
    for nworkers in (1, 2, 3, 8):
        s = S(W[:nworkers], batch_size=8)
        s.submit_job("ResNet50", 60, 1, files(100))
        s.submit_job("InceptionV3", 45, 2, files(100))
        steps = 0
        finished = 0
        while not s.idle():
            a = s.schedule()
            running = list(s.workers_tasks_dict)
            assert running, "nothing running and nothing scheduled: the leader would stall"
            finished += s.task_done(running[0])
            steps += 1
            assert steps < 1000
        assert finished == 2
