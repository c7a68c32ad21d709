"""The branch-stream schedule (csrc/graph.cc: plan_streams) checked independently on the CPU.

The Keras functional graph of InceptionV3 (/root/reference/models.py:26) has independent branches inside every mixed block; the
engine launches them on different CUDA streams.  Here the schedule is dumped by a small host program (tests/native/
schedule_dump.cc, graph.cc only, no CUDA) and this file re-derives every memory hazard from the dumped accesses -- same buffer
and overlapping channels, or two buffers the arena placed on overlapping addresses -- and checks that stream order plus the
event waits order every hazardous pair.  ResNet50's graph is a chain: it must stay on one stream.
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "distributed-machine-learning_b200", "csrc")


@pytest.fixture(scope="module")
def dump_bin(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("sched") / "schedule_dump")
    cmd = ["g++", "-O1", "-std=c++17", "-I", CSRC, "-o", out, os.path.join(ROOT, "tests", "native", "schedule_dump.cc"),
           os.path.join(CSRC, "graph.cc")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def _parse(text):
    lines = text.strip().splitlines()
    head = lines[0].split()
    info = {"streams": int(head[1]), "regions": int(head[3])}
    ops = []
    for ln in lines[1:]:
        f = ln.split()
        kv = dict(x.split("=") for x in f[3:9])
        r_at, w_at = f.index("R"), f.index("W")

        def accs(items):
            return [tuple(int(v) for v in it.split(":")) for it in items]
        ops.append(dict(i=int(f[1]), name=f[2], skip=kv["skip"] == "1", stream=int(kv["stream"]), region=int(kv["region"]),
                        solo=kv["solo"] == "1", signal=kv["signal"] == "1",
                        waits=[int(w) for w in kv["waits"].split(",") if w], reads=accs(f[r_at + 1:w_at]), writes=accs(f[w_at + 1:])))
    return info, ops


def _conflict(x, y):
    bx, c0x, c1x, a0x, a1x = x
    by, c0y, c1y, a0y, a1y = y
    if bx < 0 or by < 0:
        return bx == by
    if bx == by:
        return c0x < c1y and c0y < c1x
    return a0x < a1y and a0y < a1x


def _run(dump_bin, model, streams):
    r = subprocess.run([dump_bin, str(model), str(streams)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return _parse(r.stdout)


@pytest.mark.parametrize("streams", [2, 3, 4, 8])
def test_inception_schedule_orders_every_hazard(dump_bin, streams):
    info, ops = _run(dump_bin, 1, streams)
    ex = [o for o in ops if not o["skip"]]
    assert info["regions"] == 11 and 2 <= info["streams"] <= streams
    # happens-before: previous op of the same stream + the waited-for ops, closed transitively
    before = {}
    last_on = {}
    for o in ex:
        s = set()
        preds = list(o["waits"])
        if o["stream"] in last_on:
            preds.append(last_on[o["stream"]])
        for p in preds:
            s.add(p)
            s |= before[p]
        before[o["i"]] = s
        last_on[o["stream"]] = o["i"]
        for w in o["waits"]:
            assert ops[w]["signal"] and ops[w]["stream"] != o["stream"] and w < o["i"]
    hazards = 0
    for a, oi in enumerate(ex):
        for oj in ex[:a]:
            hz = any(_conflict(w, r) for w in oj["writes"] for r in oi["reads"]) or \
                 any(_conflict(w, w2) for w in oj["writes"] for w2 in oi["writes"]) or \
                 any(_conflict(r, w2) for r in oj["reads"] for w2 in oi["writes"])
            if hz:
                hazards += 1
                assert oj["i"] in before[oi["i"]], (oj["name"], oi["name"])
    assert hazards > 200
    # every block really forks: its non-solo ops sit on at least two streams
    for r in range(info["regions"]):
        par = {o["stream"] for o in ex if o["region"] == r and not o["solo"]}
        assert len(par) >= 2, r
    # outside the blocks nothing is concurrent
    assert all(o["solo"] for o in ex if o["region"] < 0)


def test_followers_are_covered_by_their_leader(dump_bin):
    """A skipped op (computed inside its leader's merged GEMM) writes only inside the leader's write range."""
    _, ops = _run(dump_bin, 1, 4)
    for o in ops:
        if o["skip"]:
            heads = [h for h in ops if not h["skip"] and h["region"] == o["region"] and h["writes"][0][0] == o["writes"][0][0]
                     and h["writes"][0][1] <= o["writes"][0][1] and o["writes"][0][2] <= h["writes"][0][2] and h["i"] < o["i"]]
            assert heads, o["name"]


def test_resnet50_stays_serial(dump_bin):
    info, ops = _run(dump_bin, 0, 4)
    assert info["streams"] == 1 and info["regions"] == 0
    assert all(o["stream"] == 0 and not o["waits"] for o in ops)
