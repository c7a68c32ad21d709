"""world_size-2 gloo tests (CPU) of the multi-rank host logic: slicing, weight-blob
replication and result merging.  The kernels themselves need no collective."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG, ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import importlib
    import sys
    sys.path.insert(0, ROOT)
    par = importlib.import_module(PKG + ".parallel")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # weight replica: rank 0's blob reaches everyone
        blob = torch.arange(1000, dtype=torch.int32).to(torch.uint8) if rank == 0 else torch.zeros(1000, dtype=torch.uint8)
        par.replicate_weights(blob, src=0)
        assert torch.equal(blob, torch.arange(1000, dtype=torch.int32).to(torch.uint8))
        # sharded job: every rank ends with the full, ordered result
        files = [f"/sdfs/{i}.jpeg" for i in range(1, 12)]
        seen = []

        def fake_predict(fs):
            seen.extend(fs)
            return {f.split("/")[-1]: [[("n0", "class", float(rank))]] for f in fs}

        res = par.run_sharded(fake_predict, files)
        a, b = par.shard_range(len(files), world, rank)
        assert seen == files[a:b]
        assert list(res) == [f"{i}.jpeg" for i in range(1, 12)]
        assert [v[0][0][2] for v in res.values()] == [0.0] * 6 + [1.0] * 5
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_shard_range_covers_everything():
    import importlib
    par = importlib.import_module(PKG + ".parallel")
    for n in (0, 1, 7, 8, 128, 129):
        for world in (1, 2, 4, 8):
            spans = [par.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        par.shard_range(4, 2, 2)
