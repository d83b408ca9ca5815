"""N>1 path on CPU: world_size-2 gloo run of bench.py's file sharding and timing reduction.  The
data path has no collective (files share nothing, recode.cpp:1314,1597); ranks take disjoint clip
ranges, the only exchange is max(time) / sum(bytes, bins)."""
import os
import sys
import tempfile

import pytest

from conftest import ROOT

sys.path.insert(0, ROOT)


def _worker(rank, world, port, clips, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "avrecode-ms_b200"))
    import bench
    import avr_ctypes as A
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    with tempfile.TemporaryDirectory() as td:
        blobs, _ = bench.gen_corpus(td, rank * clips, clips)
    recs = [A.Record(b) for b in blobs]
    assert all(r.rc == 0 for r in recs)
    my = dict(mb=sum(len(b) for b in blobs) / 1e6, bins=float(sum(r.bins for r in recs)), t=0.1 * (rank + 1))
    t = torch.tensor([my["t"]], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    s = torch.tensor([my["mb"], my["bins"]], dtype=torch.float64)
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    seeds = torch.tensor([bench.clip_params(rank * clips + i)["seed"] for i in range(clips)])
    allseeds = [torch.zeros_like(seeds) for _ in range(world)]
    dist.all_gather(allseeds, seeds)
    if rank == 0:
        import json
        json.dump(dict(t=t.item(), mb=s[0].item(), bins=s[1].item(), seeds=torch.cat(allseeds).tolist()),
                  open(os.path.join(out_dir, "r.json"), "w"))
    dist.destroy_process_group()


def test_two_rank_sharding_is_disjoint_and_aggregates():
    import json
    import torch.multiprocessing as mp
    import bench
    import avr_ctypes as A
    clips = 3
    with tempfile.TemporaryDirectory() as out:
        mp.spawn(_worker, args=(2, 29731, clips, out), nprocs=2, join=True)
        r = json.load(open(os.path.join(out, "r.json")))
    assert sorted(r["seeds"]) == [1000 + i for i in range(2 * clips)]          # disjoint, complete
    assert r["t"] == pytest.approx(0.2)                                         # max over ranks
    with tempfile.TemporaryDirectory() as td:
        blobs, _ = bench.gen_corpus(td, 0, 2 * clips)
    recs = [A.Record(b) for b in blobs]
    assert r["mb"] == pytest.approx(sum(len(b) for b in blobs) / 1e6)           # whole-job aggregate
    assert r["bins"] == sum(x.bins for x in recs)
