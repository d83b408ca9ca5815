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


def _strong_worker(rank, world, port, total, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # what bench.py's strong leg does: every rank computes the same greedy longest-first assignment and keeps its part
    part = bench.lpt_assign([bench.clip_weight(i) for i in range(total)], world)
    mine = [i for i in range(total) if part[i] == rank]
    load = torch.tensor([float(sum(bench.clip_weight(i) for i in mine))], dtype=torch.float64)
    loads = [torch.zeros_like(load) for _ in range(world)]
    dist.all_gather(loads, load)
    ids = torch.full((total,), -1, dtype=torch.int64)
    ids[:len(mine)] = torch.tensor(mine, dtype=torch.int64)
    allids = [torch.zeros_like(ids) for _ in range(world)]
    dist.all_gather(allids, ids)
    if rank == 0:
        import json
        json.dump(dict(loads=[x.item() for x in loads], ids=[[int(v) for v in a.tolist() if v >= 0] for a in allids]),
                  open(os.path.join(out_dir, "s.json"), "w"))
    dist.destroy_process_group()


def test_strong_leg_assignment_is_a_balanced_partition_every_rank_agrees_on():
    """bench.py `strong` (BASELINE configs[3]): N clips in total over the ranks, greedy longest-first by macroblocks
    coded; no collective on the data path, every rank derives the same partition locally"""
    import json
    import torch.multiprocessing as mp
    import bench
    total = 41
    with tempfile.TemporaryDirectory() as out:
        mp.spawn(_strong_worker, args=(2, 29741, total, out), nprocs=2, join=True)
        r = json.load(open(os.path.join(out, "s.json")))
    assert sorted(r["ids"][0] + r["ids"][1]) == list(range(total))              # a partition: disjoint and complete
    w = [bench.clip_weight(i) for i in range(total)]
    assert r["loads"] == [float(sum(w[i] for i in part)) for part in r["ids"]]
    # LPT guarantee: the loads differ by at most one item (the smallest assigned to the heavier rank)
    heavy = max(range(2), key=lambda k: r["loads"][k])
    assert abs(r["loads"][0] - r["loads"][1]) <= min(w[i] for i in r["ids"][heavy])
    # determinism: the same call gives the same answer, and one rank gets everything when there is one rank
    assert bench.lpt_assign(w, 2) == bench.lpt_assign(w, 2)
    assert set(bench.lpt_assign(w, 1)) == {0}
