"""Host logic of the data-parallel path on CPU: world_size-2 gloo run of the bucket reducer, fed the way
m2f_plan_backward_bucketed reports finished gradient ranges (top of the arena downwards, 1-D region last)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

PKG_NAME = "multimodal-emotion-recognition_b200"


def _worker(rank, world, port, n, ranges, out):
    import importlib
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    pkg = importlib.import_module(PKG_NAME)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    flat = torch.arange(n, dtype=torch.float32) * (rank + 1)
    red = pkg.GradBucketReducer(lambda: flat, average=True)
    red.begin_step()
    for lo, hi in ranges:
        red.on_bucket(lo, hi)
    red.finish()
    if rank == 0:
        torch.save({"flat": flat, "log": red.log, "bytes": red.bytes_last_step}, out)
    dist.destroy_process_group()


def test_bucket_reducer_world2_gloo(tmp_path):
    n, off_1d = 1000, 900
    ranges = [(600, 900), (250, 600), (0, 250), (900, 1000)]  # the order the backward reports them
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, 29611, n, ranges, out), nprocs=2, join=True)
    r = torch.load(out)
    assert torch.allclose(r["flat"], torch.arange(n, dtype=torch.float32) * 1.5)
    assert sorted(r["log"]) == [(0, 250), (250, 600), (600, 900), (900, 1000)] and r["bytes"] == 4 * n


def _wrap_worker(rank, world, port, out):
    import importlib
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    pkg = importlib.import_module(PKG_NAME)
    D = importlib.import_module(PKG_NAME + ".data")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = D.default_model_config(dropout=0.0, AUDIO=dict(embedding_size=16, n_head=2, n_encoder_layers=1),
                                 TEXT=dict(embedding_size=16, n_head=2, n_encoder_layers=1), FAM=dict(embedding_size=16, n_head=2, n_layers=1),
                                 CLASSIFIER=dict(hidden_size=16, output_size=7, n_layers=2))
    torch.manual_seed(10 + rank)  # different init per rank
    model = pkg.M2FNet(cfg)  # stays on the host: only the wrapper's set-up is exercised
    res = {}
    for bad, exc in (("bogus", ValueError), ("peer", RuntimeError)):
        try:
            pkg.DataParallelM2FNet(model, exchange=bad)
            res[bad] = "no error"
        except exc as e:
            res[bad] = type(e).__name__
    net = pkg.DataParallelM2FNet(model)  # auto on a host model: the generic bucket reducer
    res["exchange"], res["bucket_mb"] = net.exchange, net.bucket_mb
    flat = model.flat_parameters().clone()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    res["broadcast"] = all(torch.equal(gathered[0], g) for g in gathered)
    if rank == 0:
        torch.save(res, out)
    dist.destroy_process_group()


def test_wrapper_setup_world2_gloo(tmp_path):
    out = str(tmp_path / "w.pt")
    mp.spawn(_wrap_worker, args=(2, 29613, out), nprocs=2, join=True)
    r = torch.load(out)
    assert r["bogus"] == "ValueError" and r["peer"] == "RuntimeError"
    assert r["exchange"] == "nccl" and r["bucket_mb"] == 128.0 and r["broadcast"]


def test_shard_range_partitions_exactly(pkg):
    for n, w in ((100000, 8), (10, 4), (3, 8), (1098, 2)):
        parts = [pkg.shard_range(n, r, w) for r in range(w)]
        assert parts[0][0] == 0 and parts[-1][1] == n
        assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
        sizes = [b - a for a, b in parts]
        assert max(sizes) - min(sizes) <= 1


def test_synthetic_loader_layout_matches_collate_fn(pkg):
    D = __import__("importlib").import_module(PKG_NAME + ".data")
    dl = D.SyntheticMELD(n_dialogues=21, batch_size=8, S=12, Dt=16, Da=8, ragged=True, rank=1, world=2)
    batches = list(dl)
    assert sum(b["text"].shape[0] for b in batches) == 10  # shard sizes 11 + 10
    b = batches[0]
    assert b["text"].dtype == torch.float32 and b["padding_mask"].dtype == torch.bool and b["emotion"].dtype == torch.int64
    assert (b["emotion"][b["padding_mask"]] == -1).all() and (b["emotion"][~b["padding_mask"]] >= 0).all()
    assert (b["text"][b["padding_mask"]] == 0).all() and (~b["padding_mask"]).any(dim=1).all()



def test_every_rank_runs_the_same_number_of_steps(pkg):
    """A rank with one batch more than its peers would wait in the (collective) gradient exchange for ever: the sharded
    loaders give every rank the same step count whatever the dialogue count (uneven shards)."""
    import numpy as np
    D = __import__("importlib").import_module(PKG_NAME + ".data")
    for n, bs, world in ((97, 32, 3), (21, 8, 2), (100000, 64, 8), (7, 2, 4)):
        lens = [len(D.SyntheticMELD(n_dialogues=n, batch_size=bs, S=5, Dt=8, Da=8, rank=r, world=world)) for r in range(world)]
        assert len(set(lens)) == 1, (n, bs, world, lens)
    rng = np.random.default_rng(0)
    lengths = rng.integers(1, 34, size=1037)
    for world in (2, 3, 8):
        per_rank = [list(D.LengthBucketSampler(lengths, 16, seed=3, rank=r, world=world)) for r in range(world)]
        assert len({len(b) for b in per_rank}) == 1
        seen = np.concatenate([c for b in per_rank for c, _ in b])
        assert len(seen) == len(set(seen.tolist()))  # no dialogue served twice
        assert len(seen) >= 1037 - 16 * (world - 1) - 16 * 4  # only tail batches were dropped


def _uneven_worker(rank, world, port, out):
    """Drives the sharded sampler + reducer the way a data-parallel epoch does, with a dialogue count that does not divide."""
    import importlib
    import sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    pkg = importlib.import_module(PKG_NAME)
    D = importlib.import_module(PKG_NAME + ".data")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lengths = np.random.default_rng(1).integers(1, 34, size=203)
    sampler = D.LengthBucketSampler(lengths, 8, seed=5, rank=rank, world=world)
    flat = torch.zeros(64)
    red = pkg.GradBucketReducer(lambda: flat, average=False)
    steps = 0
    for idx, S in sampler:  # one collective per step: hangs (and the test times out) if the ranks disagree on the count
        flat.fill_(float(len(idx)))
        red.begin_step()
        red.on_bucket(0, 64)
        red.finish()
        steps += 1
    if rank == 0:
        torch.save({"steps": steps}, out)
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_uneven_dialogue_count_world2_gloo(tmp_path):
    out = str(tmp_path / "u.pt")
    mp.spawn(_uneven_worker, args=(2, 29617, out), nprocs=2, join=True)
    assert torch.load(out)["steps"] > 0
