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

