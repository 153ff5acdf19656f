"""Data-parallel path on 2 GPUs (NCCL): bucketed, overlapped gradient all-reduce must equal a single-GPU backward on
the concatenated batch (full-length dialogues: the mean of per-rank CE means equals the global mean)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
PKG_NAME = "multimodal-emotion-recognition_b200"


def _worker(rank, world, port, out):
    import importlib
    import sys
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    pkg = importlib.import_module(PKG_NAME)
    D = importlib.import_module(PKG_NAME + ".data")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    cfg = D.default_model_config(dropout=0.0, AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=2),
                                 TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=2), FAM=dict(embedding_size=64, n_head=4, n_layers=2),
                                 CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2))
    torch.manual_seed(100 + rank)  # different init per rank: the wrapper must broadcast rank 0's
    model = pkg.M2FNet(cfg, planes=3).to("cuda").train()
    net = pkg.DataParallelM2FNet(model, bucket_mb=0.05, exchange="nccl")
    full = D.synth_batch(8, 9, 128, 64, seed=5)
    lo, hi = pkg.shard_range(8, rank, world)
    b = {k: v[lo:hi].cuda() for k, v in full.items()}
    crit = pkg.FusedCrossEntropyLoss()
    loss = crit(net(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"])
    loss.backward()
    torch.cuda.synchronize()
    g_dp = model.flat_grads().clone()
    n_buckets = len(net.reducer.log)
    # single-GPU reference on the concatenated batch with the same (broadcast) weights
    solo = pkg.M2FNet(cfg, planes=3).to("cuda").train()
    solo.load_state_dict(model.state_dict())
    fb = {k: v.cuda() for k, v in full.items()}
    crit(solo(fb["text"], fb["audio"], fb["padding_mask"]).permute(0, 2, 1), fb["emotion"]).backward()
    torch.cuda.synchronize()
    err = (torch.linalg.vector_norm(g_dp - solo.flat_grads()) / torch.linalg.vector_norm(solo.flat_grads())).item()
    # several more steps through the optimizer: the replayed backward graph must keep ordering every bucket's
    # all-reduce behind the kernels that produce it (a stale event would reduce half-written gradients)
    opt_dp = pkg.FusedAdam(model, lr=1e-3, weight_decay=0.01)
    opt_solo = pkg.FusedAdam(solo, lr=1e-3, weight_decay=0.01)
    opt_dp.step()
    opt_solo.step()
    for it in range(3):
        fb2 = D.synth_batch(8, 9, 128, 64, seed=50 + it)
        b2 = {k: v[lo:hi].cuda() for k, v in fb2.items()}
        fb2 = {k: v.cuda() for k, v in fb2.items()}
        opt_dp.zero_grad()
        crit(net(b2["text"], b2["audio"], b2["padding_mask"]).permute(0, 2, 1), b2["emotion"]).backward()
        opt_dp.step()
        opt_solo.zero_grad()
        crit(solo(fb2["text"], fb2["audio"], fb2["padding_mask"]).permute(0, 2, 1), fb2["emotion"]).backward()
        opt_solo.step()
    torch.cuda.synchronize()
    perr = (torch.linalg.vector_norm(model.flat_parameters() - solo.flat_parameters()) / torch.linalg.vector_norm(solo.flat_parameters())).item()
    gathered = [torch.empty_like(g_dp) for _ in range(world)]
    dist.all_gather(gathered, g_dp)
    same = all(torch.equal(gathered[0], x) for x in gathered)
    if rank == 0:
        torch.save({"err": err, "same": same, "buckets": n_buckets, "param_err_after_4_steps": perr}, out)
    dist.destroy_process_group()


def test_dp2_gradients_equal_single_gpu(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    out = str(tmp_path / "dp.pt")
    mp.spawn(_worker, args=(2, 29633, out), nprocs=2, join=True)
    r = torch.load(out)
    assert r["same"] and r["buckets"] >= 3
    assert r["err"] < 2e-5, r
    # four Adam steps (lr 1e-3) later the two runs may differ by what the 6e-8 of summation-order noise in the first gradients grows
    # into (Adam divides by sqrt(v): small gradient entries move by lr whatever their size) -- measured 0.9e-5 .. 1.2e-5; a bucket
    # reduced before its producers finished (what this part of the test is for) shows up as 1e-2 and more
    assert r["param_err_after_4_steps"] < 5e-5, r
