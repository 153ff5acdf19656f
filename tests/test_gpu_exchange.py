"""Gradient exchange over peer memory (csrc/exchange.cu, include/m2f_b200.h: m2f_grad_allreduce_peer).

One GPU: the slicing / unaligned-end arithmetic and the self-resetting barrier words with world == 1 (peer load / store
variant on an ordinary allocation).  Two GPUs: the kernel on a symmetric allocation, multicast (multimem) and peer
variants, against torch's sum of the two buffers; then the data-parallel step through `DataParallelM2FNet(exchange=
"peer")` against a single-GPU run on the concatenated batch (the same bar as tests/test_gpu_dp.py sets for NCCL)."""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
PKG_NAME = "multimodal-emotion-recognition_b200"


def _pkg():
    import importlib
    return importlib.import_module(PKG_NAME)


def test_world1_ranges_scale_in_place():
    pkg = _pkg()
    L, lib = pkg._lib, pkg._lib.lib()
    n = 1 << 20
    sig_off = n * 4
    buf = torch.zeros(n + L.XCHG_SIGNAL_BYTES // 4, device="cuda")
    x = torch.randn(n, device="cuda")
    buf[:n] = x
    pg = L.PeerGroup()
    pg.mc_base = None
    pg.peer_base[0] = buf.data_ptr()
    pg.signal_off_bytes = sig_off
    pg.rank, pg.world = 0, 1
    st = torch.cuda.current_stream().cuda_stream
    expect = x.clone()
    # aligned, unaligned head, unaligned tail, both, shorter than one vector, empty; several CTA counts
    for lo, hi, ctas in [(0, n, 0), (4096, 8192, 1), (5, 4099, 7), (8, 1003, 64), (1001, 1003, 0), (77, 78, 0), (300, 300, 0),
                         (3, n - 1, 200)]:
        L.check(lib.m2f_grad_allreduce_peer(ctypes.byref(pg), lo, hi, 0.5, ctas, L.XCHG_NO_EXIT_BARRIER if hi < n else 0, st),
                "grad_allreduce_peer")
        expect[lo:hi] *= 0.5
    torch.cuda.synchronize()
    assert torch.equal(buf[:n], expect)
    assert int(buf[n:].view(torch.int32).abs().sum()) == 0  # every barrier word consumed


def test_bad_arguments_are_errors():
    pkg = _pkg()
    L, lib = pkg._lib, pkg._lib.lib()
    buf = torch.zeros(1024 + L.XCHG_SIGNAL_BYTES // 4, device="cuda")
    pg = L.PeerGroup()
    pg.peer_base[0] = buf.data_ptr()
    pg.signal_off_bytes = 4096
    pg.rank, pg.world = 0, 1
    assert lib.m2f_grad_allreduce_peer(ctypes.byref(pg), 0, 2048, 1.0, 0, 0, None) == L.M2F_ERR_ARG  # range overlaps the signal area
    pg.world = 17
    assert lib.m2f_grad_allreduce_peer(ctypes.byref(pg), 0, 16, 1.0, 0, 0, None) == L.M2F_ERR_ARG
    pg.world, pg.rank = 2, 0  # peer 1 missing
    assert lib.m2f_grad_allreduce_peer(ctypes.byref(pg), 0, 16, 1.0, 0, 0, None) == L.M2F_ERR_ARG
    assert lib.m2f_grad_allreduce_peer(None, 0, 16, 1.0, 0, 0, None) == L.M2F_ERR_ARG


def _worker(rank, world, port, out, in_graph, wire="fp32"):
    os.environ["M2F_EXCHANGE_IN_GRAPH"] = "1" if in_graph else "0"
    import importlib
    import sys
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    pkg = importlib.import_module(PKG_NAME)
    D = importlib.import_module(PKG_NAME + ".data")
    L, lib = pkg._lib, pkg._lib.lib()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    res = {}
    cfg = D.default_model_config(dropout=0.0, AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=2),
                                 TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=2), FAM=dict(embedding_size=64, n_head=4, n_layers=2),
                                 CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2))
    torch.manual_seed(100 + rank)
    model = pkg.M2FNet(cfg, planes=3).to("cuda").train()
    try:
        net = pkg.DataParallelM2FNet(model, bucket_mb=0.05, exchange="peer", wire=wire)
    except RuntimeError as e:
        if rank == 0:
            torch.save({"unavailable": repr(e)}, out)
        dist.destroy_process_group()
        return
    ex = net.reducer
    res["multicast"] = ex.multicast
    res["wire"] = ex.wire
    # ---- the kernel alone: random buffers, unaligned ranges, multicast and peer variants
    n = ex.grads.numel()
    kerr = {}
    for variant in (["mc"] if ex.multicast else []) + ["p2p"]:
        pg = L.PeerGroup()
        ctypes.memmove(ctypes.byref(pg), ctypes.byref(ex.pg), ctypes.sizeof(pg))
        if variant == "p2p":
            pg.mc_base = None
        g = torch.Generator(device="cuda").manual_seed(7 + rank)
        x = torch.randn(n, device="cuda", generator=g)
        ex.grads.copy_(x)
        both = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(both, x)
        total = sum(both)
        expect = x.clone()
        torch.cuda.synchronize()
        dist.barrier()
        st = torch.cuda.current_stream().cuda_stream
        # a sequence on one stream: only the last call closes with the handshake (M2F_XCHG_NO_EXIT_BARRIER on the others)
        seq = [(n // 2 + 3, n - 1, 0), (16, 4096, 3), (5, 9, 0), (4099, n // 2 + 3, 64)]
        for k, (lo, hi, ctas) in enumerate(seq):
            flags = 0 if k == len(seq) - 1 else L.XCHG_NO_EXIT_BARRIER
            L.check(lib.m2f_grad_allreduce_peer(ctypes.byref(pg), lo, hi, 0.5, ctas, flags, st), "grad_allreduce_peer")
            expect[lo:hi] = total[lo:hi] * 0.5
        torch.cuda.synchronize()
        dist.barrier()
        kerr[variant] = (ex.grads - expect).abs().max().item()
    res["kernel_max_abs_err"] = kerr
    if ex.wire == "bf16":  # the kernel alone with the bf16 wire: aligned ranges take the bf16 path, the unaligned one falls back to fp32
        g = torch.Generator(device="cuda").manual_seed(70 + rank)
        x = torch.randn(n, device="cuda", generator=g)
        ex.grads.copy_(x)
        both = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(both, x)
        expect = x.clone()
        torch.cuda.synchronize()
        dist.barrier()
        st = torch.cuda.current_stream().cuda_stream
        n8 = n // 8 * 8
        seq = [(n8 // 2, n8, 0), (16, 4096, 3), (4099, 5003, 0), (8192, n8 // 2, 64)]
        for k, (lo, hi, ctas) in enumerate(seq):
            flags = L.XCHG_WIRE_BF16 | (0 if k == len(seq) - 1 else L.XCHG_NO_EXIT_BARRIER)
            L.check(lib.m2f_grad_allreduce_peer(ctypes.byref(ex.pg), lo, hi, 0.5, ctas, flags, st), "grad_allreduce_peer (bf16 wire)")
            if lo % 8 == 0 and (hi - lo) % 8 == 0:  # what the wire does: bf16(x_r) summed in fp32, rounded to bf16, scaled
                tot = sum(t[lo:hi].bfloat16().float() for t in both)
                expect[lo:hi] = (tot.bfloat16().float() * 0.5).bfloat16().float()
            else:
                expect[lo:hi] = sum(t[lo:hi] for t in both) * 0.5
        torch.cuda.synchronize()
        dist.barrier()
        d = (ex.grads - expect).abs()
        res["wire_kernel_max_abs_err"] = d.max().item()
        res["wire_kernel_rel"] = (torch.linalg.vector_norm(ex.grads[8192:n8] - (sum(both)[8192:n8] * 0.5)) /   # ranges 1 and 4 of the sequence
                                  torch.linalg.vector_norm(sum(both)[8192:n8] * 0.5)).item()
    # ---- the data-parallel step
    full = D.synth_batch(8, 9, 128, 64, seed=5)
    lo, hi = pkg.shard_range(8, rank, world)
    b = {k: v[lo:hi].cuda() for k, v in full.items()}
    crit = pkg.FusedCrossEntropyLoss()
    crit(net(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"]).backward()
    torch.cuda.synchronize()
    g_dp = model.flat_grads().clone()
    res["buckets"] = ex.calls_per_step
    res["in_graph"] = ex.in_graph
    solo = pkg.M2FNet(cfg, planes=3).to("cuda").train()
    solo.load_state_dict(model.state_dict())
    fb = {k: v.cuda() for k, v in full.items()}
    crit(solo(fb["text"], fb["audio"], fb["padding_mask"]).permute(0, 2, 1), fb["emotion"]).backward()
    torch.cuda.synchronize()
    res["err"] = (torch.linalg.vector_norm(g_dp - solo.flat_grads()) / torch.linalg.vector_norm(solo.flat_grads())).item()
    opt_dp = pkg.FusedAdam(model, lr=1e-3, weight_decay=0.01)
    opt_solo = pkg.FusedAdam(solo, lr=1e-3, weight_decay=0.01)
    opt_dp.step()
    opt_solo.step()
    for it in range(3):  # replayed backward graph + exchange kernels behind its event nodes
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
    res["param_err_after_4_steps"] = (torch.linalg.vector_norm(model.flat_parameters() - solo.flat_parameters()) /
                                      torch.linalg.vector_norm(solo.flat_parameters())).item()
    gathered = [torch.empty_like(g_dp) for _ in range(world)]
    dist.all_gather(gathered, g_dp)
    res["same"] = all(torch.equal(gathered[0], x) for x in gathered)
    if rank == 0:
        torch.save(res, out)
    dist.destroy_process_group()


def test_dp2_peer_exchange_bf16_wire(tmp_path):
    """The bf16 wire format: the kernel alone reproduces bf16(sum of bf16(x_r)) bit for bit on aligned ranges (and falls back
    to the exact fp32 path on an unaligned one); the data-parallel step stays within 4e-3 of the single-GPU gradients."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    out = str(tmp_path / "peer_bf16.pt")
    mp.spawn(_worker, args=(2, 29651, out, True, "bf16"), nprocs=2, join=True)
    r = torch.load(out)
    if "unavailable" in r:
        pytest.skip("symmetric memory unavailable on this box: " + r["unavailable"])
    print(r)
    if not r["multicast"]:
        pytest.skip("no multicast mapping on this box: the bf16 wire is not used")
    assert r["wire"] == "bf16"
    for variant, e in r["kernel_max_abs_err"].items():  # fp32 calls on the same group are untouched
        assert e < 1e-5, (variant, r)
    assert r["wire_kernel_max_abs_err"] <= 0.04, r         # bf16(fp32 sum of the bf16 inputs) * 0.5, to one bf16 ulp of |x| <= 8
    assert r["wire_kernel_rel"] < 4e-3, r
    assert r["same"] and r["buckets"] >= 3
    assert r["err"] < 4e-3, r
    assert r["param_err_after_4_steps"] < 1e-3, r


@pytest.mark.parametrize("in_graph", [True, False], ids=["in_graph", "callbacks"])
def test_dp2_peer_exchange_equals_single_gpu(tmp_path, in_graph):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    out = str(tmp_path / "peer.pt")
    mp.spawn(_worker, args=(2, 29641 + int(in_graph), out, in_graph), nprocs=2, join=True)
    r = torch.load(out)
    if "unavailable" in r:
        pytest.skip("symmetric memory unavailable on this box: " + r["unavailable"])
    print(r)
    assert r["in_graph"] == in_graph
    for variant, e in r["kernel_max_abs_err"].items():
        assert e < 1e-5, (variant, r)
    assert r["same"] and r["buckets"] >= 3
    assert r["err"] < 2e-5, r
    # as in tests/test_gpu_dp.py: the 6e-8 of summation-order noise in the first gradients grows through four Adam steps (small
    # entries move by lr whatever their size) to 0.9e-5 .. 1.2e-5; a range exchanged before its producers finished shows as >= 1e-2
    assert r["param_err_after_4_steps"] < 5e-5, r


def _train_steps(pkg, cfg, steps, exchange_world1, overlap, graph, bucket_floats=20000, B=5, S=9):
    """A few full training steps; with `exchange_world1` the gradient arena sits in a buffer with barrier words and a world-1
    peer group is attached, so the backward runs its in-graph exchange calls (in-place scale by 1) -- the single-GPU stand-in
    for the data-parallel path."""
    from oracle import m2fnet_oracle as O
    os.environ["M2F_GRAPH"] = graph
    L = pkg._lib
    params = O.synth_params(cfg)
    model = pkg.M2FNet(cfg, precision="bf16x2f")
    model.load_state_dict({k: v.float() for k, v in params.items()}, strict=True)
    model = model.to("cuda").train()
    model._seed = 4321
    keep = None
    if exchange_world1:
        n = model.flat_parameters().numel()
        sig = (n * 4 + 127) // 128 * 128
        keep = torch.zeros((sig + L.XCHG_SIGNAL_BYTES) // 4, device="cuda")
        pg = L.PeerGroup()
        pg.mc_base = None
        pg.peer_base[0] = keep.data_ptr()
        pg.signal_off_bytes = sig
        pg.rank, pg.world = 0, 1
        model.adopt_grad_arena(keep[:n])
        model.set_exchange(pg, 1.0, 0, bucket_floats)
    crit = pkg.FusedCrossEntropyLoss().cuda()
    opt = pkg.FusedAdam(model, lr=2e-4, weight_decay=0.01, overlap_exchange=overlap)
    losses, first = [], None
    for i in range(steps):
        b = O.synth_inputs(B, S, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size)
        opt.param_groups[0]["lr"] = 2e-4 / (i + 1)
        opt.zero_grad(set_to_none=True)
        out = model(b["text"].cuda() * (1 + 0.1 * i), b["audio"].cuda(), b["padding_mask"].cuda())
        loss = crit(out.permute(0, 2, 1), b["emotion"].cuda())
        loss.backward()
        applied = opt._applied
        opt.step()
        losses.append(loss.item())
        if i == 0:
            torch.cuda.synchronize()
            first = dict(params=model.flat_parameters().clone(), m=opt._m.clone(), v=opt._v.clone(), shadow=model._shadow.clone())
    torch.cuda.synchronize()
    os.environ.pop("M2F_GRAPH", None)
    return dict(params=model.flat_parameters().clone(), m=opt._m.clone(), v=opt._v.clone(), shadow=model._shadow.clone(), losses=losses,
                applied=applied, t=opt._t, ranges=getattr(model, "exchange_ranges", 0), model=model, opt=opt, keep=keep, first=first,
                off_1d=model._off_1d)


@pytest.mark.parametrize("graph", ["1", "0"])
@pytest.mark.parametrize("name", ["both_fam", "meld_dims"])
def test_update_behind_the_exchange_is_bit_identical(name, graph):
    """FusedAdam(overlap_exchange=True) with an in-graph exchange attached: the backward enqueues the Adam update of every
    exchanged range itself (m2f_plan_set_optimizer).  It must be the same update as the ordinary backward-then-step() path: bit
    for bit on the 2-D weights, their moments and the bf16 shadow after the first step, and the same trajectory after four steps
    with dropout and a changing learning rate."""
    from oracle import m2fnet_oracle as O
    pkg = _pkg()
    over = {"both_fam": dict(AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=3), TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=3),
                             FAM=dict(embedding_size=64, n_head=4, n_layers=3), CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2)),
            "meld_dims": dict(AUDIO=dict(n_encoder_layers=2), TEXT=dict(n_encoder_layers=2), FAM=dict(n_layers=2))}[name]
    cfg = O.default_model_cfg(dropout=0.4, **over)
    bf = 20000 if name == "both_fam" else 2 << 20
    B, S = (5, 9) if name == "both_fam" else (4, 33)
    plain = _train_steps(pkg, cfg, 4, False, False, graph, bf, B, S)
    xchg = _train_steps(pkg, cfg, 4, True, False, graph, bf, B, S)    # exchange calls in the backward, update in step()
    tail = _train_steps(pkg, cfg, 4, True, True, graph, bf, B, S)     # ... and the update behind the exchange
    assert not plain["applied"] and not xchg["applied"] and tail["applied"]
    assert tail["ranges"] >= 4 and xchg["ranges"] == tail["ranges"]
    assert plain["t"] == xchg["t"] == tail["t"] == 4
    off = plain["off_1d"]
    rel = lambda a, b: (torch.linalg.vector_norm(a.double() - b.double()) / torch.linalg.vector_norm(b.double())).item()
    for other in (xchg, tail):
        # After the first step: the 2-D weights, their moments and the bf16 shadow are bit-identical (their gradients are written
        # by one tile each, deterministically).  The 1-D gradients (biases, LayerNorm) are accumulated with atomics, so two runs
        # of the SAME path already differ in their last bits there -- and Adam's first step turns a gradient that is pure
        # rounding noise (the key bias of an attention, whose true gradient is 0) into +-lr: only a loose bound holds for them.
        assert other["losses"][0] == plain["losses"][0]
        for k in ("params", "m", "v"):
            assert torch.equal(other["first"][k][:off], plain["first"][k][:off]), k
        assert rel(other["first"]["params"][off:], plain["first"]["params"][off:]) < 1e-2
        assert (other["first"]["shadow"] != plain["first"]["shadow"]).float().mean().item() < 1e-3
        # after four steps (changing learning rate, dropout): the same trajectory, up to what the +-lr noise on the
        # zero-gradient biases does to later forwards (two runs of the ordinary path differ by as much)
        assert rel(other["params"][:off], plain["params"][:off]) < 1e-3
        for k in ("m", "v"):
            assert rel(other[k][:off], plain[k][:off]) < 1e-1, (k, rel(other[k][:off], plain[k][:off]))
        assert max(abs(a - b) / abs(b) for a, b in zip(other["losses"], plain["losses"])) < 1e-2


def test_update_behind_the_exchange_misuse():
    from oracle import m2fnet_oracle as O
    pkg = _pkg()
    cfg = O.default_model_cfg(dropout=0.0, AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=1), TEXT=dict(embedding_size=64, n_head=4, n_encoder_layers=1),
                              FAM=dict(embedding_size=64, n_head=4, n_layers=1), CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2))
    r = _train_steps(pkg, cfg, 1, True, True, "1")
    model, opt = r["model"], r["opt"]
    b = O.synth_inputs(5, 9, 64, 64)
    crit = pkg.FusedCrossEntropyLoss().cuda()
    fwd = lambda: crit(model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda()).permute(0, 2, 1), b["emotion"].cuda())
    opt.zero_grad(set_to_none=True)
    fwd().backward()
    with pytest.raises(RuntimeError, match="twice without step"):
        opt.zero_grad(set_to_none=True)
        fwd().backward()
    opt.step()
    opt.detach_from_backward()  # update in step() again
    before = model.flat_parameters().clone()
    opt.zero_grad(set_to_none=True)
    fwd().backward()
    torch.cuda.synchronize()
    assert torch.equal(before, model.flat_parameters())
    opt.step()
    torch.cuda.synchronize()
    assert not torch.equal(before, model.flat_parameters())
