"""Exchange kernel alone (no backward beside it): time m2f_grad_allreduce_peer over range sizes, CTA counts and CTA
shapes on a symmetric buffer.  Run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/exchange_sweep.py

Prints per-call time and the algorithmic bandwidth (range bytes / time) per configuration; `slice GB/s` is what one rank
pulls (and pushes) through its SMs: range bytes / world / time."""
import ctypes
import importlib
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
L, lib = pkg._lib, pkg._lib.lib()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 86_251_840  # the default model's arena
sig_off = (n * 4 + 127) // 128 * 128
buf = symm_mem.empty((sig_off + L.XCHG_SIGNAL_BYTES) // 4, dtype=torch.float32, device=f"cuda:{local}")
buf.zero_()
h = symm_mem.rendezvous(buf, dist.group.WORLD)
torch.cuda.synchronize()
dist.barrier()
pg = L.PeerGroup()
pg.mc_base = int(h.multicast_ptr) or None
for r, x in enumerate(h.buffer_ptrs):
    pg.peer_base[r] = int(x)
pg.signal_off_bytes, pg.rank, pg.world = sig_off, rank, world
st = torch.cuda.current_stream().cuda_stream
variants = [("mc", t, u) for t, u in [(512, 8), (512, 4), (256, 8), (256, 16), (1024, 4), (128, 16)]] + [("p2p", 512, 8)]
sizes_mb = [0.5, 5.5, 22, 88, 329]
for kind, t, u in variants:
    L.check(lib.m2f_exchange_tune(t, u), "tune")
    g = L.PeerGroup()
    ctypes.memmove(ctypes.byref(g), ctypes.byref(pg), ctypes.sizeof(pg))
    if kind == "p2p":
        g.mc_base = None
    for mb in sizes_mb:
        floats = int(mb * (1 << 20) / 4)
        for ctas in (8, 16, 32, 64):
            reps = 20 if mb < 100 else 5
            for _ in range(2):
                L.check(lib.m2f_grad_allreduce_peer(ctypes.byref(g), 0, floats, 1.0, ctas, 0, st), "xchg")
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                L.check(lib.m2f_grad_allreduce_peer(ctypes.byref(g), 0, floats, 1.0, ctas, 0, st), "xchg")
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / reps
            if rank == 0:
                print(f"{kind} threads {t:4d} unroll {u:2d}  {mb:6.1f} MB  ctas {ctas:2d}: {us:8.1f} us  alg {floats * 4 / us / 1e3:7.1f} GB/s  "
                      f"slice {floats * 4 / world / us / 1e3:6.1f} GB/s", flush=True)
dist.barrier()
dist.destroy_process_group()
