"""Where does the data-parallel step spend its time?  Run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_trace.py [exchange] [bucket_mb] [outdir]

Each rank writes <outdir>/dp_timeline_rank<r>.tsv: the plan's per-op %globaltimer timeline of one training step
(M2F_TIMELINE) merged with the exchange kernel's own trace (m2f_exchange_trace) on the same clock, and prints a summary:
when each gradient range became final, how long its exchange waited for the peers, how long the data phase took, and
what was left after the last backward kernel."""
import ctypes
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
exchange = sys.argv[1] if len(sys.argv) > 1 else "peer"
bucket_mb = float(sys.argv[2]) if len(sys.argv) > 2 else None
outdir = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out"
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
os.makedirs(outdir, exist_ok=True)
tl_file = os.path.join(outdir, f"_tl_raw_rank{rank}.tsv")
STEP = 8
os.environ["M2F_TIMELINE"] = tl_file
os.environ["M2F_TIMELINE_STEP"] = str(STEP)

pkg = importlib.import_module("multimodal-emotion-recognition_b200")
D = importlib.import_module("multimodal-emotion-recognition_b200.data")
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.manual_seed(0)
model = pkg.M2FNet(D.default_model_config(), precision=os.environ.get("M2F_TRACE_PRECISION", "bf16x2f")).to("cuda").train()
net = pkg.DataParallelM2FNet(model, bucket_mb=bucket_mb, exchange=exchange, wire=os.environ.get("M2F_TRACE_WIRE", "fp32"))
crit = pkg.FusedCrossEntropyLoss().cuda()
opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
b = {k: v.cuda() for k, v in D.synth_batch(64, 33, 768, 768, seed=1234 + rank).items()}
L = pkg._lib.lib()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
for i in range(STEP + 2):
    if i == STEP + 1:
        ev[0].record()
    opt.zero_grad(set_to_none=True)
    loss = crit(net(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"])
    if i == STEP + 1:
        ev[1].record()
    loss.backward()
    if i == STEP + 1:
        ev[2].record()
    opt.step()
    if i == STEP + 1:
        ev[3].record()
torch.cuda.synchronize()
dist.barrier()
print(f"[rank {rank}] untraced step: fwd+loss {ev[0].elapsed_time(ev[1]):.3f} ms, bwd (+exchange wait) {ev[1].elapsed_time(ev[2]):.3f} ms, "
      f"adam {ev[2].elapsed_time(ev[3]):.3f} ms", flush=True)

# host time per step: how long the Python thread needs to enqueue a step (no device sync inside the loop)
import time
torch.cuda.synchronize()
dist.barrier()
t_h = [0.0, 0.0, 0.0, 0.0]
w0 = time.perf_counter()
for i in range(20):
    a = time.perf_counter()
    opt.zero_grad(set_to_none=True)
    loss = crit(net(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"])
    c1 = time.perf_counter()
    loss.backward()
    c2 = time.perf_counter()
    opt.step()
    c3 = time.perf_counter()
    t_h[0] += c1 - a
    t_h[1] += c2 - c1
    t_h[2] += c3 - c2
w1 = time.perf_counter()
torch.cuda.synchronize()
w2 = time.perf_counter()
print(f"[rank {rank}] host enqueue per step: fwd+loss {t_h[0] / 20 * 1e3:.3f} ms, bwd {t_h[1] / 20 * 1e3:.3f} ms, adam {t_h[2] / 20 * 1e3:.3f} ms; "
      f"loop {1e3 * (w1 - w0) / 20:.3f} ms/step, with final sync {1e3 * (w2 - w0) / 20:.3f} ms/step", flush=True)

rows, t0 = [], 0
for line in open(tl_file):
    if line.startswith("#t0_ns"):
        t0 = int(line.split()[1])
        continue
    _i, sid, us, label = line.rstrip("\n").split("\t")
    rows.append((float(us), f"s{sid}", label))
if net.exchange == "peer":
    N = 512
    buf = (ctypes.c_uint64 * (N * 6))()
    n = int(L.m2f_exchange_trace(buf))
    per_step = net.reducer.calls_per_step
    # the traced backward is call number STEP (1-based) of STEP + 2: its records are [per_step*(STEP-1), per_step*STEP)
    for k in range(per_step * (STEP - 1), per_step * STEP):
        r = [int(buf[(k % N) * 6 + j]) for j in range(6)]
        rel = [(x - t0) * 1e-3 for x in r[:4]]
        rows.append((rel[0], "xchg", f"exchange enter lo={r[4]} MB={r[5] * 4 / 2**20:.1f}"))
        rows.append((rel[1], "xchg", f"exchange peers arrived (+{rel[1] - rel[0]:.1f} us)"))
        rows.append((rel[2], "xchg", f"exchange data done (+{rel[2] - rel[1]:.1f} us)"))
        rows.append((rel[3], "xchg", f"exchange exit (+{rel[3] - rel[2]:.1f} us)"))
rows.sort()
out = os.path.join(outdir, f"dp_timeline_rank{rank}.tsv")
with open(out, "w") as f:
    for us, sid, label in rows:
        f.write(f"{us:10.1f}\t{sid}\t{label}\n")
os.remove(tl_file)
last_bwd = max((us for us, sid, label in rows if sid != "xchg"), default=0.0)
last_x = max((us for us, sid, label in rows if sid == "xchg"), default=0.0)
print(f"[rank {rank}] traced backward (stamps add overhead): last backward op at {last_bwd:.0f} us, last exchange exit at {last_x:.0f} us; "
      f"wrote {out}", flush=True)
dist.destroy_process_group()
