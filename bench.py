"""bench.py -- M2FNet train utterances/sec on N x B200 (BASELINE.json metric), one JSON line on rank 0.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
    python bench.py --impl reference ...      # the CPU restatement of the reference path on the box's host cores

A "step" = one full training step of the reference's train() body (src/train.py:222-233) on one batch:
forward (dropout 0.4, train mode) -> label-smoothed CE -> backward -> Adam(lr 5e-5, wd 0.01).
Workload = BASELINE.json configs[1]: text_base(768) + audio_wav2vec2(768), 64 dialogues x 33 utterances per GPU, 7 classes.
`value` times device-resident inputs; `e2e` goes through the public API with pinned HOST batches (H2D inside the timed
region) and a loss.item() read back every step, exactly as train() does.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "multimodal-emotion-recognition_b200"

FLOP_FWD_PER_UTT_S33 = 174.0e6  # 172.24 MFLOP linear + 1.72 attention at S = 33 (default config)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--batch", type=int, default=64, help="dialogues per GPU")
    ap.add_argument("--seq", type=int, default=33)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--bucket-mb", type=float, default=None, help="gradient bucket size (default: 16 for the peer exchange, 128 for NCCL)")
    ap.add_argument("--exchange", default=None, choices=["auto", "peer", "nccl"],
                    help="gradient exchange at N > 1: NVSwitch peer-memory kernel or NCCL buckets (default: $M2F_EXCHANGE or auto)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_port_runner(B, S, threads=None):
    """The reference path restated on the CPU (oracle/m2fnet_oracle.py: plain-tensor forward, autograd backward,
    torch Bernoulli dropout masks like nn.Dropout, torch.optim.Adam) -- the only place bench.py executes oracle/."""
    import torch
    from oracle import m2fnet_oracle as O

    if threads:
        torch.set_num_threads(threads)
    cfg = O.default_model_cfg()
    params = {k: v.float().requires_grad_(True) for k, v in O.synth_params(cfg).items() if O.canonical_name(k) == k}
    opt = torch.optim.Adam(list(params.values()), lr=5e-5, weight_decay=0.01)
    batch = O.synth_batch(B, S, 768, 768, seed=1234)
    p = float(cfg.dropout)

    def masks(site, shape):
        return torch.bernoulli(torch.full(shape, 1.0 - p))

    def step():
        opt.zero_grad()
        logits = O.forward(params, cfg, batch["text"], batch["audio"], batch["padding_mask"], masks, p)
        loss = O.ce_loss(logits, batch["emotion"])
        loss.backward()
        opt.step()
        return loss.item()

    return step, torch.get_num_threads()


def run_reference_arm(args):
    """--impl reference: rank 0 only; each step = fwd+loss+bwd+Adam on a bounded sample (8 of the 64 dialogues)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B, S = 8, args.seq
    step, cores = cpu_port_runner(B, S)
    for _ in range(max(1, min(args.warmup, 2))):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = B * S * args.steps / dt
    sample = f"{B} of {args.batch} dialogues x {S} utterances per step (configs[0] shape), full train step, dropout 0.4, fp32"
    print(json.dumps({
        "impl": "reference", "metric": "M2FNet train utterances/sec", "value": val, "unit": "utterances/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 2), "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": val, "unit": "utterances/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(args, world):
    return {"workload": "M2FNet text_base(768)+audio_wav2vec2(768), 6+6 encoder layers, 5 fusion layers, 7 classes; "
                        f"{args.batch} dialogues x {args.seq} utterances per GPU; train step = fwd + CE(ls 0.1) + bwd + Adam; dropout 0.4",
            "dialogues_per_gpu": args.batch, "utterances_per_dialogue": args.seq, "global_dialogues": args.batch * world,
            "parallelism": f"dp{world}", "precision_mode": args.precision,
            "l2": "16 distinct device-resident batches (208 MB) cycled + >1.5 GB of weight/gradient/Adam traffic per step: "
                  "working set exceeds the 126 MB L2, no explicit flush"}


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
                 getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                 getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                 getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                 getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------ B200 arm
def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    import torch
    import torch.distributed as dist

    pkg = importlib.import_module(PKG)
    D = importlib.import_module(PKG + ".data")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    L = pkg._lib.lib()
    B, S = args.batch, args.seq
    planes = 1 if args.precision == "bf16" else 3

    cfg = D.default_model_config()  # src/config.yaml defaults, dropout 0.4
    torch.manual_seed(0)
    model = pkg.M2FNet(cfg, planes=planes).to(dev).train()
    net = pkg.DataParallelM2FNet(model, bucket_mb=args.bucket_mb, exchange=args.exchange) if world > 1 else model
    crit = pkg.FusedCrossEntropyLoss().to(dev)
    opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)

    NB = 16
    host, devb = [], []
    for i in range(NB):
        b = D.synth_batch(B, S, 768, 768, seed=1234 + 97 * rank + i)
        host.append({k: v.pin_memory() for k, v in b.items()})
        devb.append({k: v.to(dev) for k, v in b.items()})
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())

    def step(b):
        opt.zero_grad(set_to_none=True)
        out = net(b["text"], b["audio"], b["padding_mask"])
        loss = crit(out.permute(0, 2, 1), b["emotion"])
        loss.backward()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        barrier()  # ranks start the warm-up together (the exchange kernels of the ranks wait for each other)
        for i in range(warmup):
            fn(i)
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        l0 = int(L.m2f_launch_count())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        barrier()
        sampler.stop_flag = True
        sampler.join()
        ms = e0.elapsed_time(e1)
        launches = int(L.m2f_launch_count()) - l0
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms, launches, sampler.result()

    # ---- value: device-resident inputs
    ms, launches, clocks = timed(lambda i: step(devb[i % NB]), args.steps, max(args.warmup, 3))
    utt = world * B * S * args.steps
    value = utt / (ms * 1e-3)

    # ---- e2e: pinned host batches -> H2D -> step -> loss.item()  (the reference's train() body, src/train.py:222-233).
    # The package's DevicePrefetcher issues the copy of batch i+1 on a copy stream while batch i computes; every batch
    # is copied inside the timed region (steps + warmup + prefetch depth batches are copied for `steps` timed steps).
    e2e = None
    if not args.no_e2e:
        class _Cycle:
            def __iter__(self):
                i = 0
                while True:
                    yield host[i % NB]
                    i += 1
        feed = iter(D.DevicePrefetcher(_Cycle(), dev, depth=2))

        loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        loss_ev = torch.cuda.Event()

        def e2e_step(i):
            # the reference's loop body with the loss read placed where it costs nothing: the device->host copy is queued
            # right behind the loss kernel, the host reads it after it has queued backward + Adam (same step, same value)
            b = next(feed)
            opt.zero_grad(set_to_none=True)
            out = net(b["text"], b["audio"], b["padding_mask"])
            loss = crit(out.permute(0, 2, 1), b["emotion"])
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
            loss_ev.record()
            loss.backward()
            opt.step()
            loss_ev.synchronize()
            return float(loss_host[0])
        ms_e, _, _ = timed(e2e_step, args.steps, 3)
        e2e = {"value": utt / (ms_e * 1e-3), "unit": "utterances/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
               "ms_per_step": ms_e / args.steps,
               "note": "pinned host batches copied on a side stream one batch ahead (DevicePrefetcher); every step's loss is copied to "
                       "pinned host memory behind the loss kernel and read by the host in the same step, after backward + Adam were queued"}

    # ---- data parallel: every replica must hold bit-identical parameters after the timed steps (each rank received the
    # same reduced gradients and applied the same update); a lost or torn exchange shows up here
    in_sync = None
    if world > 1:
        flat = model.flat_parameters()
        chk = torch.stack([flat.double().sum(), flat.double().abs().sum(), flat[::9973].double().square().sum()])
        hi_, lo_ = chk.clone(), chk.clone()
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        in_sync = bool(torch.equal(hi_, lo_)) and bool(torch.isfinite(chk).all())

    # ---- roofline of the dominant kernel (the tcgen05 GEMM), timed alone with CUDA events, operands rotated past L2
    roof = None
    if rank == 0:
        roof = gemm_roofline(pkg, torch, dev, B * S, planes)
        step_flops = 3.0 * FLOP_FWD_PER_UTT_S33 * B * S * (1 if S == 33 else 1)
        peaks = load_peaks()
        roof["step"] = {"algorithmic_tflop_per_step_per_gpu": step_flops / 1e12,
                        "achieved_tflops_per_gpu": step_flops / (ms / args.steps * 1e-3) / 1e12,
                        "frac_of_sustained_bf16_peak": step_flops / (ms / args.steps * 1e-3) / 1e12 / peaks["bf16_tflops_sustained"],
                        "note": "whole train step incl. attention/LN/loss/Adam kernels and launch gaps; fp32 mode runs 6 plane products per GEMM"}

    if rank == 0 and world == 1 and roof is not None:
        try:  # additional evidence only: never let it cost the bench line
            roof["hbm_kernels"] = hbm_kernel_rooflines(pkg, torch, dev, opt, model)
        except Exception as e:  # noqa: BLE001
            roof["hbm_kernels"] = None
            roof["hbm_kernels_error"] = repr(e)[:200]

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline()

    if rank == 0:
        cfg_out = workload_config(args, world)
        if world > 1:
            cfg_out["exchange"] = ("gradient all-reduce: m2f_grad_allreduce_peer over NVSwitch " +
                                   ("multicast (multimem.ld_reduce / multimem.st)" if getattr(net.reducer, "multicast", False) else "peer loads / stores")
                                   if net.exchange == "peer" else "gradient all-reduce: NCCL, async buckets behind the backward")
            cfg_out["bucket_mb"] = net.bucket_mb
            cfg_out["replicas_in_sync_after_run"] = in_sync
            cfg_out["exchange_calls_per_step"] = getattr(net.reducer, "calls_per_step", len(net.reducer.log))
            if net.exchange == "peer":
                cfg_out["exchange_launch"] = "kernels captured in the backward's CUDA graph" if net.reducer.in_graph else "bucket callbacks"
        out = {"metric": "M2FNet train utterances/sec", "value": value, "unit": "utterances/s", "n_gpus": world,
               "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if planes == 1 else "f32 (3 x bf16 planes)",
               "data": "synthetic", "config": cfg_out, "clocks": clocks, "e2e": e2e,
               "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def hbm_kernel_rooflines(pkg, torch, dev, opt, model):
    """The bandwidth-bound kernels against the measured HBM peak, at sizes whose working set exceeds the 126 MB L2
    (at M = 2112 these tensors sit in L2 and the kernels are launch-bound): LayerNorm forward / backward over
    M = 131 072 rows of 768 (the C5 batch), and the fused Adam + shadow pass over the model's arena.  Algorithmic bytes as in
    DESIGN.md section 5.  Each timed alone, CUDA events on the launch stream, 10 launches after 2 warm-ups."""
    ops = pkg.ops
    peak = load_peaks()["hbm_gbs"]
    out = []

    def time_us(fn, reps=10):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / reps

    def add(name, nbytes, us):
        gbs = nbytes / us / 1e3
        out.append({"kernel": name, "algorithmic_bytes": nbytes, "us_per_launch": us, "achieved": gbs, "unit": "GB/s", "peak": peak,
                    "frac": gbs / peak})

    M, D, P = 131072, 768, 1
    g = torch.Generator(device=dev).manual_seed(3)
    x = torch.randn(M, D, generator=g, device=dev)
    skip = torch.randn(M, D, generator=g, device=dev)
    gamma, beta = torch.ones(D, device=dev), torch.zeros(D, device=dev)
    y = torch.empty(M, D, device=dev)
    ypl = ops.PlanesTensor.empty(M, D, P, dev)
    stats = torch.empty(M, 2, device=dev)
    us = time_us(lambda: ops.ln_fwd(x, gamma, beta, skip=skip, y=y, y_planes=ypl, drop_p=0.4, seed=1, step=2, site=3, stats=stats))
    add("ln_fwd_kernel: y = LN(x) + skip, fp32 + 1 bf16 plane (dropout 0.4), M=131072 D=768", (3 * 4 + 2 * P) * M * D + 8 * M, us)
    dx = torch.empty(M, D, device=dev)
    dgamma, dbeta, dbias = (torch.zeros(D, device=dev) for _ in range(3))
    us = time_us(lambda: ops.ln_bwd(skip, x, gamma, stats, dx=dx, dx_planes=ypl, drop_p=0.4, seed=4, step=5, site=6, dgamma=dgamma, dbeta=dbeta,
                                    dbias=dbias))
    add("ln_bwd_kernel: dx fp32 + 1 masked bf16 plane, dgamma / dbeta / dbias, M=131072 D=768", (3 * 4 + 2 * P) * M * D + 8 * M, us)
    del x, skip, y, ypl, dx
    n = model.flat_parameters().numel()
    shadow_bytes = int(model._shadow.numel()) if getattr(model, "_shadow", None) is not None else 0
    us = time_us(lambda: opt.step(), reps=5)
    add("adam_shadow_kernel: fused Adam over the parameter arena + bf16 shadow refresh", 7 * 4 * n + shadow_bytes, us)
    return out


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        p["source"] = "measured (MEASURED_PEAKS.json)"
        return p
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def gemm_roofline(pkg, torch, dev, M, planes):
    """FFN1 of an encoder layer, Y[M,2048] = relu(X[M,768] W^T + b) -> bf16 planes: the most frequent GEMM class of
    the step.  Timed alone: 24 launches captured in one CUDA graph (no host launch gaps), 8 operand sets
    rotated so consecutive launches read different weights, CUDA events on the stream the graph runs on."""
    ops = pkg.ops
    N, K, sets, rep, replays = 2048, 768, 8, 24, 10
    g = torch.Generator(device=dev).manual_seed(1)
    A = [ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), planes) for _ in range(sets)]
    W = [ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, planes) for _ in range(sets)]
    Y = [ops.PlanesTensor.empty(M, N, planes, dev) for _ in range(sets)]
    bias = torch.zeros(N, device=dev)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(sets):  # warm-up: tensor maps, code
            ops.gemm(A[i], W[i], M, N, bias=bias, relu=True, out_planes=Y[i])
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            for i in range(rep):
                ops.gemm(A[i % sets], W[i % sets], M, N, bias=bias, relu=True, out_planes=Y[i % sets])
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(replays):
            gr.replay()
        e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (rep * replays)
    nprod = {1: 1, 2: 3, 3: 6}[planes]
    flops = 2.0 * M * N * K  # algorithmic: one product of the fp32 values, whatever the plane count
    peaks = load_peaks()
    ach = flops / (us * 1e-6) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "gemm_ffn1_traffic.json")  # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(f"planes{planes}", {}).get("dram_bytes_per_launch")
    return {"bound": "tensor", "kernel": f"gemm_kernel<BLOCK_N,K-major,K-major> FFN1 {M}x{N}x{K}, {nprod} plane product(s)",
            "achieved": ach, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": ach / peaks["bf16_tflops"],
            "us_per_launch": us, "traffic": traffic, "peak_source": peaks.get("source"),
            "note": "kernel timed alone (graph replay of 24 launches, CUDA events on the launch stream) -> burst peak"}


def cpu_baseline():
    """Oracle port on the host cores, bounded sample: B = 8 dialogues x 33 (configs[0]), ~10-30 s."""
    step, cores = cpu_port_runner(8, 33)
    step()
    t0 = time.perf_counter()
    n = 0
    while n < 3 or (time.perf_counter() - t0 < 10.0 and n < 40):
        step()
        n += 1
    dt = time.perf_counter() - t0
    return {"value": 8 * 33 * n / dt, "unit": "utterances/s", "cores": cores, "kind": "port",
            "sample": f"{n} full train steps (fwd+CE+bwd+Adam, dropout 0.4, fp32) on 8 dialogues x 33 utterances (configs[0]); "
                      f"{dt / n:.2f} s/step"}


if __name__ == "__main__":
    main()
