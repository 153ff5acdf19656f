"""bench.py -- M2FNet train utterances/sec on N x B200 (BASELINE.json metric), one JSON line on rank 0.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
    python bench.py --impl reference ...      # the UNMODIFIED reference (oracle/_ref) through its own train() on the host cores

A "step" = one full training step of the reference's train() body (src/train.py:222-233) on one batch:
forward (dropout 0.4, train mode) -> label-smoothed CE -> backward -> Adam(lr 5e-5, wd 0.01).
Workload = BASELINE.json configs[1]: text_base(768) + audio_wav2vec2(768), 64 dialogues x 33 utterances per GPU, 7 classes.

Headline precision = "bf16x2f" (bf16 operands; two planes in the forward, one in the backward): the cheapest mode whose
logits AND gradients meet BASELINE.json's 2e-2 under SURVEY.md 8c's protocol (tests/test_gpu_protocol.py).  `modes` carries
the other precision modes (plain bf16, fp32-parity) measured in the same run.

`value` times device-resident inputs; `e2e` goes through the public API with pinned HOST batches (H2D inside the timed
region) and the loss read back every step; `e2e.reference_train_loop` is the reference's own train() function, unchanged,
driving the drop-in.  `extra_workloads` = BASELINE.json configs[3] (100k-dialogue epoch through the on-device loader) and
configs[4] (B = 512, S = 256 inference through the test() body).  `context.torch_eager_b200` = the unmodified reference
module run by stock PyTorch eager on the same B200.
"""
from __future__ import annotations

import argparse
import importlib
import itertools
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
FLOP_FWD_PER_UTT_S256 = 185.6e6
DTYPE_OF = {"bf16x2f": "bf16", "bf16": "bf16", "bf16x2": "bf16", "fp32": "f32 (3 x bf16 planes)"}
MODE_NOTE = {"bf16x2f": "bf16 operands, 2 planes forward / 1 plane backward (5 plane products per Linear); meets 2e-2 on logits and gradients",
             "bf16": "bf16 operands, 1 plane everywhere (3 products per Linear); logits meet 2e-2, gradients ~2.5e-2: inference mode",
             "bf16x2": "bf16 operands, 2 planes everywhere (9 products per Linear)",
             "fp32": "3 bf16 planes everywhere (18 products per Linear); 1e-4 parity mode"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="bf16x2f", choices=sorted(DTYPE_OF))
    ap.add_argument("--batch", type=int, default=64, help="dialogues per GPU")
    ap.add_argument("--seq", type=int, default=33)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-modes", action="store_true", help="skip the other precision modes")
    ap.add_argument("--no-extra", action="store_true", help="skip extra_workloads (c4_epoch, c5_infer)")
    ap.add_argument("--no-context", action="store_true", help="skip the stock-PyTorch-eager-on-B200 context numbers")
    ap.add_argument("--no-rooflines", action="store_true")
    ap.add_argument("--c4-dialogues", type=int, default=100000)
    ap.add_argument("--no-overlap-adam", action="store_true", help="data parallel: update in optimizer.step() only, not behind the exchange")
    ap.add_argument("--bucket-mb", type=float, default=None, help="gradient bucket size (default: 16 for the peer exchange, 128 for NCCL)")
    ap.add_argument("--wire", default=None, choices=["fp32", "bf16"], help="wire format of the peer gradient exchange (default fp32)")
    ap.add_argument("--exchange", default=None, choices=["auto", "peer", "nccl"],
                    help="gradient exchange at N > 1: NVSwitch peer-memory kernel or NCCL buckets (default: $M2F_EXCHANGE or auto)")
    return ap.parse_args()


def workload_config(args, world, dialogues=None):
    b = args.batch if dialogues is None else dialogues
    return {"workload": "M2FNet text_base(768)+audio_wav2vec2(768), 6+6 encoder layers, 5 fusion layers, 7 classes; "
                        f"{b} dialogues x {args.seq} utterances per GPU; train step = fwd + CE(ls 0.1) + bwd + Adam; dropout 0.4",
            "dialogues_per_gpu": b, "utterances_per_dialogue": args.seq, "global_dialogues": b * world,
            "parallelism": f"dp{world}", "precision_mode": args.precision, "precision_note": MODE_NOTE[args.precision],
            "l2": "16 distinct device-resident batches (208 MB) cycled + >1.5 GB of weight/gradient/Adam traffic per step: "
                  "working set exceeds the 126 MB L2, no explicit flush"}


# ------------------------------------------------------------------------------------------ the reference on the host cores
def host_threads():
    """All host cores, whatever the launcher exported (torchrun sets OMP_NUM_THREADS=1 for its workers)."""
    import torch
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0)) or n
    except Exception:
        pass
    torch.set_num_threads(n)
    return torch.get_num_threads()


class ListLoader:
    """What train() needs of a DataLoader: iteration and len()."""

    def __init__(self, batches):
        self.batches = batches

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        return iter(self.batches)


def reference_cpu_runner(B, S):
    """Returns (run(n_steps) -> seconds, cores, kind, detail).  kind = "reference": the unmodified reference module
    (oracle/_ref/model.py) driven by the reference's own train() (oracle/_ref/train.py:217-243) over an in-memory loader, with
    the criterion and optimizer src/train.py:50,56 build; "port": the oracle restatement (only where oracle/_ref is absent)."""
    import torch
    from oracle import m2fnet_oracle as O
    from oracle import ref_harness as RH

    cores = host_threads()
    batch = O.synth_batch(B, S, 768, 768, seed=1234)
    if RH.available():
        train_mod, model_mod = RH.load_reference()
        torch.manual_seed(0)
        model = model_mod.M2FNet(RH.reference_config())
        crit = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)   # src/train.py:50
        opt = torch.optim.Adam(model.parameters(), lr=5e-5, weight_decay=0.01)   # src/train.py:56, config.yaml solver
        dev = torch.device("cpu")
        if train_mod is not None:
            def run(n):
                t0 = time.perf_counter()
                train_mod.train(model, ListLoader([batch] * n), crit, opt, 0, False, dev)
                return time.perf_counter() - t0
            return run, cores, "reference", "unmodified src/model.py driven by the reference's own train() (src/train.py:217-243)"

        def run(n):  # the reference's module with its loop body restated (train.py could not be imported)
            model.train()
            t0 = time.perf_counter()
            for _ in range(n):
                opt.zero_grad()
                out = model(batch["text"], batch["audio"], batch["padding_mask"])
                loss = crit(out.permute(0, 2, 1), batch["emotion"])
                loss.backward()
                opt.step()
                loss.item()
            return time.perf_counter() - t0
        return run, cores, "reference", "unmodified src/model.py, loop body of src/train.py:222-233 restated"
    # no reference sources on this machine: the oracle port (plain-tensor forward, autograd, nn.Dropout-style masks, Adam)
    cfg = O.default_model_cfg()
    params = {k: v.float().requires_grad_(True) for k, v in O.synth_params(cfg).items() if O.canonical_name(k) == k}
    opt = torch.optim.Adam(list(params.values()), lr=5e-5, weight_decay=0.01)
    p = float(cfg.dropout)

    def masks(site, shape):
        return torch.empty(shape).bernoulli_(1.0 - p)

    def run(n):
        t0 = time.perf_counter()
        for _ in range(n):
            opt.zero_grad()
            logits = O.forward(params, cfg, batch["text"], batch["audio"], batch["padding_mask"], masks, p)
            loss = O.ce_loss(logits, batch["emotion"])
            loss.backward()
            opt.step()
            loss.item()
        return time.perf_counter() - t0
    return run, cores, "port", "oracle/m2fnet_oracle.py restatement (oracle/_ref absent)"


def run_reference_arm(args):
    """--impl reference: rank 0 only.  The same workload as the B200 arm (64 dialogues x 33 utterances per step, full train
    step, dropout 0.4, fp32) unless the whole run would take more than ~4 minutes, in which case the per-step sample is cut
    to 8 dialogues -- and the config says so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    S = args.seq
    warm = max(args.warmup, 3)
    B = args.batch
    run, cores, kind, detail = reference_cpu_runner(B, S)
    t1 = run(1)  # first step (allocations, thread pool spin-up): counts as warm-up
    if t1 * (args.steps + warm) > 240.0 and B > 8:
        B = 8
        run, cores, kind, detail = reference_cpu_runner(B, S)
        run(1)
    if warm > 1:
        run(warm - 1)
    dt = run(args.steps)
    val = B * S * args.steps / dt
    sample = f"{B} dialogues x {S} utterances per step, {args.steps} full train steps (fwd + CE + bwd + Adam, dropout 0.4, fp32); {detail}"
    cfg = workload_config(args, 1, dialogues=B)
    cfg["precision_mode"], cfg["precision_note"] = "fp32", "stock PyTorch CPU kernels"
    cfg["l2"] = "n/a (host)"
    cfg["parallelism"] = "host cores"
    if B != args.batch:
        cfg["bounded_sample"] = f"per-step sample cut from {args.batch} to {B} dialogues to keep the run within minutes"
    print(json.dumps({
        "impl": "reference", "metric": "M2FNet train utterances/sec", "value": val, "unit": "utterances/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": warm, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": val, "unit": "utterances/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def cpu_baseline():
    """The reference on the host cores, bounded sample: 8 dialogues x 33 (configs[0]), ~10-30 s."""
    run, cores, kind, detail = reference_cpu_runner(8, 33)
    run(1)
    n, dt = 0, 0.0
    while n < 3 or (dt < 10.0 and n < 40):
        dt += run(1)
        n += 1
    return {"value": 8 * 33 * n / dt, "unit": "utterances/s", "cores": cores, "kind": kind,
            "sample": f"{n} full train steps (fwd+CE+bwd+Adam, dropout 0.4, fp32) on 8 dialogues x 33 utterances (configs[0]); "
                      f"{dt / n:.2f} s/step; {detail}"}


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


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        p["source"] = "measured (MEASURED_PEAKS.json)"
        return p
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def guarded(out, key, fn):
    """Optional evidence must never cost the bench line."""
    try:
        out[key] = fn()
    except Exception as e:  # noqa: BLE001
        out[key] = None
        out[key + "_error"] = repr(e)[:300]


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
    cfg = D.default_model_config()  # src/config.yaml defaults, dropout 0.4

    NB = 16
    host, devb = [], []
    for i in range(NB):
        b = D.synth_batch(B, S, 768, 768, seed=1234 + 97 * rank + i)
        host.append({k: v.pin_memory() for k, v in b.items()})
        devb.append({k: v.to(dev) for k, v in b.items()})
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())

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

    def build(precision, dp):
        torch.manual_seed(0)
        model = pkg.M2FNet(cfg, precision=precision).to(dev).train()
        net = pkg.DataParallelM2FNet(model, bucket_mb=args.bucket_mb, exchange=args.exchange, wire=args.wire) if dp else model
        crit = pkg.FusedCrossEntropyLoss().to(dev)
        opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01, overlap_exchange=not args.no_overlap_adam)

        def step(b):
            opt.zero_grad(set_to_none=True)
            out = net(b["text"], b["audio"], b["padding_mask"])
            loss = crit(out.permute(0, 2, 1), b["emotion"])
            loss.backward()
            opt.step()
            return loss
        return model, net, crit, opt, step

    model, net, crit, opt, step = build(args.precision, world > 1)
    warm = max(args.warmup, 3)

    # ---- value: device-resident inputs
    ms, launches, clocks = timed(lambda i: step(devb[i % NB]), args.steps, warm)
    utt = world * B * S * args.steps
    value = utt / (ms * 1e-3)

    # ---- e2e: pinned host batches -> H2D -> step -> loss read by the host every step (the reference's train() body).
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
        if world == 1:
            guarded(e2e, "reference_train_loop", lambda: unchanged_train_loop(torch, net, crit, opt, host, dev, args.steps, B, S))

    # ---- data parallel: (a) the exchanged gradients are the RIGHT average: one extra step is run twice with the same dropout
    # masks -- once with the exchange detached (local gradients, averaged over the ranks by NCCL) and once through the
    # in-backward exchange; (b) every replica holds bit-identical parameters after the timed steps
    in_sync, xv = None, None
    if world > 1:
        describe_exchange(net)
        xv = {}
        guarded(xv, "check", lambda: verify_exchange(torch, dist, net, model, crit, devb[0]))
        flat = model.flat_parameters()
        chk = torch.stack([flat.double().sum(), flat.double().abs().sum(), flat[::9973].double().square().sum()])
        hi_, lo_ = chk.clone(), chk.clone()
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        in_sync = bool(torch.equal(hi_, lo_)) and bool(torch.isfinite(chk).all())

    # ---- configs[3]: a 100k-dialogue epoch through the on-device loader, sharded by dialogue across the ranks
    extra = {}
    if not args.no_extra:
        guarded(extra, "c4_epoch", lambda: c4_epoch(torch, dist, pkg, D, dev, step, world, rank, args.c4_dialogues, B, barrier))

    # ---- rooflines (rank 0): the dominant kernel class alone, the whole step, the bandwidth-bound kernels beyond L2
    roof = None
    peaks = load_peaks()
    planes_f, planes_b = model.planes, model.planes_bwd
    if rank == 0 and not args.no_rooflines:
        roof = gemm_roofline(pkg, torch, dev, B * S, planes_f, planes_b)
    if rank == 0:
        roof = roof or {}
        step_flops = 3.0 * FLOP_FWD_PER_UTT_S33 * B * S
        nprod = {1: 1, 2: 3, 3: 6}
        executed = FLOP_FWD_PER_UTT_S33 * B * S * (nprod[planes_f] + 2 * nprod[planes_b])
        sec = ms / args.steps * 1e-3
        roof["step"] = {"algorithmic_tflop_per_step_per_gpu": step_flops / 1e12,
                        "achieved_tflops_per_gpu": step_flops / sec / 1e12,
                        "frac_of_sustained_bf16_peak": step_flops / sec / 1e12 / peaks["bf16_tflops_sustained"],
                        "executed_tensor_tflop_per_step_per_gpu": executed / 1e12,
                        "tensor_pipe_frac_of_sustained_peak": executed / sec / 1e12 / peaks["bf16_tflops_sustained"],
                        "note": "algorithmic = 3 x forward flops of the fp32 model (SURVEY.md 8d); executed = plane products actually "
                                f"issued to the tensor pipe ({nprod[planes_f]} forward + 2 x {nprod[planes_b]} backward per Linear); whole "
                                "train step incl. attention / LN / loss / Adam kernels and launch gaps"}

    # free the headline model before the other modes / workloads allocate theirs
    del step, opt, crit, net, model
    torch.cuda.empty_cache()

    modes = None
    if world == 1 and not args.no_modes:
        modes = {args.precision: {"ms_per_step": ms / args.steps, "utterances_per_s": value, "steps": args.steps, "dtype": DTYPE_OF[args.precision],
                                  "note": MODE_NOTE[args.precision], "headline": True}}
        for prec in ("bf16", "fp32"):
            if prec == args.precision:
                continue

            def one(prec=prec):
                m2, n2, c2, o2, s2 = build(prec, False)
                k = max(5, args.steps // 2)
                ms2, _, _ = timed(lambda i: s2(devb[i % NB]), k, 3)
                del m2, n2, c2, o2, s2
                torch.cuda.empty_cache()
                return {"ms_per_step": ms2 / k, "utterances_per_s": B * S * k / (ms2 * 1e-3), "steps": k, "dtype": DTYPE_OF[prec], "note": MODE_NOTE[prec]}
            guarded(modes, prec, one)

    if rank == 0 and world == 1 and not args.no_rooflines:
        guarded(roof, "hbm_kernels", lambda: hbm_kernel_rooflines(pkg, torch, dev, cfg))
    if rank == 0 and world == 1 and not args.no_extra:
        guarded(extra, "c5_infer", lambda: c5_infer(torch, pkg, D, dev, cfg))

    context = None
    if rank == 0 and world == 1 and not args.no_context:
        context = {}
        guarded(context, "torch_eager_b200", lambda: torch_eager_b200(torch, dev, devb[0], B, S))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = {}
        guarded(cpu, "r", cpu_baseline)
        cpu = cpu.get("r") or {"value": None, "error": cpu.get("r_error")}

    if rank == 0:
        cfg_out = workload_config(args, world)
        if world > 1:
            cfg_out["exchange"] = xchg_info["desc"]
            cfg_out.update(xchg_info["cfg"])
            cfg_out["replicas_in_sync_after_run"] = in_sync
            cfg_out["exchange_verified"] = bool(xv and xv.get("check") and xv["check"]["ok"])
            cfg_out["exchange_check"] = xv.get("check") if xv else None
            if xv and xv.get("check_error"):
                cfg_out["exchange_check_error"] = xv["check_error"]
        out = {"metric": "M2FNet train utterances/sec", "value": value, "unit": "utterances/s", "n_gpus": world,
               "steps": args.steps, "warmup": warm, "ms_per_step": ms / args.steps, "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": DTYPE_OF[args.precision],
               "data": "synthetic", "config": cfg_out, "clocks": clocks, "e2e": e2e,
               "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu, "modes": modes,
               "extra_workloads": extra or None, "context": context}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


xchg_info = {"desc": None, "cfg": {}}


def describe_exchange(net):
    xchg_info["desc"] = ("gradient all-reduce: m2f_grad_allreduce_peer over NVSwitch " +
                         ("multicast (multimem.ld_reduce / multimem.st)" if getattr(net.reducer, "multicast", False) else "peer loads / stores")
                         if net.exchange == "peer" else "gradient all-reduce: NCCL, async buckets behind the backward")
    xchg_info["cfg"] = {"optimizer_overlap": "Adam of each exchanged range enqueued by the backward behind its compute, gated by the next range's exchange"
                                             if getattr(net.module, "_tail_opt", None) is not None else "optimizer.step() after the last exchange",
                        "bucket_mb": net.bucket_mb, "exchange_calls_per_step": getattr(net.reducer, "calls_per_step", len(net.reducer.log)),
                        "exchange_wire": getattr(net.reducer, "wire", "fp32")}
    if net.exchange == "peer":
        xchg_info["cfg"]["exchange_launch"] = "kernels captured in the backward's CUDA graph" if net.reducer.in_graph else "bucket callbacks"


def verify_exchange(torch, dist, net, model, crit, b):
    """One extra step, run twice with identical dropout masks (the step counter is rewound): (1) exchange detached -> this
    rank's LOCAL gradients, averaged over the ranks by an independent NCCL all-reduce; (2) through the data-parallel path as
    benchmarked.  The exchanged arena must equal the NCCL mean (fp32 wire: the 1-D gradients are accumulated with atomics, so two
    runs of the same backward differ in the last bits -- tolerance 1e-5 relative L2 per range class; bf16 wire: two bf16
    roundings per element -- 4e-3)."""
    world = dist.get_world_size()
    saved_x, saved_hooks = model._exchange, (model._bucket_hook, model._bucket_floats, model._done_hook)
    saved_opt = model._tail_opt
    model.set_tail_optimizer(None)  # the check runs two backwards on the same weights: no updates in between

    def fwd_bwd(wrapper):
        model.zero_grad(set_to_none=True)
        model._grads_cleared = False
        out = wrapper(b["text"], b["audio"], b["padding_mask"])
        loss = crit(out.permute(0, 2, 1), b["emotion"])
        loss.backward()
        torch.cuda.synchronize()
        return model.flat_grads().detach().clone()

    step0 = model._step
    model.set_exchange(None)
    model.set_grad_hooks(None, 0, None)
    g_local = fwd_bwd(model)
    ref = g_local.clone()
    dist.all_reduce(ref, op=dist.ReduceOp.SUM)
    ref /= world
    model._step = step0  # same (seed, step) -> same Philox masks
    if saved_x is not None:
        model.set_exchange(*saved_x)
    model.set_grad_hooks(*saved_hooks)
    g_x = fwd_bwd(net)
    if saved_opt is not None:
        model.set_tail_optimizer(saved_opt)
    off = model._off_1d
    num2 = (g_x[:off].double() - ref[:off].double()).norm().item()
    den2 = ref[:off].double().norm().item()
    num1 = (g_x[off:].double() - ref[off:].double()).norm().item()
    den1 = ref[off:].double().norm().item()
    local_vs_mean = (g_local.double() - ref.double()).norm().item() / max(ref.double().norm().item(), 1e-30)
    e2d, e1d = num2 / max(den2, 1e-30), num1 / max(den1, 1e-30)
    t = torch.tensor([e2d, e1d], device=g_x.device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2d, e1d = t.tolist()
    tol = 4e-3 if getattr(net.reducer, "wire", "fp32") == "bf16" else 1e-5
    return {"ok": bool(e2d <= tol and e1d <= tol and local_vs_mean > 1e-3), "rel_err_2d_weights": e2d, "rel_err_1d_params": e1d,
            "local_vs_mean": local_vs_mean, "tolerance": tol,
            "how": "exchanged gradient arena vs NCCL all_reduce(SUM)/world of the per-rank local gradients of the same step (same dropout "
                   "masks); max over ranks; local_vs_mean > 0 shows the ranks' gradients really differed"}


def unchanged_train_loop(torch, net, crit, opt, host, dev, steps, B, S):
    """The reference's own train() function (oracle/_ref/train.py, src/train.py:217-243), byte-for-byte unchanged, driving
    the drop-in: pageable->device `.to(device)` per tensor, optimizer.zero_grad(), forward, criterion, backward, step,
    `loss.item()` after the optimizer step -- every host sync the reference has."""
    from oracle import ref_harness as RH
    if not RH.available():
        return {"unavailable": "oracle/_ref absent"}
    train_mod, _ = RH.load_reference()
    if train_mod is None:
        return {"unavailable": "the reference's train.py could not be imported"}
    NB = len(host)
    warm = ListLoader([host[i % NB] for i in range(3)])
    train_mod.train(net, warm, crit, opt, 0, False, dev)
    torch.cuda.synchronize()
    loader = ListLoader([host[i % NB] for i in range(steps)])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    mean_loss = train_mod.train(net, loader, crit, opt, 1, False, dev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    return {"value": B * S * steps / (ms * 1e-3), "unit": "utterances/s", "ms_per_step": ms / steps, "mean_loss": float(mean_loss),
            "note": "reference train() unchanged (tqdm, per-tensor .to(device) from pinned host memory, loss.item() after optimizer.step()); "
                    "criterion = FusedCrossEntropyLoss, optimizer = FusedAdam passed in as its arguments"}


def c4_epoch(torch, dist, pkg, D, dev, step, world, rank, n_dialogues, B, barrier):
    """BASELINE.json configs[3]: one epoch over `n_dialogues` synthetic dialogues with MELD-like lengths, sharded contiguously by
    dialogue across the ranks (parallel.shard_range), embedding tables resident in HBM (DeviceDialogueStore), batches grouped by
    length into the (8, 16, 24, 33) buckets (LengthBucketSampler) and assembled on the device (DeviceLoader ->
    m2f_collate_dialogues), full train step per batch.  Reference: src/dataset.py:32-65,71-89 (what it replaces)."""
    import numpy as np
    g = torch.Generator().manual_seed(4321)
    L_all = torch.exp(torch.randn(n_dialogues, generator=g) * 0.6 + 2.1).round().clamp(1, 33).long()
    lo, hi = pkg.shard_range(n_dialogues, rank, world)
    hi = lo + n_dialogues // world
    Ls = L_all[lo:hi]
    rows = int(Ls.sum())
    gd = torch.Generator(device=dev).manual_seed(99 + rank)
    text = torch.randn(rows, 768, generator=gd, device=dev) * 0.63 + 0.025
    audio = torch.randn(rows, 768, generator=gd, device=dev) * 0.233
    probs = torch.tensor(D.MELD_CLASS_P)
    labels = torch.multinomial(probs, rows, replacement=True, generator=g)
    did = torch.repeat_interleave(torch.arange(len(Ls)), Ls).numpy()
    starts = np.concatenate(([0], np.cumsum(Ls.numpy())[:-1]))
    uid = np.arange(rows) - np.repeat(starts, Ls.numpy())
    store = D.DeviceDialogueStore(text, audio, labels, did, uid, dev)
    buckets = (8, 16, 24, 33)
    sampler = D.LengthBucketSampler(store.lengths, B, buckets=buckets, shuffle=True, seed=7, drop_last=True)
    batches = list(sampler)
    n_steps = len(batches)
    if world > 1:  # every rank runs the same number of steps (the exchange is collective)
        t = torch.tensor([n_steps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        n_steps = int(t.item())
    batches = batches[:n_steps]
    # warm-up: two steps per bucket shape (plan creation, graph capture) on the B shortest dialogues (they fit every
    # bucket), identical count on every rank
    short = np.argsort(store.lengths, kind="stable")[:B]
    for Sb in buckets:
        for _ in range(2):
            step(store.collate(short, max(Sb, int(store.lengths[short].max()))))
    barrier()
    loader = D.DeviceLoader(store, batches)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for b in loader:
        step(b)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    valid = int(sum(int(store.lengths[c].sum()) for c, _ in batches))
    padded = int(sum(len(c) * s_ for c, s_ in batches))
    stats = torch.tensor([ms, valid, padded, len(batches) * B], device=dev, dtype=torch.float64)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        stats[0] = mx[0]
    ms, valid, padded, served = stats.tolist()
    return {"config": "configs[3]: synthetic epoch, MELD-like dialogue lengths (mean ~9.8, max 33), dialogues sharded contiguously across ranks",
            "dialogues_in_epoch": n_dialogues, "dialogues_served": int(served), "steps_per_rank": n_steps, "seconds": ms * 1e-3,
            "valid_utterances_per_s": valid / (ms * 1e-3), "padded_utterances_per_s": padded / (ms * 1e-3),
            "padding_fraction": 1.0 - valid / padded, "dialogues_per_batch": B, "length_buckets": list(buckets),
            "pipeline": "DeviceDialogueStore (tables in HBM) + LengthBucketSampler + DeviceLoader (m2f_collate_dialogues); one plan + CUDA graph per bucket"}


def c5_infer(torch, pkg, D, dev, cfg):
    """BASELINE.json configs[4]: the test() body (src/test.py:58-71) on 256-utterance dialogues, batch 512: eval forward +
    argmax + per-batch accuracy / weighted-F1 -- the metrics on the device (m2f_batch_metrics), one host read at the end."""
    B, S = 512, 256
    torch.manual_seed(0)
    model = pkg.M2FNet(cfg, precision="bf16").to(dev).eval()
    metrics = D.BatchMetrics(7, dev)
    batches = [{k: v.to(dev) for k, v in D.synth_batch(B, S, 768, 768, seed=77 + i).items()} for i in range(2)]
    L = pkg._lib.lib()

    def body(b):
        out = model(b["text"], b["audio"], b["padding_mask"])
        metrics.update(out, b["emotion"])

    steps = 5
    with torch.inference_mode():
        for i in range(2):
            body(batches[i % 2])
        torch.cuda.synchronize()
        l0 = int(L.m2f_launch_count())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            body(batches[i % 2])
        e1.record()
        torch.cuda.synchronize()
        acc, f1 = metrics.result()
    ms = e0.elapsed_time(e1) / steps
    ws_gb = sum(p.ws_bytes for p in model._plans.values()) / 1e9
    flops = FLOP_FWD_PER_UTT_S256 * B * S
    peaks = load_peaks()
    res = {"config": "configs[4]: inference, 512 dialogues x 256 utterances per batch (M = 131072), eval forward + on-device metrics",
           "precision_mode": "bf16", "ms_per_batch": ms, "utterances_per_s": B * S / (ms * 1e-3), "workspace_gb": ws_gb,
           "achieved_tflops": flops / (ms * 1e-3) / 1e12, "frac_of_sustained_bf16_peak": flops / (ms * 1e-3) / 1e12 / peaks["bf16_tflops_sustained"],
           "gpu_launches_per_batch": (int(L.m2f_launch_count()) - l0) / steps, "accuracy": acc, "weighted_f1": f1}
    del model, batches
    torch.cuda.empty_cache()
    return res


def torch_eager_b200(torch, dev, b, B, S):
    """Stock PyTorch eager running the UNMODIFIED reference module (oracle/_ref/model.py) on this B200: the same batch, the same
    full step (zero_grad, forward, CrossEntropyLoss, backward, torch.optim.Adam), fp32 with TF32 off (the reference's
    numerics) and under bf16 autocast.  BASELINE.md 4.4: the real competitor on the same box."""
    from oracle import ref_harness as RH
    if not RH.available():
        return {"unavailable": "oracle/_ref absent"}
    _, model_mod = RH.load_reference()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    out = {"module": "unmodified src/model.py (oracle/_ref), torch " + torch.__version__, "batch": f"{B} dialogues x {S} utterances"}
    for name, autocast in (("fp32_tf32_off", False), ("bf16_autocast", True)):
        torch.manual_seed(0)
        model = model_mod.M2FNet(RH.reference_config()).to(dev).train()
        crit = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)
        opt = torch.optim.Adam(model.parameters(), lr=5e-5, weight_decay=0.01)

        def step():
            opt.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                o = model(b["text"], b["audio"], b["padding_mask"])
                loss = crit(o.permute(0, 2, 1).float(), b["emotion"])
            loss.backward()
            opt.step()
            return loss.item()
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        k = 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / k
        out[name] = {"ms_per_step": ms, "utterances_per_s": B * S / (ms * 1e-3), "steps": k}
        del model, opt
        torch.cuda.empty_cache()
    return out


def hbm_kernel_rooflines(pkg, torch, dev, cfg):
    """The bandwidth-bound kernels against the measured HBM peak, at sizes whose working set exceeds the 126 MB L2
    (at M = 2112 these tensors sit in L2 and the kernels are launch-bound).  Algorithmic bytes as in
    DESIGN.md section 5 / SURVEY.md 8d.  Each timed alone, CUDA events on the launch stream, 10 launches after 2 warm-ups."""
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

    M, Dm, P = 131072, 768, 1
    g = torch.Generator(device=dev).manual_seed(3)
    x = torch.randn(M, Dm, generator=g, device=dev)
    skip = torch.randn(M, Dm, generator=g, device=dev)
    gamma, beta = torch.ones(Dm, device=dev), torch.zeros(Dm, device=dev)
    y = torch.empty(M, Dm, device=dev)
    ypl = ops.PlanesTensor.empty(M, Dm, P, dev)
    stats = torch.empty(M, 2, device=dev)
    us = time_us(lambda: ops.ln_fwd(x, gamma, beta, skip=skip, y=y, y_planes=ypl, drop_p=0.4, seed=1, step=2, site=3, stats=stats))
    add("LayerNorm forward (staged): y = LN(x) + skip, fp32 + 1 bf16 plane (dropout 0.4), M=131072 D=768", (3 * 4 + 2 * P) * M * Dm + 8 * M, us)
    dx = torch.empty(M, Dm, device=dev)
    dgamma, dbeta, dbias = (torch.zeros(Dm, device=dev) for _ in range(3))
    us = time_us(lambda: ops.ln_bwd(skip, x, gamma, stats, dx=dx, dx_planes=ypl, drop_p=0.4, seed=4, step=5, site=6, dgamma=dgamma, dbeta=dbeta,
                                    dbias=dbias))
    add("LayerNorm backward (staged): dx fp32 + 1 masked bf16 plane, dgamma / dbeta / dbias, M=131072 D=768", (3 * 4 + 2 * P) * M * Dm + 8 * M, us)
    del x, skip, y, ypl, dx

    # cross entropy at M = 131072 rows of 7 logits: read logits + targets, write dlogits (M * (2*C*4 + 8) bytes)
    C = 7
    logits = torch.randn(M, C, generator=g, device=dev)
    target = torch.randint(0, C, (M,), device=dev)
    dlog = torch.empty(M, C, device=dev)
    us = time_us(lambda: ops.ce_fwd_bwd(logits, target, None, eps=0.1, grad_scale=1.0, dlogits=dlog))
    add("cross entropy fwd+bwd (label smoothing 0.1, ignore_index -1), M=131072 C=7", M * (2 * C * 4 + 8), us)
    del logits, target, dlog

    # attention, training kernels (S = 33, one CTA per (dialogue, head)) at 2048 dialogues: q, k, v, o of 104 MB each
    Bq, Sq, H = 2048, 33, 8
    Mq = Bq * Sq
    mk = lambda: ops.PlanesTensor.from_float(torch.randn(Mq, Dm, generator=g, device=dev) * 0.5, 1)
    q, k, v, d_o = mk(), mk(), mk(), mk()
    o, dq, dk, dv = (ops.PlanesTensor.empty(Mq, Dm, 1, dev) for _ in range(4))
    pad = torch.zeros(Bq, Sq, dtype=torch.uint8, device=dev)
    us = time_us(lambda: ops.attn_fwd(q, k, v, pad, Bq, Sq, H, o, drop_p=0.4, seed=1, step=1, site=5))
    add("attention forward (S=33, hd=96, dropout 0.4), 2048 dialogues x 8 heads", 4 * Mq * Dm * 2 + Mq, us)
    us = time_us(lambda: ops.attn_bwd(q, k, v, d_o, pad, Bq, Sq, H, dq, dk, dv, drop_p=0.4, seed=1, step=1, site=5))
    add("attention backward (S=33, hd=96, dropout 0.4; P recomputed, O not re-read: 4 reads + 3 writes), 2048 dialogues x 8 heads",
        7 * Mq * Dm * 2 + Mq, us)
    del q, k, v, d_o, o, dq, dk, dv

    # attention, long-dialogue inference kernel (S = 256) at the C5 batch: 805 MB per MHA
    Bl, Sl = 512, 256
    Ml = Bl * Sl
    mk = lambda: ops.PlanesTensor.from_float(torch.randn(Ml, Dm, generator=g, device=dev) * 0.5, 1)
    q, k, v = mk(), mk(), mk()
    o = ops.PlanesTensor.empty(Ml, Dm, 1, dev)
    pad = torch.zeros(Bl, Sl, dtype=torch.uint8, device=dev)
    us = time_us(lambda: ops.attn_fwd(q, k, v, pad, Bl, Sl, H, o), reps=5)
    flops = 4.0 * Sl * Dm * Ml
    add("attention forward, long dialogues (S=256, hd=96, inference; tcgen05 / TMEM / TMA kernel), 512 dialogues x 8 heads", 4 * Ml * Dm * 2 + Ml, us)
    out[-1]["tensor_tflops"] = flops / (us * 1e-6) / 1e12
    out[-1]["note"] = ("103 GFLOP and 33.5 M exponentials per launch; persistent CTAs (one per SM) walk the (dialogue, head) items with the next item's "
                       "K / Q / V in flight under the current item's softmax; what is left is the softmax itself (16 ex2 per clock per SM: 2.1 us of "
                       "the ~6.5 us an item takes) and the two score tiles that have to leave TMEM before the next item's second product")
    del q, k, v, o

    # fused Adam + bf16 shadow refresh over the model arena
    torch.manual_seed(0)
    model = pkg.M2FNet(cfg, precision="bf16x2f").to(dev).train()
    opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
    model._ensure_shadow()
    model._grad_arena = torch.zeros_like(model.flat_parameters())
    n = model.flat_parameters().numel()
    shadow_bytes = int(model._shadow.numel())
    us = time_us(lambda: opt.step(), reps=5)
    add("adam_shadow_kernel: fused Adam over the parameter arena + 2-plane bf16 shadow refresh", 7 * 4 * n + shadow_bytes, us)
    del model, opt
    torch.cuda.empty_cache()
    return out


def gemm_roofline(pkg, torch, dev, M, planes_f, planes_b):
    """The GEMM classes of an encoder layer's FFN1 (the most frequent shape of the step), each timed alone: 24
    launches captured in one CUDA graph (no host launch gaps), 8 operand sets rotated so consecutive launches read different
    operands, CUDA events on the stream the graph runs on.  The headline entry is the forward kernel as the benchmarked
    precision mode launches it; its work is the plane products it is asked for (products x 2MNK)."""
    ops = pkg.ops
    N, K, sets, rep, replays = 2048, 768, 8, 24, 10
    g = torch.Generator(device=dev).manual_seed(1)
    peaks = load_peaks()
    nprod = {1: 1, 2: 3, 3: 6}

    def run(label, P, make, launch):
        bufs = [make(P) for _ in range(sets)]
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for i in range(sets):
                launch(bufs[i])
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                for i in range(rep):
                    launch(bufs[i % sets])
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
        flops = 2.0 * M * N * K * nprod[P]
        ach = flops / (us * 1e-6) / 1e12
        return {"kernel": label, "plane_products": nprod[P], "us_per_launch": us, "achieved": ach, "unit": "TFLOP/s", "peak": peaks["bf16_tflops"],
                "frac": ach / peaks["bf16_tflops"], "algorithmic_tflops_one_product": 2.0 * M * N * K / (us * 1e-6) / 1e12}

    bias = torch.zeros(N, device=dev)

    def mk_fwd(P):
        return (ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), P),
                ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, P), ops.PlanesTensor.empty(M, N, P, dev))

    def mk_dgrad(P):  # dX[M,K] = dH[M,N] . W1[N,K]   (B read MN-major), + skip-path gradient, fp32 out
        return (ops.PlanesTensor.from_float(torch.randn(M, N, generator=g, device=dev), P),
                ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, P), torch.randn(M, K, generator=g, device=dev),
                torch.empty(M, K, device=dev))

    def mk_wgrad(P):  # dW1[N,K] = dH^T[N,M] . X[M,K]  (both read MN-major), fp32 out
        return (ops.PlanesTensor.from_float(torch.randn(M, N, generator=g, device=dev), P),
                ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), P), torch.empty(N, K, device=dev))

    fwd = run(f"gemm_kernel forward FFN1 {M}x{N}x{K}: relu(X W^T + b), dropout 0.4 -> {planes_f} bf16 plane(s)", planes_f, mk_fwd,
              lambda t: ops.gemm(t[0], t[1], M, N, bias=bias, relu=True, drop_p=0.4, seed=1, step=1, site=7, out_planes=t[2]))
    classes = [fwd]
    classes.append(run(f"gemm_kernel dgrad FFN1 {M}x{K}x{N}: dX = dH W + skip -> fp32", planes_b, mk_dgrad,
                       lambda t: ops.gemm(t[0], t[1], M, K, b_mn=True, resid=t[2], out_f32=t[3])))
    classes.append(run(f"gemm_kernel wgrad FFN1 {N}x{K}x{M}: dW = dH^T X -> fp32", planes_b, mk_wgrad,
                       lambda t: ops.gemm(t[0], t[1], N, K, a_mn=True, b_mn=True, out_f32=t[2])))
    if planes_f != 1:
        classes.append(run(f"gemm_kernel forward FFN1 {M}x{N}x{K}, 1 plane (precision mode bf16)", 1, mk_fwd,
                           lambda t: ops.gemm(t[0], t[1], M, N, bias=bias, relu=True, drop_p=0.4, seed=1, step=1, site=7, out_planes=t[2])))
    # the persistent two-accumulator GEMM (gemm_persist.cu) at the C5 inference size: FFN1 of 512 dialogues x 256 utterances, one plane
    try:
        Mi = 131072
        Ai = ops.PlanesTensor.from_float(torch.randn(Mi, K, generator=g, device=dev), 1)
        Wi = ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, 1)
        Yi = ops.PlanesTensor.empty(Mi, N, 1, dev)
        for _ in range(2):
            ops.gemm(Ai, Wi, Mi, N, bias=bias, relu=True, out_planes=Yi)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            ops.gemm(Ai, Wi, Mi, N, bias=bias, relu=True, out_planes=Yi)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 5
        ach = 2.0 * Mi * N * K / (us * 1e-6) / 1e12
        classes.append({"kernel": f"gemm_persist_kernel forward FFN1 {Mi}x{N}x{K} (inference, C5 size): relu(X W^T + b) -> 1 bf16 plane, 2048 tiles over 74 CTA pairs",
                        "plane_products": 1, "us_per_launch": us, "achieved": ach, "unit": "TFLOP/s", "peak": peaks["bf16_tflops_sustained"],
                        "frac": ach / peaks["bf16_tflops_sustained"], "algorithmic_tflops_one_product": ach,
                        "note": "a 0.26 ms launch: measured against the sustained peak; operands (0.2 GB) and output (0.5 GB) exceed L2"})
        del Ai, Wi, Yi
    except Exception as e:  # never lose the bench line over an extra class
        classes.append({"kernel": "gemm_persist_kernel (C5 size)", "error": str(e)[:200]})
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "gemm_ffn1_traffic.json")  # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tpath):
        try:
            with open(tpath) as f:
                traffic = json.load(f).get(f"planes{planes_f}", {}).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    return {"bound": "tensor", "kernel": fwd["kernel"], "achieved": fwd["achieved"], "peak": fwd["peak"], "unit": "TFLOP/s", "frac": fwd["frac"],
            "us_per_launch": fwd["us_per_launch"], "plane_products": fwd["plane_products"], "traffic": traffic, "peak_source": peaks.get("source"),
            "gemm_classes": classes,
            "note": "kernel timed alone (graph replay of 24 launches, CUDA events on the launch stream) -> burst peak; achieved = plane "
                    "products x 2MNK / time, i.e. the tensor-pipe work the kernel is asked for (the 3-MMA factor of SURVEY.md 8d); "
                    "algorithmic_tflops_one_product counts 2MNK once"}


if __name__ == "__main__":
    main()
