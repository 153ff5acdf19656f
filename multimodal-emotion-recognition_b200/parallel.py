"""Data parallelism by dialogue: one process per GPU, a full replica each, gradients all-reduced over
NCCL (NVLink 5 / NVSwitch) in buckets that are launched while the rest of the backward is still running.

The reference has no multi-GPU path at all (src/train.py:20 pins cuda:0); this is the one exchange step the hot path
needs.  The model's backward (one C call) reports finished gradient ranges of the flat arena through a callback
(`m2f_plan_backward_bucketed`); each range is all-reduced asynchronously on NCCL's own stream, which waits for exactly
the kernels enqueued so far, and the compute stream joins all of them at the end of the backward.
"""
from __future__ import annotations

import ctypes
import os
import warnings
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of `n_items` dialogues for `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class GradBucketReducer:
    """All-reduces ranges of a flat gradient tensor as they are reported.  Device-agnostic (gloo on CPU tensors in the
    tests, NCCL on the GPU): `on_bucket(lo, hi)` launches an async all-reduce of flat[lo:hi]; `finish()` makes the
    current stream (or the host, for gloo) wait for all of them and applies the 1/world averaging if requested."""

    def __init__(self, get_flat: Callable[[], torch.Tensor], group=None, average: bool = True):
        self.get_flat = get_flat
        self.group = group
        self.world = dist.get_world_size(group)
        self.average = average
        self.pending: List[Tuple[object, int, int]] = []
        self.log: List[Tuple[int, int]] = []  # ranges reduced in the last step (tests, bench reporting)
        self.bytes_last_step = 0

    def begin_step(self):
        self.log = []
        self.bytes_last_step = 0

    def on_bucket(self, lo: int, hi: int):
        if hi <= lo:
            return
        flat = self.get_flat()
        chunk = flat[lo:hi]
        if self.average and flat.is_cuda:
            work = dist.all_reduce(chunk, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        else:
            work = dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.pending.append((work, lo, hi))
        self.log.append((lo, hi))
        self.bytes_last_step += (hi - lo) * flat.element_size()

    def finish(self):
        flat = self.get_flat()
        for work, lo, hi in self.pending:
            work.wait()  # NCCL: the current stream waits (no host block); gloo: host wait
            if self.average and not flat.is_cuda:
                flat[lo:hi].div_(self.world)
        self.pending = []


class PeerGradExchange:
    """The same contract as `GradBucketReducer`, over NVLink / NVSwitch peer memory instead of NCCL: the model's gradient
    arena is placed in a symmetric allocation (torch.distributed._symmetric_memory: every rank's copy mapped on every
    GPU, plus the NVSwitch multicast mapping when the fabric has one) and each reported range is reduced in place by
    `m2f_grad_allreduce_peer` (csrc/exchange.cu) -- one kernel of a few CTAs whose reduction runs inside the switch
    (`multimem.ld_reduce` / `multimem.st`), launched on the stream the backward reports the range on.  torch is used for
    the allocation and the handle exchange only.

    Two ways to drive it: `in_graph=True` (default) hands the group to the model, whose backward then launches the exchange
    kernels itself, captured in its CUDA graph (m2f_plan_set_exchange: no host work per range); `in_graph=False` launches
    them from the bucket callback like the NCCL reducer does.

    Environment: M2F_EXCHANGE_MC=0 forces the peer load / store variant, M2F_EXCHANGE_CTAS sets the CTAs per call,
    M2F_EXCHANGE_SHAPE=threads,unroll their shape (one of the pairs m2f_exchange_tune accepts),
    M2F_EXCHANGE_IN_GRAPH=0 selects the callback mode."""

    def __init__(self, model, group=None, average: bool = True, ctas: int = 0, in_graph: Optional[bool] = None, wire: Optional[str] = None):
        """`wire`: "fp32" (default, or $M2F_EXCHANGE_WIRE) or "bf16".  bf16 halves the NVLink bytes of the exchange (the reduced
        gradient then carries two bf16 roundings, relative 2^-9) at the price of a pack and an unpack pass over local memory:
        measured on 2 x B200 it does not pay (profiles/r2_exchange_sweep_2gpu.txt), so it is opt-in."""
        import torch.distributed._symmetric_memory as symm_mem

        self.model = model
        wire = (wire or os.environ.get("M2F_EXCHANGE_WIRE") or "fp32").lower()
        if wire not in ("fp32", "bf16"):
            raise ValueError(f"exchange wire format must be fp32 or bf16, got {wire!r}")
        self.in_graph = (os.environ.get("M2F_EXCHANGE_IN_GRAPH", "1") != "0") if in_graph is None else bool(in_graph)

        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        if self.world > _lib.XCHG_MAX_RANKS:
            raise RuntimeError(f"peer gradient exchange supports up to {_lib.XCHG_MAX_RANKS} ranks (one NVSwitch domain)")
        arena = model.flat_parameters()
        if not arena.is_cuda:
            raise RuntimeError("peer gradient exchange needs the model on the GPU")
        n = arena.numel()
        self.signal_off = (n * 4 + 127) // 128 * 128
        self.wire_off = self.signal_off + _lib.XCHG_SIGNAL_BYTES  # bf16 staging copy of the gradients (2 bytes per float)
        total_floats = (self.wire_off + (2 * n + 15) // 16 * 16) // 4 if wire == "bf16" else (self.signal_off + _lib.XCHG_SIGNAL_BYTES) // 4
        with torch.cuda.device(arena.device):
            self.buf = symm_mem.empty(total_floats, dtype=torch.float32, device=arena.device)
            self.buf.zero_()
            self.handle = symm_mem.rendezvous(self.buf, self.group)
            torch.cuda.synchronize()
        dist.barrier(self.group)  # every rank's barrier words are zero before anybody signals
        h = self.handle
        mc = int(h.multicast_ptr) if os.environ.get("M2F_EXCHANGE_MC", "1") != "0" else 0
        ptrs = [int(x) for x in h.buffer_ptrs]
        if int(h.world_size) != self.world or ptrs[int(h.rank)] != self.buf.data_ptr():
            raise RuntimeError("symmetric-memory rendezvous returned an inconsistent handle")
        self.pg = _lib.PeerGroup()
        self.pg.mc_base = mc or None
        for r, x in enumerate(ptrs):
            self.pg.peer_base[r] = x
        self.pg.signal_off_bytes = self.signal_off
        self.pg.rank, self.pg.world = int(h.rank), self.world
        self.multicast = bool(mc)
        self.wire = wire if (wire == "fp32" or (mc and self.world > 1)) else "fp32"  # the bf16 wire needs the multicast mapping
        self.pg.wire_off_bytes = self.wire_off if self.wire == "bf16" else 0
        self.pg.wire_bf16 = 1 if self.wire == "bf16" else 0
        self.scale = 1.0 / self.world if average else 1.0
        self.ctas = int(os.environ.get("M2F_EXCHANGE_CTAS", ctas))
        shape = os.environ.get("M2F_EXCHANGE_SHAPE")  # "threads,unroll" of the exchange CTAs (m2f_exchange_tune); every rank the same
        if shape:
            t_, u_ = (int(x) for x in shape.split(","))
            _lib.check(_lib.lib().m2f_exchange_tune(t_, u_), "exchange_tune")
        self.grads = self.buf[:n]
        self._dryrun = os.environ.get("M2F_EXCHANGE_DRYRUN", "0") == "1"
        self._ev = torch.cuda.Event()
        self._pending = False
        self.log: List[Tuple[int, int]] = []
        self.bytes_last_step = 0
        model.adopt_grad_arena(self.grads)

    def attach(self, bucket_floats: int):
        """Called by DataParallelM2FNet once the bucket size is known."""
        if self.in_graph:
            self.model.set_exchange(self.pg, self.scale, self.ctas, bucket_floats)
            self.model.set_grad_hooks(None, 0, None)
        else:
            self.model.set_exchange(None)

    @property
    def calls_per_step(self) -> int:
        return int(getattr(self.model, "exchange_ranges", 0)) if self.in_graph else len(self.log)

    def begin_step(self):
        self.log = []
        self.bytes_last_step = 0

    def on_bucket(self, lo: int, hi: int):
        if hi <= lo:
            return
        st = torch.cuda.current_stream()  # the stream the range is final on (model._launch_backward sets it)
        if self._dryrun:  # timing experiments only: the ranges are reported but not exchanged
            self.log.append((lo, hi))
            return
        _lib.check(_lib.lib().m2f_grad_allreduce_peer(ctypes.byref(self.pg), lo, hi, self.scale, self.ctas,
                                                      _lib.XCHG_WIRE_BF16 if self.wire == "bf16" else 0, st.cuda_stream),
                   "grad_allreduce_peer")
        self._ev.record(st)
        self._pending = True
        self.log.append((lo, hi))
        self.bytes_last_step += (hi - lo) * 4

    def finish(self):
        if self._pending:  # the ranges are reduced one after the other on one stream: the last event covers them all
            torch.cuda.current_stream().wait_event(self._ev)
            self._pending = False


class DataParallelM2FNet(torch.nn.Module):
    """Wraps a (CUDA) `M2FNet` for data-parallel training: broadcasts rank 0's parameters, then overlaps the bucketed
    gradient all-reduce with the backward.  `forward` is the wrapped model's; `train.train()` of the reference can be
    called with this object as `model` (it only uses __call__, train/eval, parameters, state_dict).

    `exchange`: "peer" = the NVSwitch peer-memory kernel (`PeerGradExchange`), "nccl" = bucketed NCCL all-reduce,
    "auto" (default, or $M2F_EXCHANGE) = peer when the symmetric allocation can be set up on every rank, else nccl."""

    def __init__(self, model, group=None, bucket_mb: Optional[float] = None, average: bool = True, exchange: Optional[str] = None,
                 wire: Optional[str] = None):
        super().__init__()
        self.module = model
        self.group = group
        flat = model.flat_parameters()
        exchange = (exchange or os.environ.get("M2F_EXCHANGE", "auto")).lower()
        if exchange not in ("auto", "peer", "nccl"):
            raise ValueError(f"exchange must be auto, peer or nccl, got {exchange!r}")
        if exchange == "peer" and not flat.is_cuda:
            raise RuntimeError("peer gradient exchange needs the model on the GPU (there is no CPU path)")
        dist.broadcast(flat, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        self.reducer = None
        if exchange != "nccl" and flat.is_cuda:
            # Collective-safe choice of the path: (1) every rank probes, WITHOUT any collective, whether it could take part
            # (the private symmetric-memory module imports, the device supports it); (2) the ranks vote with one all-reduce;
            # (3) only if all can, they enter the set-up (its rendezvous / barrier are collectives every rank then reaches);
            # (4) a second vote catches a rank whose set-up raised after the rendezvous.  A rank never blocks in a collective
            # its peers skipped.
            err = None
            try:
                import torch.distributed._symmetric_memory as _sm  # (private API: allocation + handle exchange only)
                with torch.cuda.device(flat.device):
                    _probe = _sm.empty(256, dtype=torch.float32, device=flat.device)  # local allocation, no collective
                del _probe
                can = 1
            except Exception as e:  # torch built without symmetric memory
                err, can = e, 0
            vote = torch.tensor([can], device=flat.device)
            dist.all_reduce(vote, op=dist.ReduceOp.MIN, group=group)
            if int(vote.item()) == 1:
                try:
                    self.reducer = PeerGradExchange(model, group, average, wire=wire)
                except Exception as e:  # no symmetric memory on this box / fabric
                    err = e
                ok = torch.tensor([0 if self.reducer is None else 1], device=flat.device)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)  # all ranks take the same path
                can = int(ok.item())
            else:
                can = 0
            if can == 0:
                if exchange == "peer":
                    raise RuntimeError(f"peer gradient exchange unavailable: {err!r}")
                if self.reducer is not None:
                    model.set_exchange(None)
                    model.adopt_grad_arena(None)
                    self.reducer = None
                if dist.get_rank(group) == 0:
                    warnings.warn(f"peer gradient exchange unavailable ({err!r}); using NCCL buckets")
        if self.reducer is None:
            self.reducer = GradBucketReducer(model.flat_grads, group, average)
        self.exchange = "peer" if isinstance(self.reducer, PeerGradExchange) else "nccl"
        if bucket_mb is None:  # NCCL kernels compete for SMs: few large buckets; the peer kernel is light: finer overlap
            # peer: 32 MB ranges (9 calls per step at the default config) measured best once the optimizer update runs behind the
            # exchange (profiles/r2_exchange_sweep_2gpu.txt: 16 MB 4.19 ms, 32 MB 4.02 ms, 8 MB 4.22 ms at 2 GPUs)
            bucket_mb = 32.0 if self.exchange == "peer" else 128.0
        self.bucket_mb = float(bucket_mb)
        model.set_grad_hooks(self._on_bucket, int(self.bucket_mb * (1 << 20) / 4), self.reducer.finish)
        if self.exchange == "peer":
            self.reducer.attach(int(self.bucket_mb * (1 << 20) / 4))
        self._first = True

    def _on_bucket(self, lo: int, hi: int):
        if self._first:
            self.reducer.begin_step()
            self._first = False
        self.reducer.on_bucket(lo, hi)

    def forward(self, text, audio, mask):
        self._first = True
        return self.module(text, audio, mask)

    def state_dict(self, *a, **k):
        return self.module.state_dict(*a, **k)

    def load_state_dict(self, *a, **k):
        return self.module.load_state_dict(*a, **k)
