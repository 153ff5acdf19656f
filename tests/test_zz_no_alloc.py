"""m2f_ce_fwd_bwd_static: the allocation-free variant of the cross-entropy entry point (scratch in a file-scope device array of
the library, csrc/loss_static.cu) against the existing m2f_ce_fwd_bwd, through the C ABI, twice in a row (the second call finds
the first call's leftovers in the shared scratch), plus the size limit.  No counterpart in the reference; the file is named so
that pytest collects it last."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 262144  # M2F_CE_STATIC_MAX_ROWS (include/m2f_b200.h)


@pytest.fixture(scope="module")
def P(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    return pkg


def make_case(M, Cc, weighted, seed):
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(M, Cc, generator=g) * 2).cuda()
    tgt = torch.randint(0, Cc, (M,), generator=g)
    tgt[torch.rand(M, generator=g) < 0.3] = -1
    w = (torch.rand(Cc, generator=g) + 0.5).cuda() if weighted else None
    return logits, tgt.cuda(), w


def run_both(P, M, Cc, weighted, seed, out, dl_static):
    """`out` and `dl_static` are allocated by the caller BEFORE the call under test: the static entry point gets only inputs and
    outputs."""
    ops = P.ops
    logits, tgt, w = make_case(M, Cc, weighted, seed)
    dl_ref = torch.empty(M, Cc, device="cuda")
    ref = ops.ce_fwd_bwd(logits, tgt, w, eps=0.1, grad_scale=1.0, dlogits=dl_ref)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    ops.ce_fwd_bwd_static(logits, tgt, out, w, eps=0.1, grad_scale=1.0, dlogits=dl_static[:M])
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before  # nothing allocated through torch either
    torch.cuda.synchronize()
    # same formulas, same fp32 arithmetic; only the order of the atomics and exp(z - lse) vs exp(lp) differ
    assert abs(out[0].item() - ref[0].item()) <= 2e-6 * abs(ref[0].item()), (out[0].item(), ref[0].item())
    assert abs(out[1].item() - ref[1].item()) <= 2e-6 * abs(ref[1].item())
    err = (torch.linalg.vector_norm(dl_static[:M].double() - dl_ref.double()) / torch.linalg.vector_norm(dl_ref.double())).item()
    assert err < 5e-6, err
    tref = torch.nn.CrossEntropyLoss(weight=w, ignore_index=-1, label_smoothing=0.1)(logits, tgt)
    assert abs(out[0].item() - tref.item()) < 1e-5


def test_static_scratch_cross_entropy_matches_the_allocating_entry_point_twice_in_a_row(P):
    out = torch.full((2,), float("nan"), device="cuda")
    dl = torch.empty(4096, 7, device="cuda")
    # first call: a large batch fills the per-row scratch; second call: a smaller, differently masked and weighted batch must not
    # see any of it (sums cleared per call, rows rewritten before they are read)
    run_both(P, 4096, 7, False, 71, out, dl)
    run_both(P, 1000, 7, True, 72, out, dl)
    run_both(P, 2112, 7, True, 73, out, dl)


def test_static_scratch_limit_is_an_error_not_a_fault(P):
    L = P._lib.lib()
    out = torch.zeros(2, device="cuda")
    one = torch.zeros(8, device="cuda")
    tgt = torch.zeros(8, dtype=torch.int64, device="cuda")
    rc = L.m2f_ce_fwd_bwd_static(one.data_ptr(), 7, tgt.data_ptr(), None, MAX_ROWS + 1, 7, 0.1, 1.0, out.data_ptr(), None, 0,
                                 C.byref(P._lib.Planes()), None)
    assert rc == -2 and b"M2F_CE_STATIC_MAX_ROWS" in L.m2f_last_error()
    torch.cuda.synchronize()
