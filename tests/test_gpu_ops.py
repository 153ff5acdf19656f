"""Op-level parity of the CUDA kernels, called through the C ABI (ctypes), against fp64 torch expressions
of the same arithmetic (the formulas of oracle/m2fnet_oracle.py).  Tolerances: bf16 planes (P=1) inputs are
exact bf16 values, so the GEMM must match an fp64 product to fp32-accumulation accuracy (1e-5 rel-L2);
P=3 must match the fp64 product of the fp32 inputs to 5e-6 (fp32 accumulation over up to 6 x 2112 terms)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    return pkg.ops


def rel(x, ref):
    x, ref = x.double().flatten(), ref.double().flatten()
    return (torch.linalg.vector_norm(x - ref) / torch.linalg.vector_norm(ref).clamp_min(1e-30)).item()


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, generator=g, device="cuda") * scale


GEMM_SHAPES = [
    # M, N, K
    (2112, 768, 768),    # out-proj / fusion projections (BLOCK_N 64)
    (2112, 2304, 768),   # packed QKV (BLOCK_N 256)
    (2112, 2048, 768),   # FFN1 (BLOCK_N 128)
    (264, 7, 768),       # classifier output at config 1 (ragged M, N < 8)
    (100, 136, 304),     # ragged everything, K not a multiple of 64
]


@pytest.mark.parametrize("planes", [1, 3])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_forward_tn(ops, M, N, K, planes):
    X, W = rnd(M, K, seed=1), rnd(N, K, seed=2, scale=K ** -0.5)
    A, B = ops.PlanesTensor.from_float(X, planes), ops.PlanesTensor.from_float(W, planes)
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(A, B, M, N, out_f32=out)
    ref = A.float().double() @ B.float().double().T
    assert rel(out, ref) < (1e-5 if planes == 1 else 5e-6)
    if planes == 3:  # the planes carry the fp32 value itself
        assert rel(out, X.double() @ W.double().T) < 5e-6


@pytest.mark.parametrize("planes", [1, 3])
@pytest.mark.parametrize("M,N,K", [(2112, 2048, 768), (264, 7, 768), (100, 136, 304)])
def test_gemm_dgrad_b_mn_major(ops, M, N, K, planes):
    """dX[M,K] = dY[M,N] . W[N,K]: contraction over N, W read MN-major."""
    dY, W = rnd(M, N, seed=3), rnd(N, K, seed=4, scale=N ** -0.5)
    ldn = (N + 7) // 8 * 8
    A = ops.PlanesTensor.from_float(dY, planes, ld=ldn)
    B = ops.PlanesTensor.from_float(W, planes)
    out = torch.full((M, K), float("nan"), device="cuda")
    ops.gemm(A, B, M, K, b_mn=True, out_f32=out)
    ref = A.float().double() @ B.float().double()
    assert rel(out, ref) < (1e-5 if planes == 1 else 5e-6)


@pytest.mark.parametrize("planes", [1, 3])
@pytest.mark.parametrize("M,N,K", [(2112, 2048, 768), (2112, 768, 2048), (264, 7, 768), (100, 136, 304)])
def test_gemm_wgrad_both_mn_major(ops, M, N, K, planes):
    """dW[N,K] = dY[M,N]^T . X[M,K]: contraction over M, both operands read MN-major."""
    dY, X = rnd(M, N, seed=5), rnd(M, K, seed=6)
    ldn = (N + 7) // 8 * 8
    A = ops.PlanesTensor.from_float(dY, planes, ld=ldn)
    B = ops.PlanesTensor.from_float(X, planes)
    out = torch.full((N, K), float("nan"), device="cuda")
    ops.gemm(A, B, N, K, a_mn=True, b_mn=True, out_f32=out)
    ref = A.float().double().T @ B.float().double()
    assert rel(out, ref) < (1e-5 if planes == 1 else 5e-6)


@pytest.mark.parametrize("M,N,K,planes,variant", [
    (70000, 768, 256, 1, "planes"),     # bias + ReLU -> one bf16 plane (FFN1 / fusion linear at inference); M not a multiple of 256
    (65600, 2304, 192, 1, "planes"),    # nine column tiles, K not a multiple of 64 rounded by the tail k-block
    (66000, 768, 320, 1, "resid"),      # bias + fp32 residual -> fp32 (out-projection / FFN2 at inference)
    (65792, 512, 128, 2, "both"),       # two operand planes (three plane products), fp32 and plane output together
    (65536, 300, 128, 1, "planes"),     # ragged N: the second column tile is clipped
    (65700, 512, 192, 3, "planes3"),    # fp32-parity inference: three operand planes (six products), three output planes
])
def test_gemm_persistent_many_row_tiles(ops, M, N, K, planes, variant):
    """M >= 65 536 (>= 512 row tiles): the persistent pair kernel of gemm_persist.cu -- every pair walks several tiles, two TMEM
    accumulators, epilogue staging of its own.  Compared with the fp64 product of the same bf16 planes."""
    X, W = rnd(M, K, seed=31), rnd(N, K, seed=32, scale=K ** -0.5)
    A, B = ops.PlanesTensor.from_float(X, planes), ops.PlanesTensor.from_float(W, planes)
    bias = rnd(N, seed=33)
    ref = A.float().double() @ B.float().double().T + bias.double()
    if variant == "planes3":
        outp = ops.PlanesTensor.empty(M, N, 3, "cuda", zero=True)
        ops.gemm(A, B, M, N, bias=bias, relu=True, out_planes=outp)
        assert rel(outp.float(), torch.relu(ref)) < 5e-6
    elif variant == "planes":
        outp = ops.PlanesTensor.empty(M, N, 1, "cuda", zero=True)
        ops.gemm(A, B, M, N, bias=bias, relu=True, out_planes=outp)
        assert rel(outp.float(), torch.relu(ref)) < 4e-3
    elif variant == "resid":
        resid = rnd(M, N, seed=34)
        out = torch.full((M, N), float("nan"), device="cuda")
        ops.gemm(A, B, M, N, bias=bias, resid=resid, out_f32=out)
        assert rel(out, ref + resid.double()) < 1e-5
    else:
        out = torch.full((M, N), float("nan"), device="cuda")
        outp = ops.PlanesTensor.empty(M, N, 1, "cuda", zero=True)
        ops.gemm(A, B, M, N, bias=bias, out_f32=out, out_planes=outp)
        assert rel(out, ref) < 1e-5 and rel(outp.float(), ref) < 4e-3


@pytest.mark.parametrize("M,N,K,planes,variant", [
    (2112, 768, 768, 2, "drop_resid"),   # out-projection forward of the benched mode: bias + dropout + residual -> fp32
    (2112, 2304, 768, 2, "planes"),      # packed QKV: twelve column tiles, two operand and two output planes
    (2112, 768, 2048, 1, "relu_drop"),   # bias + ReLU + dropout -> plane (the FFN1 epilogue on an N = 768 shape)
    (300, 500, 200, 3, "planes"),        # ragged: last column tile clipped (500 = 2 x 192 + 116), K tail, three planes
    (2112, 1536, 768, 1, "two_segs"),    # two K segments (concat operand), fp32 out
    (2112, 768, 2304, 1, "dgrad"),       # dX = dY W with W read MN-major: 96 columns per CTA as three atoms of the 64-byte swizzle
    (300, 500, 200, 3, "dgrad"),         # ... ragged, three planes
    (768, 768, 2112, 1, "wgrad"),        # dW = dY^T X, both operands MN-major
    (520, 400, 300, 2, "wgrad"),         # ... ragged, two planes
])
def test_gemm_192_wide_pair_tiles(ops, M, N, K, planes, variant):
    """256 x 192 CTA-pair tiles (the plan's choice for forward GEMMs whose N is a multiple of 192; forced here through the debug hook):
    six 32-column epilogue chunks per tile, 96 B rows per CTA, ring sized by the epilogue staging."""
    import ctypes as C
    L = ops._lib.lib()
    L.m2f_debug_force_gemm_bn.argtypes = [C.c_int]
    L.m2f_debug_force_gemm_bn.restype = None
    X, W = rnd(M, K, seed=35), rnd(N, K, seed=36, scale=K ** -0.5)
    A, B = ops.PlanesTensor.from_float(X, planes), ops.PlanesTensor.from_float(W, planes)
    bias = rnd(N, seed=37)
    ref = A.float().double() @ B.float().double().T + bias.double()
    tol_pl = {1: 4e-3, 2: 3e-5, 3: 5e-6}[planes]
    L.m2f_debug_force_gemm_bn(192)
    try:
        if variant == "dgrad":  # out[M, N] = dY[M, K] . Wt[K, N]
            Wt = ops.PlanesTensor.from_float(rnd(K, N, seed=39, scale=K ** -0.5), planes, ld=(N + 7) // 8 * 8)
            resid = rnd(M, N, seed=38)
            out = torch.full((M, N), float("nan"), device="cuda")
            ops.gemm(A, Wt, M, N, b_mn=True, resid=resid, out_f32=out)
            assert rel(out, A.float().double() @ Wt.float().double() + resid.double()) < 1e-5
        elif variant == "wgrad":  # out[M, N] = dY[K, M]^T . X[K, N]
            dY = ops.PlanesTensor.from_float(rnd(K, M, seed=40), planes, ld=(M + 7) // 8 * 8)
            Xk = ops.PlanesTensor.from_float(rnd(K, N, seed=41), planes, ld=(N + 7) // 8 * 8)
            out = torch.full((M, N), float("nan"), device="cuda")
            ops.gemm(dY, Xk, M, N, a_mn=True, b_mn=True, out_f32=out)
            assert rel(out, dY.float().double().T @ Xk.float().double()) < 1e-5
        elif variant == "planes":
            outp = ops.PlanesTensor.empty(M, N, planes, "cuda", zero=True)
            ops.gemm(A, B, M, N, bias=bias, out_planes=outp)
            assert rel(outp.float(), ref) < tol_pl
        elif variant == "drop_resid":
            resid = rnd(M, N, seed=38)
            out = torch.full((M, N), float("nan"), device="cuda")
            ops.gemm(A, B, M, N, bias=bias, resid=resid, out_f32=out, drop_p=0.4, seed=3, step=4, site=0x55)
            keep = ops.dropout_mask(0.4, 3, 4, 0x55, M * N, "cuda").reshape(M, N).double()
            assert rel(out, ref * keep / 0.6 + resid.double()) < 1e-5
        elif variant == "relu_drop":
            outp = ops.PlanesTensor.empty(M, N, 1, "cuda", zero=True)
            ops.gemm(A, B, M, N, bias=bias, relu=True, out_planes=outp, drop_p=0.4, seed=3, step=4, site=0x56)
            keep = ops.dropout_mask(0.4, 3, 4, 0x56, M * N, "cuda").reshape(M, N).double()
            assert rel(outp.float(), torch.relu(ref) * keep / 0.6) < tol_pl
        else:
            out = torch.full((M, N), float("nan"), device="cuda")
            h = K // 128 * 64
            ops.gemm(A, B, M, N, bias=bias, segs=[(0, 0, h), (h, h, K - h)], out_f32=out)
            assert rel(out, ref) < 1e-5
    finally:
        L.m2f_debug_force_gemm_bn(-1)


def test_gemm_a_mn_b_k_major(ops):
    M, N, K = 300, 200, 192
    At, W = rnd(K, M, seed=7), rnd(N, K, seed=8)
    A, B = ops.PlanesTensor.from_float(At, 1), ops.PlanesTensor.from_float(W, 1)
    out = torch.empty(M, N, device="cuda")
    ops.gemm(A, B, M, N, a_mn=True, out_f32=out)
    assert rel(out, A.float().double().T @ B.float().double().T) < 1e-5


def test_gemm_colsum_ragged(ops):
    M, N, K = 100, 136, 304
    X, W = rnd(M, K, seed=21), rnd(N, K, seed=22)
    A, B = ops.PlanesTensor.from_float(X, 1), ops.PlanesTensor.from_float(W, 1)
    bias = rnd(N, seed=23)
    outp = ops.PlanesTensor.empty(M, N, 3, "cuda", zero=True)
    cs = torch.zeros(N, device="cuda")
    ops.gemm(A, B, M, N, bias=bias, out_planes=outp, colsum=cs)
    ref = A.float().double() @ B.float().double().T + bias.double()
    assert rel(outp.float(), ref) < 1e-5 and rel(cs, ref.sum(0)) < 1e-5


def test_gemm_two_k_segments(ops):
    """Fusion linear on cat(x, text) without materialising the concat (src/model.py:16-18)."""
    M, N, Df = 500, 256, 128
    U, W = rnd(M, 2 * Df, seed=9), rnd(N, 2 * Df, seed=10)
    A, B = ops.PlanesTensor.from_float(U, 1), ops.PlanesTensor.from_float(W, 1)
    out = torch.empty(M, N, device="cuda")
    ops.gemm(A, B, M, N, segs=[(0, 0, Df), (Df, Df, Df)], out_f32=out)
    assert rel(out, A.float().double() @ B.float().double().T) < 1e-5
    # only the second half
    ops.gemm(A.cview(Df, Df), B.cview(Df, Df), M, N, out_f32=out)
    assert rel(out, A.float()[:, Df:].double() @ B.float()[:, Df:].double().T) < 1e-5


@pytest.mark.parametrize("planes_out", [1, 3])
def test_gemm_epilogue_chain(ops, planes_out):
    """bias -> relu -> gate -> dropout -> residual -> accumulate, fp32 + planes outputs."""
    M, N, K, p = 264, 192, 128, 0.4
    X, W = rnd(M, K, seed=11), rnd(N, K, seed=12, scale=K ** -0.5)
    bias, resid, prev = rnd(N, seed=13), rnd(M, N, seed=14), rnd(M, N, seed=15)
    gate_src = rnd(M, N, seed=16)
    A, B = ops.PlanesTensor.from_float(X, 1), ops.PlanesTensor.from_float(W, 1)
    gate = ops.PlanesTensor.from_float(gate_src, 1)
    out = prev.clone()
    outp = ops.PlanesTensor.empty(M, N, planes_out, "cuda", zero=True)
    cs = torch.ones(N, device="cuda")
    ops.gemm(A, B, M, N, bias=bias, relu=True, gate=gate, gate_scale=1.25, drop_p=p, seed=77, step=3, site=0x1234,
             resid=resid, out_f32=out, accumulate=True, out_planes=outp, colsum=cs)
    keep = ops.dropout_mask(p, 77, 3, 0x1234, M * N, "cuda").reshape(M, N).double()
    v = A.float().double() @ B.float().double().T + bias.double()
    v = torch.relu(v)
    v = torch.where(gate.float() > 0, v * 1.25, torch.zeros_like(v))
    v = v * keep / (1 - p) + resid.double()
    assert rel(out, v + prev.double()) < 1e-5
    assert rel(outp.float(), v + prev.double()) < (4e-3 if planes_out == 1 else 1e-6)
    assert rel(cs, (v + prev.double()).sum(0) + 1.0) < 1e-5  # bias-gradient column sums of the stored values
    frac = keep.mean().item()
    assert abs(frac - (1 - p)) < 0.01


def test_dropout_mask_statistics(ops):
    n = 1 << 20
    for p in (0.1, 0.4, 0.5):
        m = ops.dropout_mask(p, 5, 0, 9, n, "cuda").float()
        assert abs(m.mean().item() - (1 - p)) < 3e-3
    a = ops.dropout_mask(0.4, 5, 0, 9, n, "cuda")
    assert torch.equal(a, ops.dropout_mask(0.4, 5, 0, 9, n, "cuda"))          # deterministic
    for other in (ops.dropout_mask(0.4, 6, 0, 9, n, "cuda"), ops.dropout_mask(0.4, 5, 1, 9, n, "cuda"),
                  ops.dropout_mask(0.4, 5, 0, 10, n, "cuda")):                # seed / step / site decorrelate
        agree = (a == other).float().mean().item()
        assert abs(agree - (0.6 * 0.6 + 0.4 * 0.4)) < 5e-3
    assert ops.dropout_mask(0.0, 1, 2, 3, 1000, "cuda").all()


def test_relu_and_colsum(ops):
    X = rnd(333, 96, seed=20)
    for P in (1, 3):
        src = ops.PlanesTensor.from_float(X, P)
        dst = ops.PlanesTensor.empty(333, 96, P, "cuda", zero=True)
        ops.relu_planes(src, dst)
        assert rel(dst.float(), torch.relu(src.float())) < (4e-3 if P == 1 else 1e-6)
        acc = torch.ones(96, device="cuda")
        ops.colsum_planes(src, acc)
        assert rel(acc, src.float().double().sum(0) + 1.0) < 1e-5


def _ln_ref(x, g, b):
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + 1e-5) * g + b


@pytest.mark.parametrize("staged", ["0", "2"])
@pytest.mark.parametrize("M,D", [(2112, 768), (264, 1024), (77, 300), (5, 64), (4099, 768)])
def test_layernorm_forward_backward(ops, M, D, staged, monkeypatch):
    """Both kernel families (M2F_LN_V2 is read per call): 0 = register-resident rows, 2 = rows staged through shared memory with
    bulk asynchronous copies (what the size rule picks once an [M, D] fp32 operand reaches 32 MB)."""
    monkeypatch.setenv("M2F_LN_V2", staged)
    x, skip = rnd(M, D, seed=30) * 2 + 0.3, rnd(M, D, seed=31)
    gamma, beta = 1 + 0.1 * rnd(D, seed=32), 0.1 * rnd(D, seed=33)
    p = 0.4
    y = torch.empty(M, D, device="cuda")
    ypl = ops.PlanesTensor.empty(M, D, 3, "cuda", zero=True)
    stats = torch.empty(M, 2, device="cuda")
    ops.ln_fwd(x, gamma, beta, skip=skip, y=y, y_planes=ypl, drop_p=p, seed=1, step=2, site=3, stats=stats)
    ref = _ln_ref(x.double(), gamma.double(), beta.double()) + skip.double()
    assert rel(y, ref) < 2e-6
    keep = ops.dropout_mask(p, 1, 2, 3, M * D, "cuda").reshape(M, D).double()
    assert rel(ypl.float(), ref * keep / (1 - p)) < 2e-6
    assert rel(stats[:, 0], x.double().mean(-1)) < 1e-5
    # backward
    g = rnd(M, D, seed=34)
    xd = x.double().requires_grad_(True)
    gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    _ln_ref(xd, gd, bd).backward(g.double())
    dx = torch.empty(M, D, device="cuda")
    dxp = ops.PlanesTensor.empty(M, D, 3, "cuda", zero=True)
    dgamma, dbeta, dbias = (torch.zeros(D, device="cuda") for _ in range(3))
    ops.ln_bwd(g, x, gamma, stats, dx=dx, dx_planes=dxp, drop_p=p, seed=4, step=5, site=6, dgamma=dgamma, dbeta=dbeta, dbias=dbias)
    assert rel(dx, xd.grad) < 5e-6
    keep2 = ops.dropout_mask(p, 4, 5, 6, M * D, "cuda").reshape(M, D).double()
    assert rel(dxp.float(), xd.grad * keep2 / (1 - p)) < 5e-6
    assert rel(dgamma, gd.grad) < 1e-5
    assert rel(dbeta, bd.grad) < 1e-5
    assert rel(dbias, (xd.grad * keep2 / (1 - p)).sum(0)) < 1e-4


def _attn_ref(q, k, v, pad, H, keep=None, p=0.0):
    B, S, E = q.shape
    hd = E // H
    sp = lambda t: t.reshape(B, S, H, hd).permute(0, 2, 1, 3)
    s = (sp(q) / math.sqrt(hd)) @ sp(k).transpose(-1, -2)
    s = s.masked_fill(pad[:, None, None, :].bool(), float("-inf"))
    P = torch.softmax(s, dim=-1)
    if keep is not None:
        P = P * keep / (1 - p)
    return (P @ sp(v)).permute(0, 2, 1, 3).reshape(B, S, E)


@pytest.mark.parametrize("planes", [1, 3])
@pytest.mark.parametrize("B,S,H,hd,p", [(6, 33, 8, 96, 0.0), (6, 33, 8, 96, 0.4), (3, 7, 10, 30, 0.4), (2, 64, 4, 64, 0.0), (2, 1, 2, 32, 0.0),
                                        (3, 48, 2, 128, 0.4), (5, 17, 3, 16, 0.0), (2, 100, 2, 64, 0.4), (64, 33, 8, 96, 0.4)])
def test_attention_forward_backward(ops, B, S, H, hd, p, planes):
    E = H * hd
    g = torch.Generator().manual_seed(40)
    L = torch.randint(1, S + 1, (B,), generator=g)
    L[0] = S
    if B > 1:
        L[B - 1] = 1
    pad = (torch.arange(S)[None, :] >= L[:, None]).to(torch.uint8).cuda().contiguous()
    qkv = rnd(B * S, 3 * E, seed=41)
    packed = ops.PlanesTensor.from_float(qkv, planes)
    q, k, v = packed.cview(0, E), packed.cview(E, E), packed.cview(2 * E, E)
    o = ops.PlanesTensor.empty(B * S, E, planes, "cuda", zero=True)
    ops.attn_fwd(q, k, v, pad, B, S, H, o, drop_p=p, seed=9, step=1, site=0x3000)
    S8 = (S + 7) // 8 * 8  # row stride of the attention-probability dropout elements (include/m2f_b200.h)
    keep = ops.dropout_mask(p, 9, 1, 0x3000, B * H * S * S8, "cuda").reshape(B, H, S, S8)[..., :S].double() if p > 0 else None
    qd, kd, vd = (t.float().double().reshape(B, S, E).requires_grad_(True) for t in (q, k, v))
    ref = _attn_ref(qd, kd, vd, pad, H, keep, p)
    tol = 5e-3 if planes == 1 else 3e-6
    assert rel(o.float(), ref.reshape(B * S, E)) < tol
    d_o = rnd(B * S, E, seed=42)
    dop = ops.PlanesTensor.from_float(d_o, planes)
    ref.backward(dop.float().double().reshape(B, S, E))
    dqkv = ops.PlanesTensor.empty(B * S, 3 * E, planes, "cuda", zero=True)
    db = torch.ones(3 * E, device="cuda")
    ops.attn_bwd(q, k, v, dop, pad, B, S, H, dqkv.cview(0, E), dqkv.cview(E, E), dqkv.cview(2 * E, E), drop_p=p, seed=9, step=1, site=0x3000,
                 dbq=db[:E], dbk=db[E:2 * E], dbv=db[2 * E:])
    got = dqkv.float()
    ref_db = torch.cat([t.grad.reshape(B * S, E).sum(0) for t in (qd, kd, vd)]) + 1.0  # in_proj_bias gradient slices
    assert rel(db, ref_db) < (2e-2 if planes == 1 else 1e-5)
    assert rel(got[:, :E], qd.grad.reshape(B * S, E)) < tol
    assert rel(got[:, E:2 * E], kd.grad.reshape(B * S, E)) < tol
    assert rel(got[:, 2 * E:], vd.grad.reshape(B * S, E)) < tol


@pytest.mark.parametrize("planes", [1, 3])
@pytest.mark.parametrize("B,S,H,hd", [(3, 256, 8, 96), (4, 65, 2, 64), (2, 130, 3, 128), (5, 200, 4, 32), (2, 129, 2, 96),
                                      # more (dialogue, head) items than SMs: the persistent tcgen05 kernel (attention_tc.cu, one plane)
                                      # walks several items per CTA -- prefetch, barrier phases, TMEM hand-over between items; one
                                      # and two query tiles, both query-slot layouts (hd <= 96 / hd = 128)
                                      (56, 256, 8, 96), (90, 130, 4, 128), (170, 100, 4, 64), (75, 131, 8, 96),
                                      (2, 300, 2, 64), (2, 513, 1, 128)])  # beyond the tcgen05 kernel's single key tile: mma.sync tiles
def test_attention_forward_long_dialogues(ops, B, S, H, hd, planes):
    """64 < S <= 256, no dropout (inference at BASELINE.json config 5): the tcgen05 kernel of attention_tc.cu for one plane and
    head sizes 64 / 96 / 128, the online-softmax mma.sync kernel of attention_mma_long.cu otherwise.  Lengths are ragged; one
    dialogue masks a whole 64-key tile in the middle, which the reference's key_padding_mask allows (any boolean pattern,
    torch/nn/functional.py:6460-6490)."""
    if planes == 3 and B > 8:
        pytest.skip("the many-item cases target the one-plane kernel")
    E = H * hd
    g = torch.Generator().manual_seed(50)
    L = torch.randint(1, S + 1, (B,), generator=g)
    L[0] = S
    L[B - 1] = 1
    padb = torch.arange(S)[None, :] >= L[:, None]
    padb[0, 64:128] = True   # a fully masked key tile between live ones
    padb[1, :64] = True      # leading tile masked ...
    padb[1, S - 1] = False   # ... with a live key in the last tile
    pad = padb.to(torch.uint8).cuda().contiguous()
    qkv = rnd(B * S, 3 * E, seed=51) * 2.0
    packed = ops.PlanesTensor.from_float(qkv, planes)
    q, k, v = packed.cview(0, E), packed.cview(E, E), packed.cview(2 * E, E)
    o = ops.PlanesTensor.empty(B * S, E, planes, "cuda", zero=True)
    ops.attn_fwd(q, k, v, pad, B, S, H, o)
    qd, kd, vd = (t.float().double().reshape(B, S, E) for t in (q, k, v))
    ref = _attn_ref(qd, kd, vd, pad, H)
    assert rel(o.float(), ref.reshape(B * S, E)) < (5e-3 if planes == 1 else 3e-6)


@pytest.mark.parametrize("planes", [1, 2, 3])
@pytest.mark.parametrize("B,S,H,hd,p", [(3, 130, 2, 64, 0.4), (2, 256, 2, 128, 0.4), (3, 200, 4, 96, 0.0), (4, 65, 3, 32, 0.4), (2, 129, 8, 96, 0.4),
                                        (2, 300, 2, 64, 0.4), (2, 520, 1, 96, 0.4)])  # beyond 256: no structural limit in the tiled kernels
def test_attention_long_dialogues_training(ops, B, S, H, hd, p, planes):
    """64 < S <= 256 in training: forward with probability dropout (attn_fwd_long_kernel<.., DROP>) and the tiled backward
    (attn_bwd_long_kernel: row statistics, dQ, then dK / dV, tile by tile) of attention_mma_long.cu against torch autograd in fp64
    on the same operands with the library's own keep mask.  Ragged lengths, a fully masked key tile in the middle of one dialogue,
    a leading masked tile in another, a single-utterance dialogue."""
    E = H * hd
    g = torch.Generator().manual_seed(60)
    L = torch.randint(1, S + 1, (B,), generator=g)
    L[0] = S
    L[B - 1] = 1
    padb = torch.arange(S)[None, :] >= L[:, None]
    if S >= 128:
        padb[0, 64:128] = True
    padb[1, :64] = True
    padb[1, S - 1] = False
    pad = padb.to(torch.uint8).cuda().contiguous()
    qkv = rnd(B * S, 3 * E, seed=61) * 1.5
    packed = ops.PlanesTensor.from_float(qkv, planes)
    q, k, v = packed.cview(0, E), packed.cview(E, E), packed.cview(2 * E, E)
    o = ops.PlanesTensor.empty(B * S, E, planes, "cuda", zero=True)
    ops.attn_fwd(q, k, v, pad, B, S, H, o, drop_p=p, seed=19, step=3, site=0x3100)
    S8 = (S + 7) // 8 * 8
    keep = ops.dropout_mask(p, 19, 3, 0x3100, B * H * S * S8, "cuda").reshape(B, H, S, S8)[..., :S].double() if p > 0 else None
    qd, kd, vd = (t.float().double().reshape(B, S, E).requires_grad_(True) for t in (q, k, v))
    ref = _attn_ref(qd, kd, vd, pad, H, keep, p)
    tol = {1: 5e-3, 2: 5e-5, 3: 3e-6 if S <= 256 else 6e-6}[planes]  # fp32 sums over S terms: 3.7e-6 measured at S = 520
    assert rel(o.float(), ref.reshape(B * S, E)) < tol
    d_o = rnd(B * S, E, seed=62)
    dop = ops.PlanesTensor.from_float(d_o, planes)
    ref.backward(dop.float().double().reshape(B, S, E))
    dqkv = ops.PlanesTensor.empty(B * S, 3 * E, planes, "cuda", zero=True)
    db = torch.ones(3 * E, device="cuda")
    ops.attn_bwd(q, k, v, dop, pad, B, S, H, dqkv.cview(0, E), dqkv.cview(E, E), dqkv.cview(2 * E, E), drop_p=p, seed=19, step=3, site=0x3100,
                 dbq=db[:E], dbk=db[E:2 * E], dbv=db[2 * E:])
    got = dqkv.float()
    ref_db = torch.cat([t.grad.reshape(B * S, E).sum(0) for t in (qd, kd, vd)]) + 1.0
    assert rel(db, ref_db) < (2e-2 if planes == 1 else 1e-4)
    assert rel(got[:, :E], qd.grad.reshape(B * S, E)) < tol
    assert rel(got[:, E:2 * E], kd.grad.reshape(B * S, E)) < tol
    assert rel(got[:, 2 * E:], vd.grad.reshape(B * S, E)) < tol


def test_attention_limits_are_errors(ops):
    q = ops.PlanesTensor.empty(300, 60, 1, "cuda", zero=True)  # head size 30: the generic kernel, whose limit is S = 256
    pad = torch.zeros(1, 300, dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match="exceeds the single-CTA limit"):
        ops.attn_fwd(q, q, q, pad, 1, 300, 2, q)
    with pytest.raises(RuntimeError, match="exceeds the single-CTA limit"):
        ops.attn_bwd(q, q, q, q, pad, 1, 300, 2, q, q, q)
    big = ops.PlanesTensor.empty(4100, 64, 1, "cuda", zero=True)  # head size 32: tiled kernels up to S = 4096
    pad = torch.zeros(1, 4100, dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match="exceeds the single-CTA limit"):
        ops.attn_fwd(big, big, big, pad, 1, 4100, 2, big)


@pytest.mark.parametrize("weighted", [False, True])
def test_cross_entropy(ops, weighted):
    from oracle import m2fnet_oracle as O
    M, Cc = 2112, 7
    logits = rnd(M, Cc, seed=50) * 2
    g = torch.Generator().manual_seed(51)
    tgt = torch.randint(0, Cc, (M,), generator=g)
    tgt[torch.rand(M, generator=g) < 0.3] = -1
    tgt = tgt.cuda()
    w = (torch.rand(Cc, generator=g) + 0.5).cuda() if weighted else None
    dl = torch.empty(M, Cc, device="cuda")
    dlp = ops.PlanesTensor.empty(M, 8, 3, "cuda")
    out = ops.ce_fwd_bwd(logits, tgt, w, eps=0.1, grad_scale=1.0, dlogits=dl, dl_planes=dlp)
    ld = logits.double().requires_grad_(True)
    ref = O.ce_loss(ld[None], tgt[None], w.double() if w is not None else None)
    ref.backward()
    assert abs(out[0].item() - ref.item()) < 2e-6 * abs(ref.item())
    assert rel(dl, ld.grad) < 5e-6
    assert rel(dlp.float()[:, :Cc], ld.grad) < 5e-6
    assert (dlp.float()[:, Cc:] == 0).all()
    tref = torch.nn.CrossEntropyLoss(weight=w, ignore_index=-1, label_smoothing=0.1)(logits, tgt)
    assert abs(out[0].item() - tref.item()) < 1e-5
