// K1: TMA-fed tcgen05 GEMM for sm_100a with a fused epilogue.
//
//   D[M,N] = sum over K-segments and bf16-plane products of  A[M,K] . B[N,K]^T      (fp32 accumulate in TMEM)
//
// Replaces every nn.Linear / MHA in-projection of the reference model (src/model.py:9,68,80,90-98) and
// their autograd backward: forward Y = X W^T is "TN" (both operands K-major); dgrad dX = dY W reads W
// MN-major; wgrad dW = dY^T X reads both operands MN-major -- no transposes are materialised, the UMMA
// shared-memory descriptors / instruction-descriptor major bits select the layout.
//
// Warp roles (192 threads): warp 0 = TMA producer (one elected lane), warp 1 = MMA issuer (one elected lane),
// warps 2..5 = epilogue (TMEM -> registers -> global; warp w owns TMEM lanes 32*(w%4)..+31).
// Pipeline: STAGES-deep smem ring with full/empty mbarriers (TMA <-> MMA), one tmem_full mbarrier
// (MMA -> epilogue).  One 128 x BLOCK_N output tile per CTA.
#include <mutex>

#include "common.cuh"
#include "gemm.cuh"
#include "ptx.cuh"

namespace m2f {

template <int BLOCK_N>
struct GemmCfg {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = BLOCK_N * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (BLOCK_N == 256) ? 4 : (BLOCK_N == 128 ? 6 : 8);
    static constexpr int BAR_BYTES = (2 * STAGES + 1) * 8 + 16;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;  // + alignment slack
    static constexpr int TMEM_COLS = BLOCK_N < 32 ? 32 : BLOCK_N;
};

template <int BLOCK_N, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
    using Cfg = GemmCfg<BLOCK_N>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_blk = blockIdx.x, n_blk = blockIdx.y;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tma_a);
        ptx::prefetch_tensormap(&tma_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        ptx::mbar_init(tmem_full_bar, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    ptx::tcgen05_fence_before();
    __syncthreads();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (ptx::elect_one()) {
            int it = 0;
            for (int pr = 0; pr < p.nprod; ++pr) {
                const int pa = p.prod_a[pr], pb = p.prod_b[pr];
                for (int sg = 0; sg < p.nsegs; ++sg) {
                    for (int kb = 0; kb < p.seg_nkb[sg]; ++kb, ++it) {
                        const int stage = it % STAGES;
                        const uint32_t phase = (it / STAGES) & 1;
                        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                        ptx::mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                        const int ka = p.seg_a_k0[sg] + kb * BLOCK_K;
                        const int kbb = p.seg_b_k0[sg] + kb * BLOCK_K;
                        uint8_t* sa = smem_a + stage * Cfg::A_BYTES;
                        uint8_t* sb = smem_b + stage * Cfg::B_BYTES;
                        if (!A_MN) {
                            ptx::tma_load_3d(sa, &tma_a, &full_bar[stage], ka, m_blk * BLOCK_M, pa);
                        } else {
#pragma unroll
                            for (int j = 0; j < BLOCK_M / 64; ++j)
                                ptx::tma_load_3d(sa + j * (64 * BLOCK_K * 2), &tma_a, &full_bar[stage],
                                                 m_blk * BLOCK_M + j * 64, ka, pa);
                        }
                        if (!B_MN) {
                            ptx::tma_load_3d(sb, &tma_b, &full_bar[stage], kbb, n_blk * BLOCK_N, pb);
                        } else {
#pragma unroll
                            for (int j = 0; j < BLOCK_N / 64; ++j)
                                ptx::tma_load_3d(sb + j * (64 * BLOCK_K * 2), &tma_b, &full_bar[stage],
                                                 n_blk * BLOCK_N + j * 64, kbb, pb);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::instr_desc_bf16(BLOCK_M, BLOCK_N, A_MN, B_MN);
            // K-major SW128: 8-row groups 1024 B apart (SBO), one atom along K (LBO unused);
            //   advancing K by 16 elements = +32 B inside the swizzle atom.
            // MN-major SW128: 64-element MN atoms BLOCK_K*128 B apart (LBO), 8-k-row groups 1024 B apart (SBO);
            //   advancing K by 16 rows = +2048 B.
            constexpr uint32_t A_LBO = A_MN ? BLOCK_K * 128 : 0, B_LBO = B_MN ? BLOCK_K * 128 : 0;
            constexpr uint32_t A_KSTEP = A_MN ? UMMA_K * 128 : UMMA_K * 2, B_KSTEP = B_MN ? UMMA_K * 128 : UMMA_K * 2;
            for (int it = 0; it < p.total_kb; ++it) {
                const int stage = it % STAGES;
                const uint32_t phase = (it / STAGES) & 1;
                ptx::mbar_wait(&full_bar[stage], phase);
                ptx::tcgen05_fence_after();
                const uint32_t sa = ptx::smem_u32(smem_a + stage * Cfg::A_BYTES);
                const uint32_t sb = ptx::smem_u32(smem_b + stage * Cfg::B_BYTES);
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                    const uint64_t ad = ptx::smem_desc_sw128(sa + k * A_KSTEP, A_LBO, 1024);
                    const uint64_t bd = ptx::smem_desc_sw128(sb + k * B_KSTEP, B_LBO, 1024);
                    ptx::umma_bf16(tmem_base, ad, bd, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                ptx::umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
            }
            ptx::umma_commit(tmem_full_bar);
        }
    } else {
        // ------------------------------------------------------------------ epilogue
        ptx::mbar_wait(tmem_full_bar, 0);
        ptx::tcgen05_fence_after();
        const int lg = warp & 3;  // TMEM lane group this warp may access
        const int64_t m = (int64_t)m_blk * BLOCK_M + lg * 32 + lane;
        const bool row_ok = m < p.M;
        const bool vec4 = (p.out_f32 == nullptr || (p.out_f32_ld & 3) == 0) && (p.resid == nullptr || (p.resid_ld & 3) == 0) &&
                          ((p.drop_ld & 3) == 0);
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 32; ++c) {
            const int n0 = n_blk * BLOCK_N + c * 32;
            if (n0 >= p.N) break;  // warp-uniform
            uint32_t r[32];
            ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(c * 32), r);
            ptx::tmem_ld_wait();
            const bool full = (n0 + 32 <= p.N) && vec4;
            if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                float v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[j + i]);
                const int n = n0 + j;
                if (full) {
                    if (p.bias) {
                        const float4 b = *reinterpret_cast<const float4*>(p.bias + n);
                        v[0] += b.x; v[1] += b.y; v[2] += b.z; v[3] += b.w;
                    }
                    if (p.relu) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) v[i] = fmaxf(v[i], 0.f);
                    }
                    if (p.gate) {
                        const uint2 raw = *reinterpret_cast<const uint2*>(p.gate + m * p.gate_ld + n);
                        const __nv_bfloat162 g0 = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
                        const __nv_bfloat162 g1 = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
                        v[0] = __low2float(g0) > 0.f ? v[0] * p.gate_scale : 0.f;
                        v[1] = __high2float(g0) > 0.f ? v[1] * p.gate_scale : 0.f;
                        v[2] = __low2float(g1) > 0.f ? v[2] * p.gate_scale : 0.f;
                        v[3] = __high2float(g1) > 0.f ? v[3] * p.gate_scale : 0.f;
                    }
                    if (p.drop.enabled) {
                        float dm[4];
                        drop_mult4(p.drop, (uint64_t)(m * p.drop_ld + n) >> 2, dm);
#pragma unroll
                        for (int i = 0; i < 4; ++i) v[i] *= dm[i];
                    }
                    if (p.resid) {
                        const float4 q = *reinterpret_cast<const float4*>(p.resid + m * p.resid_ld + n);
                        v[0] += q.x; v[1] += q.y; v[2] += q.z; v[3] += q.w;
                    }
                    if (p.out_f32) {
                        float4* o = reinterpret_cast<float4*>(p.out_f32 + m * p.out_f32_ld + n);
                        if (p.accumulate) {
                            const float4 q = *o;
                            v[0] += q.x; v[1] += q.y; v[2] += q.z; v[3] += q.w;
                        }
                        *o = make_float4(v[0], v[1], v[2], v[3]);
                    }
                    if (p.out_pl) store_planes4(p.out_pl, p.out_pl_stride, p.out_pl_n, m * p.out_pl_ld + n, v);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int nn = n + i;
                        if (nn >= p.N) break;
                        float x = v[i];
                        if (p.bias) x += p.bias[nn];
                        if (p.relu) x = fmaxf(x, 0.f);
                        if (p.gate) x = __bfloat162float(p.gate[m * p.gate_ld + nn]) > 0.f ? x * p.gate_scale : 0.f;
                        if (p.drop.enabled) x *= drop_mult1(p.drop, (uint64_t)(m * p.drop_ld + nn));
                        if (p.resid) x += p.resid[m * p.resid_ld + nn];
                        if (p.out_f32) {
                            float* o = p.out_f32 + m * p.out_f32_ld + nn;
                            if (p.accumulate) x += *o;
                            *o = x;
                        }
                        if (p.out_pl) store_planes(p.out_pl, p.out_pl_stride, p.out_pl_n, m * p.out_pl_ld + nn, x);
                    }
                }
            }
            }  // row_ok
        }
        ptx::tcgen05_fence_before();
    }
    __syncthreads();
    if (warp == 2) {
        ptx::tcgen05_fence_after();
        ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------- host side

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}

// 3-D bf16 tensor map over planes: dims {cols, rows, planes}; box {64, box_rows, 1}; 128-byte swizzle;
// out-of-bounds elements read as zero (ragged M / N / K need no special casing in the kernel).
int make_planes_tmap(CUtensorMap* map, const m2f_planes_t* t, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    M2F_REQUIRE(enc != nullptr, M2F_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    M2F_REQUIRE(t->ptr != nullptr && t->rows > 0 && t->cols > 0 && t->planes >= 1 && t->planes <= 3, M2F_ERR_ARG,
                "tensor map: bad planes view (rows %lld cols %lld planes %d)", (long long)t->rows, (long long)t->cols,
                t->planes);
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(t->ptr) & 15) == 0, M2F_ERR_ALIGN, "tensor map: base not 16-byte aligned");
    M2F_REQUIRE((t->ld % 8) == 0 && t->ld >= t->cols, M2F_ERR_ALIGN, "tensor map: ld %lld must be a multiple of 8 and >= cols %lld",
                (long long)t->ld, (long long)t->cols);
    M2F_REQUIRE(t->planes == 1 || (t->plane_stride % 8) == 0, M2F_ERR_ALIGN, "tensor map: plane stride must be a multiple of 8");
    cuuint64_t dims[3] = {(cuuint64_t)t->cols, (cuuint64_t)t->rows, (cuuint64_t)t->planes};
    const int64_t ps = t->planes > 1 ? t->plane_stride : t->rows * t->ld;
    cuuint64_t strides[2] = {(cuuint64_t)t->ld * 2, (cuuint64_t)ps * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, t->ptr, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    M2F_REQUIRE(r == CUDA_SUCCESS, M2F_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return M2F_OK;
}

template <int BLOCK_N, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t stream) {
    using Cfg = GemmCfg<BLOCK_N>;
    auto kern = gemm_kernel<BLOCK_N, A_MN, B_MN>;
    static bool attr_set = false;  // per instantiation
    if (!attr_set) {
        M2F_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        attr_set = true;
    }
    dim3 grid(ceil_div(p.M, BLOCK_M), ceil_div(p.N, BLOCK_N));
    kern<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

int gemm_block_n(int64_t M, int64_t N) {
    // one 128 x BLOCK_N tile per CTA: prefer the widest tile that still yields >= 148 CTAs
    const int64_t mt = (M + BLOCK_M - 1) / BLOCK_M;
    if (N > 128 && mt * ((N + 255) / 256) >= 148) return 256;
    if (N > 64 && mt * ((N + 127) / 128) >= 120) return 128;
    return N > 64 && mt * ((N + 63) / 64) > 296 ? 128 : 64;
}

int gemm_fill_params(GemmParams& p, const m2f_planes_t* A, int a_mn, const m2f_planes_t* B, int b_mn,
                     const m2f_kseg_t* segs, int nsegs, int64_t M, int64_t N, const m2f_epilogue_t* e) {
    M2F_REQUIRE(A && B && segs && e, M2F_ERR_ARG, "gemm: null argument");
    M2F_REQUIRE(M > 0 && N > 0 && M < (1ll << 31) && N < (1ll << 31), M2F_ERR_ARG, "gemm: bad M/N");
    M2F_REQUIRE(nsegs >= 1 && nsegs <= MAX_SEGS, M2F_ERR_UNSUPPORTED, "gemm: 1..%d K segments supported, got %d", MAX_SEGS, nsegs);
    M2F_REQUIRE(A->planes == B->planes, M2F_ERR_ARG, "gemm: A has %d planes, B has %d", A->planes, B->planes);
    const int64_t a_mn_ext = a_mn ? A->cols : A->rows, a_k_ext = a_mn ? A->rows : A->cols;
    const int64_t b_mn_ext = b_mn ? B->cols : B->rows, b_k_ext = b_mn ? B->rows : B->cols;
    M2F_REQUIRE(a_mn_ext >= M && b_mn_ext >= N, M2F_ERR_ARG, "gemm: operands smaller than M x N (%lld x %lld vs %lld x %lld)",
                (long long)a_mn_ext, (long long)b_mn_ext, (long long)M, (long long)N);
    p.M = (int)M;
    p.N = (int)N;
    p.nsegs = nsegs;
    int kb_per_prod = 0;
    for (int i = 0; i < nsegs; ++i) {
        M2F_REQUIRE(segs[i].klen > 0 && segs[i].a_k0 >= 0 && segs[i].b_k0 >= 0, M2F_ERR_ARG, "gemm: bad K segment %d", i);
        M2F_REQUIRE(segs[i].a_k0 + segs[i].klen <= a_k_ext && segs[i].b_k0 + segs[i].klen <= b_k_ext, M2F_ERR_ARG,
                    "gemm: K segment %d exceeds an operand", i);
        // a ragged tail would pull in columns of the next segment; it is only legal at the end of the operand
        const bool ragged = (segs[i].klen % BLOCK_K) != 0;
        M2F_REQUIRE(!ragged || (segs[i].a_k0 + segs[i].klen == a_k_ext && segs[i].b_k0 + segs[i].klen == b_k_ext),
                    M2F_ERR_UNSUPPORTED, "gemm: K segment %d has a ragged tail that is not at the end of both operands", i);
        p.seg_a_k0[i] = segs[i].a_k0;
        p.seg_b_k0[i] = segs[i].b_k0;
        p.seg_nkb[i] = ceil_div(segs[i].klen, BLOCK_K);
        kb_per_prod += p.seg_nkb[i];
    }
    // plane products, least significant first (fp32 accumulation order)
    static const int P1[][2] = {{0, 0}};
    static const int P2[][2] = {{1, 0}, {0, 1}, {0, 0}};
    static const int P3[][2] = {{2, 0}, {1, 1}, {0, 2}, {1, 0}, {0, 1}, {0, 0}};
    const int (*tab)[2] = A->planes == 1 ? P1 : (A->planes == 2 ? P2 : P3);
    p.nprod = A->planes == 1 ? 1 : (A->planes == 2 ? 3 : 6);
    for (int i = 0; i < p.nprod; ++i) {
        p.prod_a[i] = tab[i][0];
        p.prod_b[i] = tab[i][1];
    }
    p.total_kb = kb_per_prod * p.nprod;
    p.bias = e->bias;
    p.relu = e->relu;
    p.accumulate = e->accumulate;
    p.gate = reinterpret_cast<const __nv_bfloat16*>(e->gate);
    p.gate_ld = e->gate_ld;
    p.gate_scale = e->gate_scale;
    p.drop = drop_make(e->drop_p, e->seed, e->step, e->site);
    p.drop_ld = e->drop_p > 0.f ? e->drop_ld : 0;
    p.resid = e->resid;
    p.resid_ld = e->resid_ld;
    p.out_f32 = e->out_f32;
    p.out_f32_ld = e->out_f32_ld;
    p.out_pl = reinterpret_cast<__nv_bfloat16*>(e->out_planes.ptr);
    p.out_pl_ld = e->out_planes.ld;
    p.out_pl_stride = e->out_planes.plane_stride;
    p.out_pl_n = e->out_planes.planes;
    M2F_REQUIRE(p.out_f32 || p.out_pl, M2F_ERR_ARG, "gemm: no output given");
    M2F_REQUIRE(!p.out_pl || (p.out_pl_n >= 1 && p.out_pl_n <= 3 && (p.out_pl_ld % 4) == 0 &&
                              (reinterpret_cast<uintptr_t>(p.out_pl) & 7) == 0 && (p.out_pl_n == 1 || (p.out_pl_stride % 4) == 0)),
                M2F_ERR_ALIGN, "gemm: planes output needs ld %% 4 == 0 and an 8-byte aligned base");
    M2F_REQUIRE(!p.gate || ((p.gate_ld % 4) == 0 && (reinterpret_cast<uintptr_t>(p.gate) & 7) == 0), M2F_ERR_ALIGN,
                "gemm: gate needs ld %% 4 == 0 and an 8-byte aligned base");
    M2F_REQUIRE(!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0, M2F_ERR_ALIGN, "gemm: bias must be 16-byte aligned");
    M2F_REQUIRE(!p.out_f32 || (reinterpret_cast<uintptr_t>(p.out_f32) & 15) == 0, M2F_ERR_ALIGN, "gemm: out_f32 must be 16-byte aligned");
    M2F_REQUIRE(!p.resid || (reinterpret_cast<uintptr_t>(p.resid) & 15) == 0, M2F_ERR_ALIGN, "gemm: resid must be 16-byte aligned");
    M2F_REQUIRE(e->drop_p >= 0.f && e->drop_p < 1.f, M2F_ERR_ARG, "gemm: dropout p must be in [0,1)");
    return M2F_OK;
}

int gemm_launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int block_n, int a_mn, int b_mn,
                cudaStream_t stream) {
#define M2F_GEMM_CASE(BN)                                                              \
    if (block_n == BN) {                                                               \
        if (!a_mn && !b_mn) return launch_gemm<BN, false, false>(ta, tb, p, stream);   \
        if (!a_mn && b_mn) return launch_gemm<BN, false, true>(ta, tb, p, stream);     \
        if (a_mn && b_mn) return launch_gemm<BN, true, true>(ta, tb, p, stream);       \
        return launch_gemm<BN, true, false>(ta, tb, p, stream);                        \
    }
    M2F_GEMM_CASE(64)
    M2F_GEMM_CASE(128)
    M2F_GEMM_CASE(256)
#undef M2F_GEMM_CASE
    set_error("gemm: unsupported BLOCK_N %d", block_n);
    return M2F_ERR_UNSUPPORTED;
}

}  // namespace m2f

extern "C" int m2f_gemm(const m2f_planes_t* A, int a_mn_major, const m2f_planes_t* B, int b_mn_major,
                        const m2f_kseg_t* segs, int nsegs, int64_t M, int64_t N, const m2f_epilogue_t* epi,
                        void* stream) {
    using namespace m2f;
    GemmParams p;
    M2F_TRY(gemm_fill_params(p, A, a_mn_major, B, b_mn_major, segs, nsegs, M, N, epi));
    const int bn = gemm_block_n(M, N);
    CUtensorMap ta, tb;
    M2F_TRY(make_planes_tmap(&ta, A, a_mn_major ? BLOCK_K : BLOCK_M));
    M2F_TRY(make_planes_tmap(&tb, B, b_mn_major ? BLOCK_K : bn));
    return gemm_launch(ta, tb, p, bn, a_mn_major, b_mn_major, reinterpret_cast<cudaStream_t>(stream));
}
