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

#include <cstdlib>
#include <cstring>
#include "common.cuh"
#include "gemm.cuh"
#include "ptx.cuh"

namespace m2f {

static unsigned long long* g_trace = nullptr;
static int g_dbg = 0;  // experiment switches (m2f_debug_set_gemm_flags): 1 skip global stores, 2 skip tcgen05.ld, 4 skip staging  // set through m2f_debug_set_gemm_trace (profiling aid)

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)::"memory");
    return t;
}
#define M2F_TRACE(slot)                                                                                          \
    do {                                                                                                         \
        if (p.trace) p.trace[(size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 8 + (slot)] = gtimer();             \
    } while (0)

// CTA2: the CTA is half of a cta_group::2 pair -- the pair computes a 256 x BLOCK_N tile, this CTA owns 128 rows of it
// and stages only BLOCK_N / 2 rows of B (twice the flops per byte an SM ingests from L2).
#ifndef M2F_PAIR_RING_KB
#define M2F_PAIR_RING_KB 192
#endif
#ifndef M2F_PAIR_EPI_WARPS
#define M2F_PAIR_EPI_WARPS 8
#endif
template <int BLOCK_N, bool CTA2>
struct GemmCfg {
    static constexpr int B_ROWS = CTA2 ? BLOCK_N / 2 : BLOCK_N;
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    // Single CTAs: 96 KB of operand ring, so that TWO CTAs are co-resident per SM and one's epilogue overlaps the other's
    // prologue / main loop (TMEM: 2 x BLOCK_N <= 512 columns); 4 epilogue warps each.
    // Pairs: one CTA per SM with a 192 KB ring (enough bytes in flight to cover the L2 latency at the rate the
    // 256-row MMA consumes them) and 8 epilogue warps -- the epilogue of a 128 x 256 tile is instruction-latency bound,
    // two warps per TMEM lane group halve it.
    // Build-time experiment (M2F_NVCC_DEFS="-DM2F_PAIR_RING_KB=96 -DM2F_PAIR_EPI_WARPS=4", see _build.py): pairs with a
    // half-size ring and four epilogue warps, so that two pairs share an SM and one's epilogue overlaps the other's main loop.
    static constexpr int PAIR_RING_KB = M2F_PAIR_RING_KB;
    static constexpr int CTAS_PER_SM = CTA2 ? (PAIR_RING_KB <= 100 ? 2 : 1) : 2;
    static constexpr int EPI_WARPS = CTA2 ? M2F_PAIR_EPI_WARPS : 4;
    static constexpr int THREADS = 64 + 32 * EPI_WARPS;
    static constexpr int STAGES = ((CTA2 ? PAIR_RING_KB : 96) * 1024) / STAGE_BYTES;
    static constexpr int BAR_BYTES = (2 * STAGES + 1) * 8 + 16 + 8 * 8;  // ring, accumulator-ready, TMEM slot, 8 epilogue-load barriers
    static constexpr int BIAS_BYTES = BLOCK_N * 4;  // the tile's bias values, staged while the main loop runs (so not in the ring)
    // the ring doubles as the epilogue's staging area: 6 stages of a 192-wide pair tile (168 KB) are less than the 8 x 23 KB the
    // TMA epilogue stages, so the region is sized by the larger of the two
    static constexpr int STG_BYTES = EPI_WARPS * (EPI2_WARP_BYTES > EPI_STG_FLOATS * 4 ? EPI2_WARP_BYTES : EPI_STG_FLOATS * 4);
    static constexpr int RING_BYTES = STAGES * STAGE_BYTES > STG_BYTES ? STAGES * STAGE_BYTES : STG_BYTES;
    static constexpr int SMEM_BYTES = RING_BYTES + BAR_BYTES + BIAS_BYTES + 16 + 1024;  // + alignment slack
    static constexpr int TMEM_COLS = BLOCK_N < 32 ? 32 : (BLOCK_N == 192 ? 256 : BLOCK_N);  // allocations are powers of two
    static_assert(BLOCK_N != 192 || CTA2, "192-wide tiles exist for CTA pairs only (K-major B: 96 rows per CTA)");
    static_assert(STAGES >= 2, "pipeline too shallow");
    static_assert(STG_BYTES <= RING_BYTES && SMEM_BYTES <= 227 * 1024, "ring / epilogue staging must fit in shared memory");
};

__device__ __forceinline__ uint2 pack_bf16x4(const float v[4]) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
    const __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 r;
    r.x = *reinterpret_cast<const uint32_t*>(&a);
    r.y = *reinterpret_cast<const uint32_t*>(&b);
    return r;
}
__device__ __forceinline__ void unpack_sub(const uint2 h, float v[4]) {  // v -= float(h)
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&h.x);
    const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&h.y);
    v[0] -= __low2float(a); v[1] -= __high2float(a); v[2] -= __low2float(b); v[3] -= __high2float(b);
}

// Epilogue parameters held in registers by the epilogue warps (loaded once, before the accumulator wait).
struct EpiRegs {
    const float* bias;
    const __nv_bfloat16* gate;
    const float* resid;
    float* out_f32;
    __nv_bfloat16* out_pl;
    float* colsum;
    int64_t gate_ld, resid_ld, out_f32_ld, out_pl_ld, out_pl_stride, drop_ld;
    float gate_scale, floor_;
    int M, N, out_pl_n, accumulate, vec4, drop_on;
};

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Epilogue of one 32-row x 32-column chunk.  `stg` (shared-space address) holds the raw accumulators
// [32 rows][EPI_STG_LD floats]; lane l handles columns 4*(l%8)..+3 of rows it*4 + l/8, it = 0..7, so every global
// access of the warp covers 4 full 128-byte rows.  Fast path in phases (all shared loads, all global loads, math,
// all stores) so that the memory latencies of the 8 row groups overlap; the rare ragged path goes element by element.
__device__ __forceinline__ void epilogue_chunk(const EpiRegs& e, const Drop& drop, uint32_t stg, int lane, int64_t m0, int n0,
                                               const float4 b, const uint2 (&g)[8], const float4 (&q)[8]) {
    const int cg = (lane & 7) * 4;
    const int n = n0 + cg;
    const int r0 = lane >> 3;
    const bool vec = e.vec4 && (n + 4 <= e.N);
    float cs0 = 0.f, cs1 = 0.f, cs2 = 0.f, cs3 = 0.f;
    if (vec) {
        float4 v[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) v[it] = lds128(stg + (uint32_t)(((it * 4 + r0) * EPI_STG_LD + cg) * 4));
        bool ok[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) ok[it] = (m0 + it * 4 + r0) < e.M;
        if (e.bias) {  // b: this lane's four bias values, loaded by the caller ahead of the accumulator read
#pragma unroll
            for (int it = 0; it < 8; ++it) { v[it].x += b.x; v[it].y += b.y; v[it].z += b.z; v[it].w += b.w; }
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            v[it].x = fmaxf(v[it].x, e.floor_); v[it].y = fmaxf(v[it].y, e.floor_);
            v[it].z = fmaxf(v[it].z, e.floor_); v[it].w = fmaxf(v[it].w, e.floor_);
        }
        if (e.gate) {  // g: prefetched by the caller
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const __nv_bfloat162 g0 = *reinterpret_cast<const __nv_bfloat162*>(&g[it].x);
                const __nv_bfloat162 g1 = *reinterpret_cast<const __nv_bfloat162*>(&g[it].y);
                v[it].x = __low2float(g0) > 0.f ? v[it].x * e.gate_scale : 0.f;
                v[it].y = __high2float(g0) > 0.f ? v[it].y * e.gate_scale : 0.f;
                v[it].z = __low2float(g1) > 0.f ? v[it].z * e.gate_scale : 0.f;
                v[it].w = __high2float(g1) > 0.f ? v[it].w * e.gate_scale : 0.f;
            }
        }
        if (e.drop_on) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                float dm[4];
                drop_mult4(drop, (uint64_t)((m0 + it * 4 + r0) * e.drop_ld + n) >> 2, dm);
                v[it].x *= dm[0]; v[it].y *= dm[1]; v[it].z *= dm[2]; v[it].w *= dm[3];
            }
        }
        if (e.resid) {  // q: prefetched by the caller
#pragma unroll
            for (int it = 0; it < 8; ++it) { v[it].x += q[it].x; v[it].y += q[it].y; v[it].z += q[it].z; v[it].w += q[it].w; }
        }
        if (e.out_f32) {
            if (e.accumulate) {
                float4 q[8];
#pragma unroll
                for (int it = 0; it < 8; ++it)
                    q[it] = ok[it] ? *reinterpret_cast<const float4*>(e.out_f32 + (m0 + it * 4 + r0) * e.out_f32_ld + n) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int it = 0; it < 8; ++it) { v[it].x += q[it].x; v[it].y += q[it].y; v[it].z += q[it].z; v[it].w += q[it].w; }
            }
#pragma unroll
            for (int it = 0; it < 8; ++it)
                if (ok[it]) *reinterpret_cast<float4*>(e.out_f32 + (m0 + it * 4 + r0) * e.out_f32_ld + n) = v[it];
        }
        if (e.colsum) {
#pragma unroll
            for (int it = 0; it < 8; ++it)
                if (ok[it]) { cs0 += v[it].x; cs1 += v[it].y; cs2 += v[it].z; cs3 += v[it].w; }
        }
        if (e.out_pl) {  // hi / mid / lo bf16 planes, registers only
            for (int pl = 0; pl < e.out_pl_n; ++pl) {
                __nv_bfloat16* base = e.out_pl + pl * e.out_pl_stride + n;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    float t[4] = {v[it].x, v[it].y, v[it].z, v[it].w};
                    const uint2 h = pack_bf16x4(t);
                    if (ok[it]) *reinterpret_cast<uint2*>(base + (m0 + it * 4 + r0) * e.out_pl_ld) = h;
                    if (pl + 1 < e.out_pl_n) {
                        unpack_sub(h, t);  // residual feeds the next plane
                        v[it] = make_float4(t[0], t[1], t[2], t[3]);
                    }
                }
            }
        }
    } else if (n < e.N) {
        // rare: ragged N or unaligned leading dimensions, element by element
#pragma unroll 1
        for (int it = 0; it < 8; ++it) {
            const int64_t m = m0 + it * 4 + r0;
            if (m >= e.M) continue;
            const float4 acc = lds128(stg + (uint32_t)(((it * 4 + r0) * EPI_STG_LD + cg) * 4));
            const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int nn = n + i;
                float x = 0.f;
                if (nn < e.N) {
                    x = a4[i];
                    if (e.bias) x += e.bias[nn];
                    x = fmaxf(x, e.floor_);
                    if (e.gate) x = __bfloat162float(e.gate[m * e.gate_ld + nn]) > 0.f ? x * e.gate_scale : 0.f;
                    if (e.drop_on) x *= drop_mult1(drop, (uint64_t)(m * e.drop_ld + nn));
                    if (e.resid) x += e.resid[m * e.resid_ld + nn];
                    if (e.out_f32) {
                        float* o = e.out_f32 + m * e.out_f32_ld + nn;
                        if (e.accumulate) x += *o;
                        *o = x;
                    }
                    if (e.out_pl) store_planes(e.out_pl, e.out_pl_stride, e.out_pl_n, m * e.out_pl_ld + nn, x);
                }
                if (i == 0) cs0 += x; else if (i == 1) cs1 += x; else if (i == 2) cs2 += x; else cs3 += x;
            }
        }
    }
    if (e.colsum) {  // bias gradient: column sums of what was stored, 4 row-groups of the warp folded first
        cs0 += __shfl_xor_sync(0xffffffffu, cs0, 8);  cs0 += __shfl_xor_sync(0xffffffffu, cs0, 16);
        cs1 += __shfl_xor_sync(0xffffffffu, cs1, 8);  cs1 += __shfl_xor_sync(0xffffffffu, cs1, 16);
        cs2 += __shfl_xor_sync(0xffffffffu, cs2, 8);  cs2 += __shfl_xor_sync(0xffffffffu, cs2, 16);
        cs3 += __shfl_xor_sync(0xffffffffu, cs3, 8);  cs3 += __shfl_xor_sync(0xffffffffu, cs3, 16);
        if (lane < 8) {
            if (n < e.N) atomicAdd(e.colsum + n, cs0);
            if (n + 1 < e.N) atomicAdd(e.colsum + n + 1, cs1);
            if (n + 2 < e.N) atomicAdd(e.colsum + n + 2, cs2);
            if (n + 3 < e.N) atomicAdd(e.colsum + n + 3, cs3);
        }
    }
}

__device__ __forceinline__ float4 lds128u(uint32_t addr) { return lds128(addr); }

// TMA epilogue of one warp.  The warp owns TMEM lanes 32*lg..+31 = 32 output rows; lane t holds row t of each
// 32-column chunk exactly as tcgen05.ld delivers it, so nothing is transposed: per-element inputs (gate, residual,
// previous value when accumulating) arrive as TMA tiles in shared memory, results leave as TMA tile stores, and the
// hardware does the address arithmetic, the coalescing and the clipping at M / N.  Swizzled tiles keep the
// "lane = row" shared-memory accesses conflict-free: 16-byte chunk j of row t sits at chunk j ^ (t & 7) (128-byte
// rows, fp32) or j ^ ((t >> 1) & 3) (64-byte rows, bf16).
// FEAT: the epilogue's feature set as a compile-time mask for the combinations the training step launches most (the
// feature tests then fold away: ~180 instead of ~330 instructions per 32x32 chunk), or -1 = decide everything at run time.
enum : int { F_BIAS = 1, F_GATE = 2, F_DROP = 4, F_RES = 8, F_ACC = 16, F_OF32 = 32, F_OPL = 64, F_COLSUM = 128, F_RELU = 256, F_P1 = 512, F_P2 = 1024 };
#define M2F_FEAT(bit, runtime) (FEAT < 0 ? (runtime) : ((FEAT & (bit)) != 0))

template <int BLOCK_N, int EPI_WARPS, int FEAT>
__device__ __forceinline__ void epilogue_tma(const GemmParams& p, const EpiMaps& em, uint8_t* smem, uint32_t s_bias, uint64_t* epi_bar,
                                             uint64_t* tmem_full_bar, uint32_t tmem_base, int warp, int lane, int m_blk, int n_blk) {
    const int lg = warp & 3, ew = warp - 2;
    // staging of this warp: with at most one output plane the output tiles are double-buffered (a tile store's read of
    // shared memory takes about as long as producing the next chunk), with 2-3 planes there is room for one set only
    const uint32_t s_base = ptx::smem_u32(smem) + (uint32_t)(ew * EPI2_WARP_BYTES);
    const bool f_bias = M2F_FEAT(F_BIAS, p.bias != nullptr), f_relu = M2F_FEAT(F_RELU, p.relu != 0);
    const bool f_of32 = M2F_FEAT(F_OF32, p.out_f32 != nullptr), f_opl = M2F_FEAT(F_OPL, p.out_pl != nullptr);
    const bool f_colsum = M2F_FEAT(F_COLSUM, p.colsum != nullptr), f_drop = M2F_FEAT(F_DROP, p.drop.enabled != 0);
    // P2: exactly two output planes and nothing else staged (the forward GEMMs of the 2-plane-forward precision modes): the
    // two plane tiles of a chunk are double-buffered in the region the fp32 tiles would take (2 x (2 x 2 KB))
    constexpr bool P2 = FEAT >= 0 && (FEAT & F_P2) != 0;
    static_assert(!P2 || (FEAT & (F_OF32 | F_COLSUM | F_P1)) == 0, "F_P2 is for planes-only outputs");
    const int n_pl = f_opl ? ((FEAT >= 0 && (FEAT & F_P1)) ? 1 : (P2 ? 2 : p.out_pl_n)) : 0;
    const bool dbl = n_pl <= 1 || P2;
    const uint32_t s_pl0 = P2 ? s_base : s_base + (dbl ? 8192u : 4096u);
    const uint32_t s_gate = s_base + (dbl ? 12288u : 10240u), s_res = s_gate + 2048u, s_acc = s_res + 4096u;
    uint64_t* bar = &epi_bar[ew];
    const int m0 = m_blk * BLOCK_M + lg * 32;
    const int nt0 = n_blk * BLOCK_N;
    const bool rows_live = m0 < p.M;
    const bool has_gate = M2F_FEAT(F_GATE, p.gate != nullptr), has_res = M2F_FEAT(F_RES, p.resid != nullptr);
    const bool has_acc = M2F_FEAT(F_ACC, p.accumulate && p.out_f32);
    const uint32_t in_bytes = (has_gate ? 2048u : 0u) + (has_res ? 4096u : 0u) + (has_acc ? 4096u : 0u);
    constexpr int CSTEP = EPI_WARPS / 4;
    const int c_first = ew >> 2;
    auto issue_loads = [&](int c) {
        const int n0 = nt0 + c * 32;
        if (lane == 0 && in_bytes && n0 < p.N) {
            ptx::mbar_expect_tx(bar, in_bytes);
            if (has_gate) ptx::tma_load_2d(s_gate, &em.gate, bar, n0, m0);
            if (has_res) ptx::tma_load_2d(s_res, &em.resid, bar, n0, m0);
            if (has_acc) ptx::tma_load_2d(s_acc, &em.out_f32, bar, n0, m0);
        }
    };
    // bias of this CTA's columns -> shared (read back as broadcasts); its own region: the operand ring is still live
    if (f_bias) {
        for (int i = (warp - 2) * 32 + lane; i < BLOCK_N; i += EPI_WARPS * 32) {
            const int n = nt0 + i;
            const float b = n < p.N ? p.bias[n] : 0.f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(s_bias + (uint32_t)i * 4u), "f"(b) : "memory");
        }
    }
    const Drop drop = drop_resolve(p.drop);
    if (f_bias) ptx::named_bar_sync(1, EPI_WARPS * 32);  // bias visible to every epilogue warp
    ptx::mbar_wait(tmem_full_bar, 0);
    ptx::tcgen05_fence_after();
    if (threadIdx.x == 64) M2F_TRACE(4);
    if (!rows_live) return;
    issue_loads(c_first);  // the ring is free from here on: it becomes the staging area
    uint32_t phase = 0, buf = 0;
    const int row = m0 + lane;
    const uint32_t sw128 = (uint32_t)(lane & 7), sw64 = (uint32_t)((lane >> 1) & 3);
#pragma unroll 1
    for (int c = c_first; c < BLOCK_N / 32; c += CSTEP) {
        const int n0 = nt0 + c * 32;
        if (n0 >= p.N) break;
        float x[32];
        {
            uint32_t r[32];
            ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(c * 32), r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(r[j]);
        }
        if (f_bias) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 b = lds128u(s_bias + (uint32_t)(c * 32 + j * 4) * 4u);
                x[4 * j] += b.x; x[4 * j + 1] += b.y; x[4 * j + 2] += b.z; x[4 * j + 3] += b.w;
            }
        }
        if (f_relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
        }
        if (in_bytes) {
            ptx::mbar_wait(bar, phase);
            phase ^= 1;
        }
        if (has_gate) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 g = lds128u(s_gate + (uint32_t)lane * 64u + (((uint32_t)j ^ sw64) << 4));
                const uint32_t w[4] = {__float_as_uint(g.x), __float_as_uint(g.y), __float_as_uint(g.z), __float_as_uint(g.w)};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const __nv_bfloat162 gg = *reinterpret_cast<const __nv_bfloat162*>(&w[q]);
                    x[8 * j + 2 * q] = __low2float(gg) > 0.f ? x[8 * j + 2 * q] * p.gate_scale : 0.f;
                    x[8 * j + 2 * q + 1] = __high2float(gg) > 0.f ? x[8 * j + 2 * q + 1] * p.gate_scale : 0.f;
                }
            }
        }
        if (f_drop) {
            // drop_ld % 8 == 0 here (gemm_make_epi_maps): the row's 32 columns are four whole Philox calls
            const uint64_t e8 = (uint64_t)((int64_t)row * p.drop_ld + n0) >> 3;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float dm[8];
                drop_mult8(drop, e8 + j, dm);
#pragma unroll
                for (int i = 0; i < 8; ++i) x[8 * j + i] *= dm[i];
            }
        }
        if (has_res) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 q = lds128u(s_res + (uint32_t)lane * 128u + (((uint32_t)j ^ sw128) << 4));
                x[4 * j] += q.x; x[4 * j + 1] += q.y; x[4 * j + 2] += q.z; x[4 * j + 3] += q.w;
            }
        }
        if (has_acc) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 q = lds128u(s_acc + (uint32_t)lane * 128u + (((uint32_t)j ^ sw128) << 4));
                x[4 * j] += q.x; x[4 * j + 1] += q.y; x[4 * j + 2] += q.z; x[4 * j + 3] += q.w;
            }
        }
        __syncwarp();                                   // every lane has consumed the input tiles
        if (c + CSTEP < BLOCK_N / 32) issue_loads(c + CSTEP);
        const uint32_t s_of = s_base + buf * 4096u, s_pl = s_pl0 + buf * (P2 ? 4096u : 2048u);
        if (lane == 0) {  // the stores that last used this staging set have read it
            if (dbl) ptx::bulk_wait_read1();
            else ptx::bulk_wait_read0();
        }
        if (dbl) buf ^= 1u;
        __syncwarp();
        if (f_colsum && row >= p.M) {
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = 0.f;    // rows past M are clipped on store but must not reach the column sums
        }
        if (f_of32 || f_colsum) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                sts128(s_of + (uint32_t)lane * 128u + (((uint32_t)j ^ sw128) << 4), __float_as_uint(x[4 * j]), __float_as_uint(x[4 * j + 1]),
                       __float_as_uint(x[4 * j + 2]), __float_as_uint(x[4 * j + 3]));
        }
        if (f_opl) {
            for (int pl = 0; pl < n_pl; ++pl) {
                uint32_t h[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const __nv_bfloat162 b2 = __floats2bfloat162_rn(x[2 * j], x[2 * j + 1]);
                    h[j] = *reinterpret_cast<const uint32_t*>(&b2);
                    if (pl + 1 < n_pl) {
                        x[2 * j] -= __low2float(b2);
                        x[2 * j + 1] -= __high2float(b2);
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    sts128(s_pl + (uint32_t)pl * 2048u + (uint32_t)lane * 64u + (((uint32_t)j ^ sw64) << 4), h[4 * j], h[4 * j + 1], h[4 * j + 2],
                           h[4 * j + 3]);
            }
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
            if (f_of32) ptx::tma_store_2d(&em.out_f32, s_of, n0, m0);
            if (f_opl)
                for (int pl = 0; pl < n_pl; ++pl) ptx::tma_store_3d(&em.out_pl, s_pl + (uint32_t)pl * 2048u, n0, m0, pl);
            ptx::bulk_commit();
        }
        if (f_colsum) {  // bias gradient: lane j sums column j of the staged fp32 tile (planes were taken from the same values)
            float cs = 0.f;
            const uint32_t cj = (uint32_t)(lane >> 2), wj = (uint32_t)(lane & 3) * 4u;
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                float v;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(s_of + (uint32_t)r * 128u + ((cj ^ (uint32_t)(r & 7)) << 4) + wj));
                cs += v;
            }
            if (n0 + lane < p.N) atomicAdd(p.colsum + n0 + lane, cs);
        }
    }
    if (lane == 0) ptx::bulk_wait_read0();  // the staging tiles must outlive the stores' reads; the writes complete on their own
    __syncwarp();
    if (threadIdx.x == 64) M2F_TRACE(5);
}

template <int BLOCK_N, bool A_MN, bool B_MN, bool CTA2>
__global__ void __launch_bounds__(GemmCfg<BLOCK_N, CTA2>::THREADS, GemmCfg<BLOCK_N, CTA2>::CTAS_PER_SM)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const __grid_constant__ EpiMaps em,
            const GemmParams p) {
    using Cfg = GemmCfg<BLOCK_N, CTA2>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::RING_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
    uint64_t* epi_bar = tmem_full_bar + 3;  // after the 16-byte TMEM slot

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_blk = blockIdx.x, n_blk = blockIdx.y;
    const uint32_t cta_rank = CTA2 ? ptx::cluster_ctarank() : 0u;  // (2,1,1) cluster: rank = blockIdx.x & 1
    if (threadIdx.x == 0) M2F_TRACE(0);

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
        for (int s = 0; s < 8; ++s) ptx::mbar_init(&epi_bar[s], 1);
        ptx::fence_barrier_init();
    }
    if (warp == 2) {
        if constexpr (CTA2) ptx::tmem_alloc_2cta(tmem_slot, Cfg::TMEM_COLS);
        else ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    }
    ptx::tcgen05_fence_before();
    if constexpr (CTA2) ptx::cluster_sync();  // the peer's barriers must exist before anything is signalled across
    else __syncthreads();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Programmatic dependent launch: everything above overlapped the previous kernel's tail.  The producer additionally
    // starts the loads of a B operand that no kernel of the step writes (weights) before it waits.
    if (!(warp == 0 && p.b_static)) {
        pdl_wait();
        pdl_trigger();
    }
    if (threadIdx.x == 0) M2F_TRACE(1);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (ptx::elect_one()) {
            // issue one k-block's B tile(s); in a pair both CTAs' bytes are credited to the leader's barrier
            auto load_b = [&](int stage, int kbb, int pb) {
                uint8_t* sb = smem_b + stage * Cfg::B_BYTES;
                const int nb0 = n_blk * BLOCK_N + (int)cta_rank * Cfg::B_ROWS;  // this CTA's share of the B tile
                if constexpr (CTA2) {
                    const uint32_t fb = ptx::mapa(ptx::smem_u32(&full_bar[stage]), 0);
                    if (!B_MN) {
                        ptx::tma_load_3d_2cta(sb, &tma_b, fb, kbb, nb0, pb);
                    } else if constexpr (BLOCK_N == 192) {
                        // 96 MN-major columns per CTA: three 32-column atoms of the 64-byte swizzle (make_planes_tmap_sw64)
#pragma unroll
                        for (int j = 0; j < 3; ++j)
                            ptx::tma_load_3d_2cta(sb + j * (32 * BLOCK_K * 2), &tma_b, fb, nb0 + j * 32, kbb, pb);
                    } else {
#pragma unroll
                        for (int j = 0; j < Cfg::B_ROWS / 64; ++j)
                            ptx::tma_load_3d_2cta(sb + j * (64 * BLOCK_K * 2), &tma_b, fb, nb0 + j * 64, kbb, pb);
                    }
                } else {
                    if (!B_MN) {
                        ptx::tma_load_3d(sb, &tma_b, &full_bar[stage], kbb, nb0, pb);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_N / 64; ++j)
                            ptx::tma_load_3d(sb + j * (64 * BLOCK_K * 2), &tma_b, &full_bar[stage], nb0 + j * 64, kbb, pb);
                    }
                }
            };
            auto arm = [&](int stage) {
                if constexpr (CTA2) {
                    if (cta_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
                } else {
                    ptx::mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                }
            };
            int pre = 0;  // k-blocks whose B tile is already in flight
            if (p.b_static) {
                int it0 = 0;
                for (int pr = 0; pr < p.nprod && it0 < STAGES; ++pr)
                    for (int sg = 0; sg < p.nsegs && it0 < STAGES; ++sg)
                        for (int kb = 0; kb < p.seg_nkb[sg] && it0 < STAGES; ++kb, ++it0) {
                            arm(it0);
                            load_b(it0, p.seg_b_k0[sg] + kb * BLOCK_K, p.prod_b[pr]);
                        }
                pre = it0;
                pdl_wait();
                pdl_trigger();
            }
            int it = 0;
            for (int pr = 0; pr < p.nprod; ++pr) {
                const int pa = p.prod_a[pr], pb = p.prod_b[pr];
                for (int sg = 0; sg < p.nsegs; ++sg) {
                    for (int kb = 0; kb < p.seg_nkb[sg]; ++kb, ++it) {
                        const int stage = it % STAGES;
                        const uint32_t phase = (it / STAGES) & 1;
                        const int ka = p.seg_a_k0[sg] + kb * BLOCK_K;
                        const int kbb = p.seg_b_k0[sg] + kb * BLOCK_K;
                        uint8_t* sa = smem_a + stage * Cfg::A_BYTES;
                        if (it >= pre) {
                            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                            arm(stage);
                            load_b(stage, kbb, pb);
                        }
                        if constexpr (CTA2) {
                            const uint32_t fb = ptx::mapa(ptx::smem_u32(&full_bar[stage]), 0);
                            if (!A_MN) {
                                ptx::tma_load_3d_2cta(sa, &tma_a, fb, ka, m_blk * BLOCK_M, pa);
                            } else {
#pragma unroll
                                for (int j = 0; j < BLOCK_M / 64; ++j)
                                    ptx::tma_load_3d_2cta(sa + j * (64 * BLOCK_K * 2), &tma_a, fb, m_blk * BLOCK_M + j * 64, ka, pa);
                            }
                        } else {
                            if (!A_MN) {
                                ptx::tma_load_3d(sa, &tma_a, &full_bar[stage], ka, m_blk * BLOCK_M, pa);
                            } else {
#pragma unroll
                                for (int j = 0; j < BLOCK_M / 64; ++j)
                                    ptx::tma_load_3d(sa + j * (64 * BLOCK_K * 2), &tma_a, &full_bar[stage],
                                                     m_blk * BLOCK_M + j * 64, ka, pa);
                            }
                        }
                    }
                }
            }
        }
        else if (p.b_static) {  // the other lanes of the producer warp
            pdl_wait();
            pdl_trigger();
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (cta_rank == 0 && ptx::elect_one()) {  // in a pair only the leader issues; its MMAs read both CTAs' smem
            constexpr uint32_t idesc = ptx::instr_desc_bf16(CTA2 ? 2 * BLOCK_M : BLOCK_M, BLOCK_N, A_MN, B_MN);
            // K-major SW128: 8-row groups 1024 B apart (SBO), one atom along K (LBO unused);
            //   advancing K by 16 elements = +32 B inside the swizzle atom.
            // MN-major SW128: 64-element MN atoms BLOCK_K*128 B apart (LBO), 8-k-row groups 1024 B apart (SBO);
            //   advancing K by 16 rows = +2048 B.
            // MN-major SW64 (B of the 192-wide pair tiles: 96 columns per CTA): 32-element atoms BLOCK_K*64 B apart (LBO), 8-k-row
            //   groups 512 B apart (SBO); advancing K by 16 rows = +1024 B.
            constexpr bool B_SW64 = B_MN && BLOCK_N == 192;
            constexpr uint32_t A_LBO = A_MN ? BLOCK_K * 128 : 0, B_LBO = B_SW64 ? BLOCK_K * 64 : (B_MN ? BLOCK_K * 128 : 0);
            constexpr uint32_t A_KSTEP = A_MN ? UMMA_K * 128 : UMMA_K * 2, B_KSTEP = B_SW64 ? UMMA_K * 64 : (B_MN ? UMMA_K * 128 : UMMA_K * 2);
            for (int it = 0; it < p.total_kb; ++it) {
                const int stage = it % STAGES;
                const uint32_t phase = (it / STAGES) & 1;
                ptx::mbar_wait(&full_bar[stage], phase);
                if (it == 0) M2F_TRACE(2);
                ptx::tcgen05_fence_after();
                const uint32_t sa = ptx::smem_u32(smem_a + stage * Cfg::A_BYTES);
                const uint32_t sb = ptx::smem_u32(smem_b + stage * Cfg::B_BYTES);
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                    const uint64_t ad = ptx::smem_desc_sw128(sa + k * A_KSTEP, A_LBO, 1024);
                    const uint64_t bd = B_SW64 ? ptx::smem_desc_sw64(sb + k * B_KSTEP, B_LBO, 512) : ptx::smem_desc_sw128(sb + k * B_KSTEP, B_LBO, 1024);
                    if constexpr (CTA2) ptx::umma_bf16_2cta(tmem_base, ad, bd, idesc, (it > 0 || k > 0) ? 1u : 0u);
                    else ptx::umma_bf16(tmem_base, ad, bd, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                // frees the smem slot (in both CTAs of a pair) once these MMAs retire
                if constexpr (CTA2) ptx::umma_commit_2cta(&empty_bar[stage], 3);
                else ptx::umma_commit(&empty_bar[stage]);
            }
            if constexpr (CTA2) ptx::umma_commit_2cta(tmem_full_bar, 3);
            else ptx::umma_commit(tmem_full_bar);
            M2F_TRACE(3);
        }
        __syncwarp();
    } else if (p.epi_tma) {
        // ------------------------------------------------------------------ epilogue, TMA tiles in and out
        const uint32_t s_bias = (ptx::smem_u32(epi_bar + 8) + 15u) & ~15u;
        const int feat = (p.bias ? F_BIAS : 0) | (p.gate ? F_GATE : 0) | (p.drop.enabled ? F_DROP : 0) | (p.resid ? F_RES : 0) |
                         ((p.accumulate && p.out_f32) ? F_ACC : 0) | (p.out_f32 ? F_OF32 : 0) | (p.out_pl ? F_OPL : 0) |
                         (p.colsum ? F_COLSUM : 0) | (p.relu ? F_RELU : 0) | ((p.out_pl && p.out_pl_n == 1) ? F_P1 : 0) |
                         ((p.out_pl && p.out_pl_n == 2 && !p.out_f32 && !p.colsum) ? F_P2 : 0);
#define M2F_EPI_CASE(MASK)                                                                                                            \
    case (MASK):                                                                                                                      \
        epilogue_tma<BLOCK_N, Cfg::EPI_WARPS, (MASK)>(p, em, smem, s_bias, epi_bar, tmem_full_bar, tmem_base, warp, lane, m_blk, n_blk); \
        break;
        switch (feat) {  // warp-uniform
            M2F_EPI_CASE(F_OF32)                                             // weight gradients
            M2F_EPI_CASE(F_BIAS | F_OPL | F_P1)                              // QKV / K / QV projections
            M2F_EPI_CASE(F_BIAS | F_DROP | F_RES | F_OF32)                   // attention out-projection, FFN2 (pre-LN sums)
            M2F_EPI_CASE(F_BIAS | F_RELU | F_DROP | F_OPL | F_P1)            // FFN1, fusion linear
            M2F_EPI_CASE(F_GATE | F_OPL | F_P1 | F_COLSUM)                   // d(hidden) through ReLU / dropout
            M2F_EPI_CASE(F_RES | F_OF32)                                     // data gradients joining a skip path
            M2F_EPI_CASE(F_OPL | F_P1)                                       // d(attention output)
            M2F_EPI_CASE(F_BIAS | F_RELU | F_OPL | F_P1)                     // fusion out-projection; FFN1 without dropout (eval)
            M2F_EPI_CASE(F_BIAS | F_RES | F_OF32)                            // out-projection / FFN2 without dropout (eval)
            M2F_EPI_CASE(F_BIAS | F_OPL | F_P2)                              // 2-plane forward: QKV / K / QV projections
            M2F_EPI_CASE(F_BIAS | F_RELU | F_DROP | F_OPL | F_P2)            // 2-plane forward: FFN1, fusion linear, classifier hidden
            M2F_EPI_CASE(F_BIAS | F_RELU | F_OPL | F_P2)                     // 2-plane forward: fusion out-projection
            default:
                epilogue_tma<BLOCK_N, Cfg::EPI_WARPS, -1>(p, em, smem, s_bias, epi_bar, tmem_full_bar, tmem_base, warp, lane, m_blk, n_blk);
        }
#undef M2F_EPI_CASE
        ptx::tcgen05_fence_before();
    } else {
        // ------------------------------------------------------------------ epilogue, pointer path (ragged / unaligned operands)
        // TMEM -> registers (thread = row) -> shared (transpose) -> registers (lane = 4 columns) -> global, coalesced.
        // The operand ring is free by now: the accumulator-ready commit fires after the last MMA has read it.
        EpiRegs e;
        e.bias = p.bias; e.gate = p.gate; e.resid = p.resid; e.out_f32 = p.out_f32; e.out_pl = p.out_pl; e.colsum = p.colsum;
        e.gate_ld = p.gate_ld; e.resid_ld = p.resid_ld; e.out_f32_ld = p.out_f32_ld; e.out_pl_ld = p.out_pl_ld;
        e.out_pl_stride = p.out_pl_stride; e.drop_ld = p.drop_ld; e.gate_scale = p.gate_scale; e.floor_ = p.relu ? 0.f : -INFINITY;
        e.M = p.M; e.N = p.N; e.out_pl_n = p.out_pl_n; e.accumulate = p.accumulate; e.vec4 = p.vec4; e.drop_on = p.drop.enabled;
        const Drop drop = drop_resolve(p.drop);
        if (p.dbg & 1) { e.out_f32 = nullptr; e.out_pl = nullptr; }  // experiment: epilogue without its global stores
        ptx::mbar_wait(tmem_full_bar, 0);
        if (threadIdx.x == 64) M2F_TRACE(4);
        ptx::tcgen05_fence_after();
        const int lg = warp & 3;  // TMEM lane group this warp may access
        const int ew = warp - 2;  // epilogue warps of the same lane group interleave the 32-column chunks
        const uint32_t stg = ptx::smem_u32(smem) + (uint32_t)(ew * EPI_STG_FLOATS * 4);
        const int64_t m0 = (int64_t)m_blk * BLOCK_M + lg * 32;
        if (m0 < p.M) {
#pragma unroll 1
            for (int c = ew >> 2; c < BLOCK_N / 32; c += Cfg::EPI_WARPS / 4) {
                const int n0 = n_blk * BLOCK_N + c * 32;
                if (n0 >= p.N) break;  // warp-uniform
                uint32_t r[32];
                long long c0 = 0, c1 = 0, c2 = 0;
                float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
                uint2 gate8[8];
                float4 res8[8];
                {   // issued before the TMEM read and the staging round trip so that its latency hides behind them
                    const int nb = n0 + (lane & 7) * 4;
                    const bool vecp = e.vec4 && nb + 4 <= e.N;
                    if (e.bias && vecp) bias4 = *reinterpret_cast<const float4*>(e.bias + nb);
                    const int rr = lane >> 3;
                    if (e.gate && vecp) {
#pragma unroll
                        for (int it = 0; it < 8; ++it)
                            gate8[it] = (m0 + it * 4 + rr) < e.M ? *reinterpret_cast<const uint2*>(e.gate + (m0 + it * 4 + rr) * e.gate_ld + nb)
                                                                 : make_uint2(0u, 0u);
                    }
                    if (e.resid && vecp) {
#pragma unroll
                        for (int it = 0; it < 8; ++it)
                            res8[it] = (m0 + it * 4 + rr) < e.M ? *reinterpret_cast<const float4*>(e.resid + (m0 + it * 4 + rr) * e.resid_ld + nb)
                                                                : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
                if (p.trace) c0 = clock64();
                ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(c * 32), r);
                ptx::tmem_ld_wait();
                if (p.trace) c1 = clock64();
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    sts128(stg + (uint32_t)((lane * EPI_STG_LD + j * 4) * 4), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                __syncwarp();
                if (p.trace) c2 = clock64();
                epilogue_chunk(e, drop, stg, lane, m0, n0, bias4, gate8, res8);
                if (p.trace && threadIdx.x == 64 && c < 2 * (Cfg::EPI_WARPS / 4)) {
                    unsigned long long* t2 = p.trace + (size_t)4096 * 8 + (size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 8 + (c / (Cfg::EPI_WARPS / 4)) * 4;
                    t2[0] = c1 - c0; t2[1] = c2 - c1; t2[2] = clock64() - c2; t2[3] = c0;
                }
            }
        }
        ptx::tcgen05_fence_before();
        if (threadIdx.x == 64) M2F_TRACE(5);
    }
    if constexpr (CTA2) ptx::cluster_sync();  // neither CTA may release its TMEM / exit while the pair is in flight
    else __syncthreads();
    if (threadIdx.x == 0) M2F_TRACE(6);
    if (warp == 2) {
        ptx::tcgen05_fence_after();
        if constexpr (CTA2) ptx::tmem_dealloc_2cta(tmem_base, Cfg::TMEM_COLS);
        else ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
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

// the same view with 32-column boxes and the 64-byte swizzle (attention_tc.cu: head sizes that are multiples of 32, not of 64)
int make_planes_tmap_sw64(CUtensorMap* map, const m2f_planes_t* t, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    M2F_REQUIRE(enc != nullptr, M2F_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    M2F_REQUIRE(t->ptr != nullptr && t->rows > 0 && t->cols > 0 && t->planes >= 1 && t->planes <= 3, M2F_ERR_ARG, "tensor map: bad planes view");
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(t->ptr) & 15) == 0 && (t->ld % 8) == 0 && t->ld >= t->cols, M2F_ERR_ALIGN,
                "tensor map: base must be 16-byte aligned and ld a multiple of 8");
    cuuint64_t dims[3] = {(cuuint64_t)t->cols, (cuuint64_t)t->rows, (cuuint64_t)t->planes};
    const int64_t ps = t->planes > 1 ? t->plane_stride : t->rows * t->ld;
    cuuint64_t strides[2] = {(cuuint64_t)t->ld * 2, (cuuint64_t)ps * 2};
    cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, t->ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    M2F_REQUIRE(r == CUDA_SUCCESS, M2F_ERR_CUDA, "cuTensorMapEncodeTiled (64-byte swizzle) failed with CUresult %d", (int)r);
    return M2F_OK;
}

// 2-D / 3-D tensor map with a 32 x 32 box for the TMA epilogue
static int make_epi_tmap(CUtensorMap* map, CUtensorMapDataType dt, int esize, void* ptr, int64_t rows, int64_t cols, int64_t ld, int planes,
                         int64_t plane_stride) {
    EncodeTiledFn enc = get_encode_fn();
    M2F_REQUIRE(enc != nullptr, M2F_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)planes};
    cuuint64_t strides[2] = {(cuuint64_t)ld * esize, (cuuint64_t)plane_stride * esize};
    cuuint32_t box[3] = {32, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    const int rank = planes > 0 ? 3 : 2;
    CUresult r = enc(map, dt, rank, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     esize == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    M2F_REQUIRE(r == CUDA_SUCCESS, M2F_ERR_CUDA, "cuTensorMapEncodeTiled (epilogue operand) failed with CUresult %d", (int)r);
    return M2F_OK;
}

int gemm_make_epi_maps(GemmParams& p, EpiMaps* em) {
    static const bool off = [] { const char* e = std::getenv("M2F_GEMM_EPI_TMA"); return e && e[0] == '0'; }();
    p.epi_tma = 0;
    auto ok = [](const void* ptr, int64_t ld, int esize) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld * esize) % 16 == 0; };
    if (off) return M2F_OK;
    if (p.out_f32 && !ok(p.out_f32, p.out_f32_ld, 4)) return M2F_OK;
    if (p.resid && !ok(p.resid, p.resid_ld, 4)) return M2F_OK;
    if (p.gate && !ok(p.gate, p.gate_ld, 2)) return M2F_OK;
    if (p.out_pl && (!ok(p.out_pl, p.out_pl_ld, 2) || (p.out_pl_n > 1 && (p.out_pl_stride * 2) % 16 != 0))) return M2F_OK;
    if (p.drop.enabled && (p.drop_ld & 7)) return M2F_OK;  // the TMA epilogue draws eight keep/drop decisions per Philox call
    if (p.out_f32) M2F_TRY(make_epi_tmap(&em->out_f32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p.out_f32, p.M, p.N, p.out_f32_ld, 0, 0));
    if (p.resid) M2F_TRY(make_epi_tmap(&em->resid, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(p.resid), p.M, p.N, p.resid_ld, 0, 0));
    if (p.gate) M2F_TRY(make_epi_tmap(&em->gate, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(p.gate), p.M, p.N, p.gate_ld, 0, 0));
    if (p.out_pl)
        M2F_TRY(make_epi_tmap(&em->out_pl, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.out_pl, p.M, p.N, p.out_pl_ld, p.out_pl_n,
                              p.out_pl_n > 1 ? p.out_pl_stride : (int64_t)p.M * p.out_pl_ld));
    p.epi_tma = 1;
    return M2F_OK;
}

template <int BLOCK_N, bool A_MN, bool B_MN, bool CTA2>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const EpiMaps& em, const GemmParams& p, cudaStream_t stream) {
    using Cfg = GemmCfg<BLOCK_N, CTA2>;
    auto kern = gemm_kernel<BLOCK_N, A_MN, B_MN, CTA2>;
    static bool attr_set = false;  // per instantiation
    if (!attr_set) {
        M2F_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        attr_set = true;
    }
    dim3 grid(ceil_div(p.M, BLOCK_M), ceil_div(p.N, BLOCK_N));
    if constexpr (CTA2) {
        grid.x = (grid.x + 1) & ~1u;  // whole pairs; a row tile past M loads zeros and stores nothing
        M2F_CUDA(launch_chain(kern, grid, dim3(Cfg::THREADS), Cfg::SMEM_BYTES, stream, 2, ta, tb, em, p));
    } else {
        M2F_CUDA(launch_chain(kern, grid, dim3(Cfg::THREADS), Cfg::SMEM_BYTES, stream, 1, ta, tb, em, p));
    }
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

// Tile code: BLOCK_N, plus GEMM_TILE_CTA2 when the tile is computed by a CTA pair (256 x BLOCK_N per pair).
static int g_force_bn = -1;
int gemm_b_box_rows(int tile) { return (tile & GEMM_TILE_CTA2) ? (tile & 0xfff) / 2 : (tile & 0xfff); }

int gemm_block_n(int64_t M, int64_t N, int64_t K) {
    // one 128 x BLOCK_N tile per CTA: prefer the widest tile that still yields >= 148 CTAs
    const int64_t mt = (M + BLOCK_M - 1) / BLOCK_M;
    static const int forced_env = [] { const char* e = std::getenv("M2F_GEMM_BN"); return e ? std::atoi(e) : 0; }();  // tuning aid
    const int forced = g_force_bn >= 0 ? g_force_bn : forced_env;  // m2f_debug_force_gemm_bn: the tests' way to reach every tile shape
    // M2F_GEMM_CTA2: 0 never, 1 always (default), 2 by shape.  Timed alone, a short-K GEMM is as fast with two small CTAs per SM
    // as with CTA pairs (rule 2 below was tuned that way); inside the step four to six GEMMs share the GPU, latency hiding comes
    // from the neighbours, and what counts is the L2 -> SM operand traffic per flop, which pairs halve for B: the whole train
    // step is 3.7 % (bf16) / 6.5 % (fp32-parity) faster with pairs everywhere (profiles/r1_ab_gemm_pairs.txt).
    static const int pairs = [] { const char* e = std::getenv("M2F_GEMM_CTA2"); return e ? std::atoi(e) : 1; }();
    // ... except with very many row tiles (M >= 65 536: the C5 inference batch has 1024): there every launch is many waves, and two
    // single CTAs per SM, one in its epilogue while the other is in its main loop, beat one pair per SM whose epilogue overlaps
    // nothing (C5 batch 43.5 -> 41.5 ms, tools/gpu_r2_c5_tiles.sh).
    // Launches of two and more 256 x 256 tiles per pair of SMs (and everything with >= 512 row tiles) go to the persistent pair kernel
    // (gemm_persist.cu: tile loop, two TMEM accumulators) when its epilogue applies, else to the plain pair kernel;
    // M2F_GEMM_PERSIST=0 restores the previous policy.
    static const int persist_mode = [] { const char* e = std::getenv("M2F_GEMM_PERSIST"); return e ? std::atoi(e) : 1; }();  // 2: tuning aid
    const bool persist_off = persist_mode == 0;
    const int64_t pair_tiles = ((M + 2 * BLOCK_M - 1) / (2 * BLOCK_M)) * ((N + 255) / 256);
    if (pairs == 1 && !forced && !persist_off && N >= 256 &&
        (mt >= 512 || pair_tiles >= 148 || (persist_mode == 2 && pair_tiles >= 2)))  // two tiles and more per pair of SMs
        return 256 | GEMM_TILE_CTA2 | GEMM_TILE_PERSIST;
    if (pairs == 1 && N > 64 && mt < 512) {
        if (forced == 192 && N > 128) return 192 | GEMM_TILE_CTA2;  // K-major operands only: m2f_gemm falls back to 256 otherwise
        const int bn = (forced == 128 || forced == 256) ? forced : 256;
        return ((N <= 128) ? 128 : bn) | GEMM_TILE_CTA2;
    }
    if (pairs == 2 && !forced && N >= 256) {
        // CTA pairs halve the B bytes each SM ingests: they win where the main loop dominates (long K) and, with
        // 256-wide tiles, once there are several waves of row tiles; short-K GEMMs keep two small CTAs per SM.
        if (mt >= 48) return 256 | GEMM_TILE_CTA2;
        if (K >= 1536) return 128 | GEMM_TILE_CTA2;
    }
    if (forced == 64 || forced == 128 || forced == 256) return (N <= 64) ? 64 : ((N <= 128 && forced == 256) ? 128 : forced);
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
    p.trace = g_trace;
    p.b_static = 0;
    p.epi_tma = 0;
    p.dbg = g_dbg;
    p.colsum = e->colsum;
    p.vec4 = (p.out_f32 == nullptr || (p.out_f32_ld & 3) == 0) && (p.resid == nullptr || (p.resid_ld & 3) == 0) && ((p.drop_ld & 3) == 0);
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

int gemm_launch(const CUtensorMap& ta, const CUtensorMap& tb, const EpiMaps& em, const GemmParams& p, int tile, int a_mn, int b_mn,
                cudaStream_t stream) {
    const int block_n = tile & 0xfff;
    const bool cta2 = (tile & GEMM_TILE_CTA2) != 0;
    if ((tile & GEMM_TILE_PERSIST) && gemm_persist_eligible(p, a_mn, b_mn)) return gemm_launch_persist(ta, tb, em, p, stream);
#define M2F_GEMM_CASE(BN, C2)                                                              \
    if (block_n == BN && cta2 == C2) {                                                     \
        if (!a_mn && !b_mn) return launch_gemm<BN, false, false, C2>(ta, tb, em, p, stream);   \
        if (!a_mn && b_mn) return launch_gemm<BN, false, true, C2>(ta, tb, em, p, stream);     \
        if (a_mn && b_mn) return launch_gemm<BN, true, true, C2>(ta, tb, em, p, stream);       \
        return launch_gemm<BN, true, false, C2>(ta, tb, em, p, stream);                        \
    }
    M2F_GEMM_CASE(64, false)
    M2F_GEMM_CASE(128, false)
    M2F_GEMM_CASE(256, false)
    M2F_GEMM_CASE(128, true)
    M2F_GEMM_CASE(256, true)
#undef M2F_GEMM_CASE
    // 256 x 192 pair tiles: K-major A; B K-major, or MN-major through 32-column atoms of the 64-byte swizzle (the caller builds
    // that tensor map with make_planes_tmap_sw64: 96 columns per CTA are 1.5 atoms of the 128-byte swizzle); MN-major A with both
    if (block_n == 192 && cta2 && !a_mn && !b_mn) return launch_gemm<192, false, false, true>(ta, tb, em, p, stream);
    if (block_n == 192 && cta2 && !a_mn && b_mn) return launch_gemm<192, false, true, true>(ta, tb, em, p, stream);
    if (block_n == 192 && cta2 && a_mn && b_mn) return launch_gemm<192, true, true, true>(ta, tb, em, p, stream);
    set_error("gemm: unsupported tile code %d", tile);
    return M2F_ERR_UNSUPPORTED;
}

}  // namespace m2f

// profiling aid (not part of the public header): 8 x uint64 globaltimer stamps per CTA are written to `buf`
extern "C" void m2f_debug_set_gemm_flags(int f) { m2f::g_dbg = f; }
extern "C" void m2f_debug_force_gemm_bn(int bn) { m2f::g_force_bn = bn; }  // -1: back to the policy
extern "C" void m2f_debug_set_gemm_trace(void* buf) { m2f::g_trace = reinterpret_cast<unsigned long long*>(buf); }

extern "C" int m2f_gemm(const m2f_planes_t* A, int a_mn_major, const m2f_planes_t* B, int b_mn_major,
                        const m2f_kseg_t* segs, int nsegs, int64_t M, int64_t N, const m2f_epilogue_t* epi,
                        void* stream) {
    using namespace m2f;
    GemmParams p;
    M2F_TRY(gemm_fill_params(p, A, a_mn_major, B, b_mn_major, segs, nsegs, M, N, epi));
    int64_t ktot = 0;
    for (int i = 0; i < nsegs; ++i) ktot += segs[i].klen;
    int bn = gemm_block_n(M, N, ktot);
    if ((bn & 0xfff) == 192 && a_mn_major && !b_mn_major) bn = 256 | GEMM_TILE_CTA2;
    CUtensorMap ta, tb;
    EpiMaps em;
    memset(&em, 0, sizeof(em));
    M2F_TRY(gemm_make_epi_maps(p, &em));
    M2F_TRY(make_planes_tmap(&ta, A, a_mn_major ? BLOCK_K : BLOCK_M));
    if (b_mn_major && (bn & 0xfff) == 192) M2F_TRY(make_planes_tmap_sw64(&tb, B, BLOCK_K));
    else M2F_TRY(make_planes_tmap(&tb, B, b_mn_major ? BLOCK_K : gemm_b_box_rows(bn)));
    return gemm_launch(ta, tb, em, p, bn, a_mn_major, b_mn_major, reinterpret_cast<cudaStream_t>(stream));
}
