// K1p: persistent variant of the tcgen05 GEMM for launches with very many row tiles (M >= 65 536: the inference batch of
// BASELINE.json's config 5, src/test.py:58-71).  Same product as gemm_sm100.cu -- Y = X W^T (both operands K-major), CTA pairs
// (cta_group::2: 256 x 256 tile per pair, each CTA stages its own 128 rows of A and half of the B tile) -- but a pair WALKS the
// tile list (tile = pair, pair + pairs, ...) with TWO accumulators in TMEM (2 x 256 columns): while the eight epilogue warps of
// each CTA drain accumulator a (bias / ReLU / residual, fp32 and one bf16 plane out through TMA tile stores), the MMA lane is
// already filling accumulator a ^ 1 from the operand ring, which the producer keeps full across tile boundaries.  The epilogue
// therefore has shared memory of its own (10 KB per warp) instead of borrowing the ring, and the ring is four stages of 32 KB.
// One launch = one wave of pairs: no prologue / tear-down / launch gap per tile, and the pair tile halves the L2 -> SM bytes
// per flop of the two single CTAs per SM that these shapes used before.
// Epilogue features: bias, ReLU, fp32 residual, fp32 output, 1 - 3 bf16 output planes as far as the 10 KB of staging per warp hold
// them (what inference launches, in every precision mode); anything else (dropout, gates, accumulation, column sums, MN-major
// operands) stays with gemm_sm100.cu.
#include <cstdlib>

#include "common.cuh"
#include "gemm.cuh"
#include "ptx.cuh"

namespace m2f {

namespace {

constexpr int PS_BN = 256;
constexpr int PS_STAGES = 4;
constexpr int PS_A_BYTES = BLOCK_M * BLOCK_K * 2;          // this CTA's 128 rows of A
constexpr int PS_B_BYTES = (PS_BN / 2) * BLOCK_K * 2;      // this CTA's half of the B tile
constexpr int PS_STAGE_BYTES = PS_A_BYTES + PS_B_BYTES;
constexpr int PS_EPI_WARPS = 8;
constexpr int PS_THREADS = 64 + 32 * PS_EPI_WARPS;
constexpr int PS_WARP_STG = 10240;                         // per epilogue warp: fp32 output tile 4 KB | bf16 plane tiles 2 KB each | residual tile 4 KB, whichever are present
constexpr int PS_OFF_STG = PS_STAGES * PS_STAGE_BYTES;
constexpr int PS_OFF_BIAS = PS_OFF_STG + PS_EPI_WARPS * PS_WARP_STG;  // 2 x 256 floats (one per accumulator)
constexpr int PS_OFF_BAR = PS_OFF_BIAS + 2 * PS_BN * 4;
constexpr int PS_NBAR = 2 * PS_STAGES + 4 + PS_EPI_WARPS;  // ring full / empty, accumulator full / empty x 2, residual loads
constexpr int PS_SMEM = PS_OFF_BAR + PS_NBAR * 8 + 16 + 1024;
static_assert(PS_SMEM <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ float4 ps_lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void ps_sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(PS_THREADS, 1)
gemm_persist_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const __grid_constant__ EpiMaps em,
                    const GemmParams p, const int tiles_n, const int total_tiles) {
    extern __shared__ uint8_t ps_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ps_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + PS_STAGES * PS_A_BYTES;
    float* s_bias = reinterpret_cast<float*>(smem + PS_OFF_BIAS);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + PS_OFF_BAR);
    uint64_t* empty_bar = full_bar + PS_STAGES;
    uint64_t* acc_full = empty_bar + PS_STAGES;   // [2]
    uint64_t* acc_empty = acc_full + 2;           // [2] (the leader's instance is the one in use)
    uint64_t* res_bar = acc_empty + 2;            // [PS_EPI_WARPS]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + PS_EPI_WARPS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta_rank = ptx::cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tma_a);
        ptx::prefetch_tensormap(&tma_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < PS_STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            ptx::mbar_init(&acc_full[a], 1);
            ptx::mbar_init(&acc_empty[a], 2 * PS_EPI_WARPS);  // one arrival per epilogue warp of both CTAs
        }
        for (int s = 0; s < PS_EPI_WARPS; ++s) ptx::mbar_init(&res_bar[s], 1);
        ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc_2cta(tmem_slot, 2 * PS_BN);
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_trigger();

    if (warp == 0) {
        // ---------------------------------------------------------------------- TMA producer (one lane in each CTA)
        if (ptx::elect_one()) {
            int it = 0;
            for (int t = pair; t < total_tiles; t += npairs) {
                const int m_blk = (t / tiles_n) * 2 + (int)cta_rank, n_blk = t % tiles_n;
                const int nb0 = n_blk * PS_BN + (int)cta_rank * (PS_BN / 2);
                for (int pr = 0; pr < p.nprod; ++pr) {
                    const int pa = p.prod_a[pr], pb = p.prod_b[pr];
                    for (int sg = 0; sg < p.nsegs; ++sg) {
                        for (int kb = 0; kb < p.seg_nkb[sg]; ++kb, ++it) {
                            const int stage = it % PS_STAGES;
                            const uint32_t phase = (uint32_t)(it / PS_STAGES) & 1u;
                            ptx::mbar_wait(&empty_bar[stage], phase ^ 1u);
                            if (cta_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2 * PS_STAGE_BYTES);
                            const uint32_t fb = ptx::mapa(ptx::smem_u32(&full_bar[stage]), 0);  // both CTAs credit the leader's barrier
                            ptx::tma_load_3d_2cta(smem_b + stage * PS_B_BYTES, &tma_b, fb, p.seg_b_k0[sg] + kb * BLOCK_K, nb0, pb);
                            ptx::tma_load_3d_2cta(smem_a + stage * PS_A_BYTES, &tma_a, fb, p.seg_a_k0[sg] + kb * BLOCK_K, m_blk * BLOCK_M, pa);
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---------------------------------------------------------------------- MMA issuer (one lane of the leader CTA)
        if (cta_rank == 0 && ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::instr_desc_bf16(2 * BLOCK_M, PS_BN, false, false);
            int it = 0, i = 0;
            for (int t = pair; t < total_tiles; t += npairs, ++i) {
                const int acc = i & 1;
                if (i >= 2) ptx::mbar_wait(&acc_empty[acc], (uint32_t)((i >> 1) + 1) & 1u);  // both CTAs have drained this accumulator
                ptx::tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * PS_BN);
                for (int kbi = 0; kbi < p.total_kb; ++kbi, ++it) {
                    const int stage = it % PS_STAGES;
                    const uint32_t phase = (uint32_t)(it / PS_STAGES) & 1u;
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tcgen05_fence_after();
                    const uint64_t ad = ptx::smem_desc_sw128(ptx::smem_u32(smem_a + stage * PS_A_BYTES), 0, 1024);
                    const uint64_t bd = ptx::smem_desc_sw128(ptx::smem_u32(smem_b + stage * PS_B_BYTES), 0, 1024);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k)  // 16 columns = 32 bytes inside the swizzle atom: address field + 2
                        ptx::umma_bf16_2cta(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kbi > 0 || k > 0) ? 1u : 0u);
                    ptx::umma_commit_2cta(&empty_bar[stage], 3);  // frees the slot in both CTAs once these MMAs retire
                }
                ptx::umma_commit_2cta(&acc_full[acc], 3);
            }
        }
        __syncwarp();
    } else {
        // ---------------------------------------------------------------------- epilogue warps
        const int ew = warp - 2, lg = warp & 3;
        const bool f_bias = p.bias != nullptr, f_relu = p.relu != 0, f_res = p.resid != nullptr, f_of32 = p.out_f32 != nullptr,
                   f_opl = p.out_pl != nullptr;
        const int n_pl = f_opl ? p.out_pl_n : 0;
        const uint32_t s_of = ptx::smem_u32(smem + PS_OFF_STG) + (uint32_t)(ew * PS_WARP_STG);
        const uint32_t s_pl = s_of + (f_of32 ? 4096u : 0u), s_res = s_pl + (uint32_t)n_pl * 2048u;
        uint64_t* rbar = &res_bar[ew];
        const uint32_t sw128 = (uint32_t)(lane & 7), sw64 = (uint32_t)((lane >> 1) & 3);
        const uint32_t acc_empty_leader = ptx::mapa(ptx::smem_u32(&acc_empty[0]), 0);
        uint32_t rphase = 0;
        int i = 0;
        for (int t = pair; t < total_tiles; t += npairs, ++i) {
            const int acc = i & 1;
            const int m_blk = (t / tiles_n) * 2 + (int)cta_rank, n_blk = t % tiles_n;
            const int m0 = m_blk * BLOCK_M + lg * 32, nt0 = n_blk * PS_BN;
            const uint32_t sb = ptx::smem_u32(s_bias + acc * PS_BN);
            if (f_bias) {  // this tile's bias values -> shared (one copy per accumulator: a slow warp may still read the other)
                for (int j = ew * 32 + lane; j < PS_BN; j += PS_EPI_WARPS * 32) {
                    const int n = nt0 + j;
                    const float b = n < p.N ? p.bias[n] : 0.f;
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(sb + (uint32_t)j * 4u), "f"(b) : "memory");
                }
                ptx::named_bar_sync(1, PS_EPI_WARPS * 32);
            }
            ptx::mbar_wait(&acc_full[acc], (uint32_t)(i >> 1) & 1u);
            ptx::tcgen05_fence_after();
            if (m0 < p.M) {
                const uint32_t t_lane = tmem_base + (uint32_t)(acc * PS_BN) + ((uint32_t)(lg * 32) << 16);
#pragma unroll 1
                for (int c = ew >> 2; c < PS_BN / 32; c += PS_EPI_WARPS / 4) {
                    const int n0 = nt0 + c * 32;
                    if (n0 >= p.N) break;
                    if (f_res && lane == 0) {
                        ptx::mbar_expect_tx(rbar, 4096u);
                        ptx::tma_load_2d(s_res, &em.resid, rbar, n0, m0);
                    }
                    float x[32];
                    {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32(t_lane + (uint32_t)(c * 32), r);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(r[j]);
                    }
                    if (f_bias) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 b = ps_lds128(sb + (uint32_t)(c * 32 + j * 4) * 4u);
                            x[4 * j] += b.x; x[4 * j + 1] += b.y; x[4 * j + 2] += b.z; x[4 * j + 3] += b.w;
                        }
                    }
                    if (f_relu) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
                    }
                    if (f_res) {
                        ptx::mbar_wait(rbar, rphase);
                        rphase ^= 1u;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 q = ps_lds128(s_res + (uint32_t)lane * 128u + (((uint32_t)j ^ sw128) << 4));
                            x[4 * j] += q.x; x[4 * j + 1] += q.y; x[4 * j + 2] += q.z; x[4 * j + 3] += q.w;
                        }
                    }
                    if (f_of32) {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            ps_sts128(s_of + (uint32_t)lane * 128u + (((uint32_t)j ^ sw128) << 4), __float_as_uint(x[4 * j]), __float_as_uint(x[4 * j + 1]),
                                      __float_as_uint(x[4 * j + 2]), __float_as_uint(x[4 * j + 3]));
                    }
                    for (int pl = 0; pl < n_pl; ++pl) {  // hi plane, then the rounded residuals
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
                            ps_sts128(s_pl + (uint32_t)pl * 2048u + (uint32_t)lane * 64u + (((uint32_t)j ^ sw64) << 4), h[4 * j], h[4 * j + 1],
                                      h[4 * j + 2], h[4 * j + 3]);
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        if (f_of32) ptx::tma_store_2d(&em.out_f32, s_of, n0, m0);
                        for (int pl = 0; pl < n_pl; ++pl) ptx::tma_store_3d(&em.out_pl, s_pl + (uint32_t)pl * 2048u, n0, m0, pl);
                        ptx::bulk_commit();
                        ptx::bulk_wait_read0();  // the tiles are re-used by the next chunk (single staging set)
                    }
                    __syncwarp();
                }
            }
            ptx::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(acc_empty_leader + (uint32_t)(acc * 8));  // this warp has drained accumulator `acc`
        }
    }
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();  // neither CTA may release its TMEM / exit while the pair is in flight
    if (warp == 2) {
        ptx::tcgen05_fence_after();
        ptx::tmem_dealloc_2cta(tmem_base, 2 * PS_BN);
    }
}

}  // namespace

// M2F_GEMM_PERSIST=0 keeps these shapes on gemm_sm100.cu (A/B, debugging)
bool gemm_persist_eligible(const GemmParams& p, int a_mn, int b_mn) {
    static const bool off = [] { const char* e = std::getenv("M2F_GEMM_PERSIST"); return e && e[0] == '0'; }();
    const int stg = (p.out_f32 ? 4096 : 0) + (p.out_pl ? p.out_pl_n * 2048 : 0) + (p.resid ? 4096 : 0);  // per epilogue warp
    return !off && !a_mn && !b_mn && p.epi_tma && !p.gate && !p.drop.enabled && !p.accumulate && !p.colsum && stg <= PS_WARP_STG &&
           (p.out_f32 || p.out_pl);
}

int gemm_launch_persist(const CUtensorMap& ta, const CUtensorMap& tb, const EpiMaps& em, const GemmParams& p, cudaStream_t stream) {
    static bool attr_set = false;
    static int sms = 0;
    if (!attr_set) {
        M2F_CUDA(cudaFuncSetAttribute(gemm_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
        int dev = 0;
        M2F_CUDA(cudaGetDevice(&dev));
        M2F_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        attr_set = true;
    }
    const int tiles_n = ceil_div(p.N, PS_BN), tiles_m = ceil_div(p.M, 2 * BLOCK_M);
    const int total = tiles_n * tiles_m;
    const int pairs = total < sms / 2 ? total : sms / 2;
    M2F_CUDA(launch_chain(gemm_persist_kernel, dim3(2 * pairs), dim3(PS_THREADS), (size_t)PS_SMEM, stream, 2, ta, tb, em, p, tiles_n, total));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

}  // namespace m2f
