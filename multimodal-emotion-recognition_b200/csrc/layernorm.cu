// K3: LayerNorm forward / backward with fused residual (outer skip), dropout-mask regeneration,
// bf16-plane operand copies and bias / gamma / beta gradient column sums.
// Reference arithmetic: nn.LayerNorm(eps=1e-5, biased variance) as used by
// nn.TransformerEncoderLayer norm1/norm2 (torch/nn/modules/transformer.py:977-982), the shared final norm
// of nn.TransformerEncoder (:550-551) and the outer skip of src/model.py:107,119.
// One warp per row; a row of D <= 1024 fp32 lives in registers (NCH float4 chunks per lane).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"

namespace m2f {

constexpr float LN_EPS = 1e-5f;
constexpr int LN_THREADS = 128;   // default block of the register-resident forward (M2F_LN_FWD_THREADS = 64 / 128 / 256: tuning aid)
constexpr int LN_WARPS = LN_THREADS / 32;
constexpr int LN_MAX_THREADS = 256;

template <int NCH>
__global__ void __launch_bounds__(LN_MAX_THREADS)
ln_fwd_kernel(const float* __restrict__ x, int64_t x_ld, const float* __restrict__ gamma, const float* __restrict__ beta,
              const float* __restrict__ skip, int64_t skip_ld, int64_t M, int D, float* __restrict__ y, int64_t y_ld,
              __nv_bfloat16* __restrict__ ypl, int64_t ypl_ld, int64_t ypl_stride, int P, Drop drop, float* __restrict__ stats) {
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    const int lane = threadIdx.x & 31;
    const int wpb = (int)(blockDim.x >> 5);
    const int64_t warp0 = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * wpb;
    const float invD = 1.f / (float)D;
    for (int64_t m = warp0; m < M; m += nwarps) {
        float4 v[NCH];
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int col = (c * 32 + lane) * 4;
            v[c] = col < D ? *reinterpret_cast<const float4*>(x + m * x_ld + col) : make_float4(0.f, 0.f, 0.f, 0.f);
            s += v[c].x + v[c].y + v[c].z + v[c].w;
        }
        const float mean = warp_sum(s) * invD;
        float q = 0.f;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int col = (c * 32 + lane) * 4;
            if (col < D) {
                const float a = v[c].x - mean, b = v[c].y - mean, cc = v[c].z - mean, d = v[c].w - mean;
                q += a * a + b * b + cc * cc + d * d;
            }
        }
        const float rstd = rsqrtf(warp_sum(q) * invD + LN_EPS);
        if (lane == 0 && stats) {
            stats[2 * m] = mean;
            stats[2 * m + 1] = rstd;
        }
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int col = (c * 32 + lane) * 4;
            if (col < D) {
                const float4 g = *reinterpret_cast<const float4*>(gamma + col);
                const float4 b = *reinterpret_cast<const float4*>(beta + col);
                float o[4] = {(v[c].x - mean) * rstd * g.x + b.x, (v[c].y - mean) * rstd * g.y + b.y,
                              (v[c].z - mean) * rstd * g.z + b.z, (v[c].w - mean) * rstd * g.w + b.w};
                if (skip) {
                    const float4 k = *reinterpret_cast<const float4*>(skip + m * skip_ld + col);
                    o[0] += k.x; o[1] += k.y; o[2] += k.z; o[3] += k.w;
                }
                if (y) *reinterpret_cast<float4*>(y + m * y_ld + col) = make_float4(o[0], o[1], o[2], o[3]);
                if (ypl) {
                    if (drop.enabled) {
                        float dm[4];
                        drop_mult4(drop, (uint64_t)(m * D + col) >> 2, dm);
#pragma unroll
                        for (int i = 0; i < 4; ++i) o[i] *= dm[i];
                    }
                    store_planes4(ypl, ypl_stride, P, m * ypl_ld + col, o);
                }
            }
        }
    }
}

// Backward.  A block owns groups of 8 rows.  Phase 1, one row per warp: dx (fp32 and / or dropout-masked planes); what the
// column sums need of the row -- g * xhat (dgamma), g (dbeta), the masked dx (dbias) -- is parked in shared memory as it is
// produced (3 x 8 rows x D floats).  Phase 2, one float4 column per thread: the block adds up its 8 parked rows in 12 registers;
// the sums leave the block as vector atomics (red.global.add.v4.f32) once, at the end.  (Until late in round 2 phase 2 re-read the
// rows from global memory and recomputed dx, Philox included: 11 -> 8 us per launch at the training size.)
constexpr int LNB_ROWS = 8;
constexpr int LNB_THREADS = LNB_ROWS * 32;

template <int NCH>
__global__ void __launch_bounds__(LNB_THREADS)
ln_bwd_kernel(const float* __restrict__ g, int64_t g_ld, const float* __restrict__ x, int64_t x_ld,
              const float* __restrict__ gamma, const float* __restrict__ stats, int64_t M, int D, float* __restrict__ dx,
              int64_t dx_ld, __nv_bfloat16* __restrict__ dxpl, int64_t dxpl_ld, int64_t dxpl_stride, int P, Drop drop,
              float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias) {
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    extern __shared__ __align__(16) float lnb_park[];  // [3][LNB_ROWS][D]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float invD = 1.f / (float)D;
    const int ccol = threadIdx.x * 4;  // phase-2 column of this thread (D <= 1024 = 256 threads x 4)
    float* park_gx = lnb_park + (size_t)warp * D;
    float* park_g = lnb_park + (size_t)(LNB_ROWS + warp) * D;
    float* park_m = lnb_park + (size_t)(2 * LNB_ROWS + warp) * D;
    float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sb = sg, sbias = sg;
    for (int64_t m0 = (int64_t)blockIdx.x * LNB_ROWS; m0 < M; m0 += (int64_t)gridDim.x * LNB_ROWS) {
        const int64_t m = m0 + warp;
        if (m < M) {
            float4 xv[NCH], av[NCH];
#pragma unroll
            for (int c = 0; c < NCH; ++c) {  // all loads of the row first
                const int col = (c * 32 + lane) * 4;
                if (col < D) {
                    xv[c] = *reinterpret_cast<const float4*>(x + m * x_ld + col);
                    av[c] = *reinterpret_cast<const float4*>(g + m * g_ld + col);
                } else {
                    xv[c] = av[c] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            const float mean = stats[2 * m], rstd = stats[2 * m + 1];
            float s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const int col = (c * 32 + lane) * 4;
                if (col < D) {
                    const float4 gm = *reinterpret_cast<const float4*>(gamma + col);
                    xv[c] = make_float4((xv[c].x - mean) * rstd, (xv[c].y - mean) * rstd, (xv[c].z - mean) * rstd, (xv[c].w - mean) * rstd);
                    *reinterpret_cast<float4*>(park_gx + col) = make_float4(av[c].x * xv[c].x, av[c].y * xv[c].y, av[c].z * xv[c].z, av[c].w * xv[c].w);
                    *reinterpret_cast<float4*>(park_g + col) = av[c];
                    av[c] = make_float4(av[c].x * gm.x, av[c].y * gm.y, av[c].z * gm.z, av[c].w * gm.w);
                    s1 += av[c].x + av[c].y + av[c].z + av[c].w;
                    s2 += av[c].x * xv[c].x + av[c].y * xv[c].y + av[c].z * xv[c].z + av[c].w * xv[c].w;
                }
            }
            const float c1 = warp_sum(s1) * invD, c2 = warp_sum(s2) * invD;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const int col = (c * 32 + lane) * 4;
                if (col < D) {
                    float o[4] = {rstd * (av[c].x - c1 - xv[c].x * c2), rstd * (av[c].y - c1 - xv[c].y * c2),
                                  rstd * (av[c].z - c1 - xv[c].z * c2), rstd * (av[c].w - c1 - xv[c].w * c2)};
                    if (dx) *reinterpret_cast<float4*>(dx + m * dx_ld + col) = make_float4(o[0], o[1], o[2], o[3]);
                    if (dxpl || dbias) {
                        if (drop.enabled) {
                            float dm[4];
                            drop_mult4(drop, (uint64_t)(m * D + col) >> 2, dm);
                            o[0] *= dm[0]; o[1] *= dm[1]; o[2] *= dm[2]; o[3] *= dm[3];
                        }
                        if (dxpl) store_planes4(dxpl, dxpl_stride, P, m * dxpl_ld + col, o);
                        if (dbias) *reinterpret_cast<float4*>(park_m + col) = make_float4(o[0], o[1], o[2], o[3]);
                    }
                }
            }
        }
        __syncthreads();
        if (ccol < D) {
            const int nrow = (int)min((int64_t)LNB_ROWS, M - m0);
            for (int r = 0; r < nrow; ++r) {
                const float4 a = *reinterpret_cast<const float4*>(lnb_park + (size_t)r * D + ccol);
                const float4 b = *reinterpret_cast<const float4*>(lnb_park + (size_t)(LNB_ROWS + r) * D + ccol);
                sg.x += a.x; sg.y += a.y; sg.z += a.z; sg.w += a.w;
                sb.x += b.x; sb.y += b.y; sb.z += b.z; sb.w += b.w;
                if (dbias) {
                    const float4 cm = *reinterpret_cast<const float4*>(lnb_park + (size_t)(2 * LNB_ROWS + r) * D + ccol);
                    sbias.x += cm.x; sbias.y += cm.y; sbias.z += cm.z; sbias.w += cm.w;
                }
            }
        }
        __syncthreads();
    }
    if (ccol < D) {
        if (dgamma) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dgamma + ccol), "f"(sg.x), "f"(sg.y), "f"(sg.z), "f"(sg.w) : "memory");
        if (dbeta) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbeta + ccol), "f"(sb.x), "f"(sb.y), "f"(sb.z), "f"(sb.w) : "memory");
        if (dbias) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbias + ccol), "f"(sbias.x), "f"(sbias.y), "f"(sbias.z), "f"(sbias.w) : "memory");
    }
}

// ------------------------------------------------------------------------------------------ forward, staged through shared memory
// Large-M variant (selected by ln_v2_level below): the register-resident kernel above keeps 16 - 20 warps per SM with one 3 KB
// row in flight each and reaches 0.48 of the HBM peak at M = 131 072.  Here a block moves tiles of 8 rows
// of x (and skip) into a 2-deep shared-memory ring with bulk asynchronous copies issued by one thread (cp.async.bulk,
// completion on an mbarrier): 48 KB per stage and 96 KB per block are in flight without costing registers, and the
// arithmetic of tile t overlaps the landing of tile t + 1.  Same arithmetic and the same Philox element indices as above.
constexpr int LN2_ROWS = 8;
constexpr int LN2_THREADS = LN2_ROWS * 32;
constexpr int LN2_STAGES = 2;

__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ptx::smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(ptx::smem_u32(bar))
                 : "memory");
}

template <int NCH>
__global__ void __launch_bounds__(LN2_THREADS)
ln_fwd_staged_kernel(const float* __restrict__ x, int64_t x_ld, const float* __restrict__ gamma, const float* __restrict__ beta,
                     const float* __restrict__ skip, int64_t skip_ld, int64_t M, int D, float* __restrict__ y, int64_t y_ld,
                     __nv_bfloat16* __restrict__ ypl, int64_t ypl_ld, int64_t ypl_stride, int P, Drop drop, float* __restrict__ stats) {
    extern __shared__ __align__(128) uint8_t ln2_smem[];
    __shared__ __align__(8) uint64_t full_bar[LN2_STAGES];
    float* ring = reinterpret_cast<float*>(ln2_smem);
    const bool has_skip = skip != nullptr;
    const int tile_floats = LN2_ROWS * D;
    const int stage_floats = tile_floats * (has_skip ? 2 : 1);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float invD = 1.f / (float)D;
    const int64_t ntiles = (M + LN2_ROWS - 1) / LN2_ROWS;
    if (threadIdx.x == 0) {
        for (int s2 = 0; s2 < LN2_STAGES; ++s2) ptx::mbar_init(&full_bar[s2], 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    // thread 0: request one tile into stage s2 (rows past M are not requested; their warps skip the arithmetic)
    auto issue = [&](int64_t tile, int s2) {
        const int64_t m0 = tile * LN2_ROWS;
        const int rows = (int)((M - m0) < LN2_ROWS ? (M - m0) : LN2_ROWS);
        const uint32_t row_bytes = (uint32_t)D * 4u;
        ptx::mbar_expect_tx(&full_bar[s2], (uint32_t)rows * row_bytes * (has_skip ? 2u : 1u));
        float* dx = ring + (size_t)s2 * stage_floats;
        if (x_ld == D) {
            bulk_g2s(dx, x + m0 * x_ld, (uint32_t)rows * row_bytes, &full_bar[s2]);
        } else {
            for (int r = 0; r < rows; ++r) bulk_g2s(dx + (size_t)r * D, x + (m0 + r) * x_ld, row_bytes, &full_bar[s2]);
        }
        if (has_skip) {
            float* ds = dx + tile_floats;
            if (skip_ld == D) {
                bulk_g2s(ds, skip + m0 * skip_ld, (uint32_t)rows * row_bytes, &full_bar[s2]);
            } else {
                for (int r = 0; r < rows; ++r) bulk_g2s(ds + (size_t)r * D, skip + (m0 + r) * skip_ld, row_bytes, &full_bar[s2]);
            }
        }
    };
    if (threadIdx.x == 0) {
        for (int s2 = 0; s2 < LN2_STAGES; ++s2) {
            const int64_t t = (int64_t)blockIdx.x + (int64_t)s2 * gridDim.x;
            if (t < ntiles) issue(t, s2);
        }
    }
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int s2 = it % LN2_STAGES;
        ptx::mbar_wait(&full_bar[s2], (uint32_t)((it / LN2_STAGES) & 1));
        const int64_t m = tile * LN2_ROWS + warp;
        if (m < M) {
            const float* xs = ring + (size_t)s2 * stage_floats + (size_t)warp * D;
            const float* ks = xs + tile_floats;
            float4 v[NCH];
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const int col = (c * 32 + lane) * 4;
                v[c] = col < D ? *reinterpret_cast<const float4*>(xs + col) : make_float4(0.f, 0.f, 0.f, 0.f);
                s += v[c].x + v[c].y + v[c].z + v[c].w;
            }
            const float mean = warp_sum(s) * invD;
            float q = 0.f;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const int col = (c * 32 + lane) * 4;
                if (col < D) {
                    const float a = v[c].x - mean, b = v[c].y - mean, cc = v[c].z - mean, d = v[c].w - mean;
                    q += a * a + b * b + cc * cc + d * d;
                }
            }
            const float rstd = rsqrtf(warp_sum(q) * invD + LN_EPS);
            if (lane == 0 && stats) {
                stats[2 * m] = mean;
                stats[2 * m + 1] = rstd;
            }
            // keep/drop multipliers of the row, before the column loop: with (m * D) % 8 == 0 the 4-groups of lanes 2k, 2k+1 are
            // the two halves of one Philox call, so a lane pair shares the calls of two chunks (drop_mult4_pair)
            float dmul[NCH][4];
            if (ypl && drop.enabled) {
                const uint64_t row4 = (uint64_t)(m * D) >> 2;
                const bool pairs = ((m * D) & 7) == 0;
#pragma unroll
                for (int c = 0; c < NCH; c += 2) {
                    if (pairs && c + 1 < NCH) drop_mult4_pair(drop, row4 + c * 32 + lane, row4 + (c + 1) * 32 + lane, lane, dmul[c], dmul[c + 1 < NCH ? c + 1 : c]);
                    else {
                        drop_mult4(drop, row4 + c * 32 + lane, dmul[c]);
                        if (c + 1 < NCH) drop_mult4(drop, row4 + (c + 1) * 32 + lane, dmul[c + 1 < NCH ? c + 1 : c]);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const int col = (c * 32 + lane) * 4;
                if (col < D) {
                    const float4 g = *reinterpret_cast<const float4*>(gamma + col);
                    const float4 b = *reinterpret_cast<const float4*>(beta + col);
                    float o[4] = {(v[c].x - mean) * rstd * g.x + b.x, (v[c].y - mean) * rstd * g.y + b.y,
                                  (v[c].z - mean) * rstd * g.z + b.z, (v[c].w - mean) * rstd * g.w + b.w};
                    if (has_skip) {
                        const float4 k = *reinterpret_cast<const float4*>(ks + col);
                        o[0] += k.x; o[1] += k.y; o[2] += k.z; o[3] += k.w;
                    }
                    if (y) *reinterpret_cast<float4*>(y + m * y_ld + col) = make_float4(o[0], o[1], o[2], o[3]);
                    if (ypl) {
                        if (drop.enabled) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) o[i] *= dmul[c][i];
                        }
                        store_planes4(ypl, ypl_stride, P, m * ypl_ld + col, o);
                    }
                }
            }
        }
        __syncthreads();  // every warp has read stage s2: it can be refilled
        const int64_t nxt = tile + (int64_t)LN2_STAGES * gridDim.x;
        if (threadIdx.x == 0 && nxt < ntiles) issue(nxt, s2);
    }
}

// Backward, same staging: tiles of 8 rows of g and x land in shared memory.  Phase 1 (one row per warp) only reduces: the
// row's c1 = mean(g*gamma), c2 = mean(g*gamma*xhat).  Phase 2 (one float4 column per thread, 8 rows) does everything else
// exactly once per element: dx, the Philox mask, the fp32 and plane stores (a warp writes 512 contiguous bytes of a row) and the
// dgamma / dbeta / dbias partial sums -- the first version of this kernel computed dx and the mask in both phases and ran at
// 0.46 of the HBM peak, bound by the duplicated Philox rounds rather than by memory.  Same arithmetic and Philox element
// indices as ln_bwd_kernel; every row is fetched from global memory once.
template <int NCH>
__global__ void __launch_bounds__(LN2_THREADS)
ln_bwd_staged_kernel(const float* __restrict__ g, int64_t g_ld, const float* __restrict__ x, int64_t x_ld,
                     const float* __restrict__ gamma, const float* __restrict__ stats, int64_t M, int D, float* __restrict__ dx,
                     int64_t dx_ld, __nv_bfloat16* __restrict__ dxpl, int64_t dxpl_ld, int64_t dxpl_stride, int P, Drop drop,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias) {
    extern __shared__ __align__(128) uint8_t ln2_smem[];
    __shared__ __align__(8) uint64_t full_bar[LN2_STAGES];
    __shared__ __align__(16) float rowc[LN2_ROWS][4];  // mean, rstd, c1, c2
    float* ring = reinterpret_cast<float*>(ln2_smem);
    const int tile_floats = LN2_ROWS * D;
    const int stage_floats = 2 * tile_floats;  // x tile, then g tile
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float invD = 1.f / (float)D;
    const int ccol = threadIdx.x * 4;
    const int64_t ntiles = (M + LN2_ROWS - 1) / LN2_ROWS;
    if (threadIdx.x == 0) {
        for (int s2 = 0; s2 < LN2_STAGES; ++s2) ptx::mbar_init(&full_bar[s2], 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    auto issue = [&](int64_t tile, int s2) {
        const int64_t m0 = tile * LN2_ROWS;
        const int rows = (int)((M - m0) < LN2_ROWS ? (M - m0) : LN2_ROWS);
        const uint32_t row_bytes = (uint32_t)D * 4u;
        ptx::mbar_expect_tx(&full_bar[s2], (uint32_t)rows * row_bytes * 2u);
        float* sx = ring + (size_t)s2 * stage_floats;
        float* sgr = sx + tile_floats;
        if (x_ld == D) {
            bulk_g2s(sx, x + m0 * x_ld, (uint32_t)rows * row_bytes, &full_bar[s2]);
        } else {
            for (int r = 0; r < rows; ++r) bulk_g2s(sx + (size_t)r * D, x + (m0 + r) * x_ld, row_bytes, &full_bar[s2]);
        }
        if (g_ld == D) {
            bulk_g2s(sgr, g + m0 * g_ld, (uint32_t)rows * row_bytes, &full_bar[s2]);
        } else {
            for (int r = 0; r < rows; ++r) bulk_g2s(sgr + (size_t)r * D, g + (m0 + r) * g_ld, row_bytes, &full_bar[s2]);
        }
    };
    if (threadIdx.x == 0) {
        for (int s2 = 0; s2 < LN2_STAGES; ++s2) {
            const int64_t t = (int64_t)blockIdx.x + (int64_t)s2 * gridDim.x;
            if (t < ntiles) issue(t, s2);
        }
    }
    float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sb = sg, sbias = sg, gmc = sg;
    if (ccol < D) gmc = *reinterpret_cast<const float4*>(gamma + ccol);
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int s2 = it % LN2_STAGES;
        const int64_t m0 = tile * LN2_ROWS;
        const int64_t m = m0 + warp;
        float mean = 0.f, rstd = 0.f;
        if (m < M) {  // the row statistics do not come through the ring: fetch them while the tile lands
            mean = stats[2 * m];
            rstd = stats[2 * m + 1];
        }
        ptx::mbar_wait(&full_bar[s2], (uint32_t)((it / LN2_STAGES) & 1));
        const float* sx = ring + (size_t)s2 * stage_floats;
        const float* sgr = sx + tile_floats;
        if (m < M) {  // phase 1: the two row means
            float s1 = 0.f, s2sum = 0.f;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const int col = (c * 32 + lane) * 4;
                if (col < D) {
                    const float4 gm = *reinterpret_cast<const float4*>(gamma + col);
                    const float4 xr = *reinterpret_cast<const float4*>(sx + (size_t)warp * D + col);
                    const float4 gr = *reinterpret_cast<const float4*>(sgr + (size_t)warp * D + col);
                    const float4 xh = make_float4((xr.x - mean) * rstd, (xr.y - mean) * rstd, (xr.z - mean) * rstd, (xr.w - mean) * rstd);
                    const float4 a = make_float4(gr.x * gm.x, gr.y * gm.y, gr.z * gm.z, gr.w * gm.w);
                    s1 += a.x + a.y + a.z + a.w;
                    s2sum += a.x * xh.x + a.y * xh.y + a.z * xh.z + a.w * xh.w;
                }
            }
            const float c1 = warp_sum(s1) * invD, c2 = warp_sum(s2sum) * invD;
            if (lane == 0) { rowc[warp][0] = mean; rowc[warp][1] = rstd; rowc[warp][2] = c1; rowc[warp][3] = c2; }
        }
        __syncthreads();  // rowc complete
        {   // phase 2: dx, mask, stores and the column sums, once per element.  Rows go in pairs: with D % 8 == 0 the 4-groups
            // of threads 2k, 2k+1 in a row are the two halves of one Philox call, so a thread pair shares the calls of two rows.
            const bool live = ccol < D;
            const int nrow = (int)min((int64_t)LN2_ROWS, M - m0);
            const bool masked = drop.enabled && (dxpl || dbias);
            const bool pairs = masked && (D & 7) == 0;
            auto element = [&](int r, const float dm[4]) {
                const int64_t mm = m0 + r;
                const float4 xr = *reinterpret_cast<const float4*>(sx + (size_t)r * D + ccol);
                const float4 gr = *reinterpret_cast<const float4*>(sgr + (size_t)r * D + ccol);
                const float4 rc = *reinterpret_cast<const float4*>(rowc[r]);  // mean, rstd, c1, c2
                const float4 xh = make_float4((xr.x - rc.x) * rc.y, (xr.y - rc.x) * rc.y, (xr.z - rc.x) * rc.y, (xr.w - rc.x) * rc.y);
                sg.x = fmaf(gr.x, xh.x, sg.x); sg.y = fmaf(gr.y, xh.y, sg.y); sg.z = fmaf(gr.z, xh.z, sg.z); sg.w = fmaf(gr.w, xh.w, sg.w);
                sb.x += gr.x; sb.y += gr.y; sb.z += gr.z; sb.w += gr.w;
                float o[4] = {rc.y * (gr.x * gmc.x - rc.z - xh.x * rc.w), rc.y * (gr.y * gmc.y - rc.z - xh.y * rc.w),
                              rc.y * (gr.z * gmc.z - rc.z - xh.z * rc.w), rc.y * (gr.w * gmc.w - rc.z - xh.w * rc.w)};
                if (dx) *reinterpret_cast<float4*>(dx + mm * dx_ld + ccol) = make_float4(o[0], o[1], o[2], o[3]);
                if (dxpl || dbias) {
                    if (masked) { o[0] *= dm[0]; o[1] *= dm[1]; o[2] *= dm[2]; o[3] *= dm[3]; }
                    if (dxpl) store_planes4(dxpl, dxpl_stride, P, mm * dxpl_ld + ccol, o);
                    sbias.x += o[0]; sbias.y += o[1]; sbias.z += o[2]; sbias.w += o[3];
                }
            };
            const bool warp_live = warp * 128 < D;  // a whole warp past D (D = 768: warps 6, 7) has nothing to do in this phase
            int r = warp_live ? 0 : nrow;
            for (; r + 1 < nrow; r += 2) {
                float da[4] = {1.f, 1.f, 1.f, 1.f}, db[4] = {1.f, 1.f, 1.f, 1.f};
                const uint64_t ia = ((uint64_t)((m0 + r) * D) >> 2) + threadIdx.x, ib = ia + ((uint64_t)D >> 2);
                if (pairs) drop_mult4_pair(drop, ia, ib, lane, da, db);  // whole warps: also the lanes past D
                else if (masked && live) { drop_mult4(drop, ia, da); drop_mult4(drop, ib, db); }
                if (live) { element(r, da); element(r + 1, db); }
            }
            if (r < nrow && live) {
                float da[4] = {1.f, 1.f, 1.f, 1.f};
                if (masked) drop_mult4(drop, ((uint64_t)((m0 + r) * D) >> 2) + threadIdx.x, da);
                element(r, da);
            }
        }
        __syncthreads();  // every thread has read stage s2 and rowc: both can be overwritten
        const int64_t nxt = tile + (int64_t)LN2_STAGES * gridDim.x;
        if (threadIdx.x == 0 && nxt < ntiles) issue(nxt, s2);
    }
    if (ccol < D) {
        if (dgamma) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dgamma + ccol), "f"(sg.x), "f"(sg.y), "f"(sg.z), "f"(sg.w) : "memory");
        if (dbeta) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbeta + ccol), "f"(sb.x), "f"(sb.y), "f"(sb.z), "f"(sb.w) : "memory");
        if (dbias) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbias + ccol), "f"(sbias.x), "f"(sbias.y), "f"(sbias.z), "f"(sbias.w) : "memory");
    }
}

// Which kernel family: the register-resident kernels keep a row per warp in registers -- best while the tensors sit in
// L2 (training sizes: 9 - 13 us per launch, and their 0 bytes of dynamic shared memory let them share an SM with a GEMM CTA
// pair) -- but only 16 - 20 warps fit an SM, too few loads in flight for HBM: 0.48 / 0.44 of the HBM peak at M = 131 072.
// The staged kernels (bulk asynchronous copies into a 2-deep ring, 96 KB in flight per block) reach 0.74 there.  Rule: staged
// once one fp32 [M, D] operand is 32 MB or more (the kernel's three to four operands then exceed the 126 MB L2).
// M2F_LN_V2 overrides: 0 = never staged, 1 = staged forward always, 2 = staged forward and backward always.
constexpr int64_t LN_STAGED_MIN_BYTES = 32ll << 20;
static int ln_v2_level(int64_t M, int64_t D) {
    if (const char* e = std::getenv("M2F_LN_V2")) return std::atoi(e);  // read per call: tests switch it inside one process
    return M * D * 4 >= LN_STAGED_MIN_BYTES ? 2 : 0;
}
static bool ln_v2_enabled(int64_t M, int64_t D) { return ln_v2_level(M, D) >= 1; }

static int ln_check(const void* x, int64_t x_ld, int64_t M, int64_t D, const char* who) {
    M2F_REQUIRE(x != nullptr && M > 0 && D > 0, M2F_ERR_ARG, "%s: bad argument", who);
    M2F_REQUIRE(D <= 1024 && (D % 4) == 0, M2F_ERR_UNSUPPORTED, "%s: D = %lld unsupported (need D %% 4 == 0 and D <= 1024)", who, (long long)D);
    M2F_REQUIRE((x_ld % 4) == 0 && (reinterpret_cast<uintptr_t>(x) % 16) == 0, M2F_ERR_ALIGN, "%s: rows must be 16-byte aligned", who);
    return M2F_OK;
}

#define M2F_LN_DISPATCH(D, CALL)                              \
    do {                                                      \
        const int nch_ = (int)((D + 127) / 128);              \
        switch (nch_) {                                       \
            case 1: { constexpr int NCH = 1; CALL; } break;   \
            case 2: { constexpr int NCH = 2; CALL; } break;   \
            case 3: { constexpr int NCH = 3; CALL; } break;   \
            case 4: { constexpr int NCH = 4; CALL; } break;   \
            case 5: { constexpr int NCH = 5; CALL; } break;   \
            case 6: { constexpr int NCH = 6; CALL; } break;   \
            case 7: { constexpr int NCH = 7; CALL; } break;   \
            default: { constexpr int NCH = 8; CALL; } break;  \
        }                                                     \
    } while (0)

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_ln_fwd(const float* x, int64_t x_ld, const float* gamma, const float* beta, const float* skip, int64_t skip_ld,
                          int64_t M, int64_t D, float* y_f32, int64_t y_ld, const m2f_planes_t* y_planes, float drop_p, uint64_t seed,
                          uint64_t step, uint32_t site, float* stats, void* stream) {
    M2F_TRY(ln_check(x, x_ld, M, D, "ln_fwd"));
    M2F_REQUIRE(gamma && beta, M2F_ERR_ARG, "ln_fwd: null gamma/beta");
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(gamma) % 16) == 0 && (reinterpret_cast<uintptr_t>(beta) % 16) == 0, M2F_ERR_ALIGN,
                "ln_fwd: gamma/beta must be 16-byte aligned");
    M2F_REQUIRE(!skip || ((skip_ld % 4) == 0 && (reinterpret_cast<uintptr_t>(skip) % 16) == 0), M2F_ERR_ALIGN, "ln_fwd: skip alignment");
    M2F_REQUIRE(!y_f32 || ((y_ld % 4) == 0 && (reinterpret_cast<uintptr_t>(y_f32) % 16) == 0), M2F_ERR_ALIGN, "ln_fwd: y alignment");
    M2F_REQUIRE(drop_p >= 0.f && drop_p < 1.f, M2F_ERR_ARG, "ln_fwd: dropout p must be in [0,1)");
    __nv_bfloat16* ypl = nullptr;
    int64_t ypl_ld = 0, ypl_stride = 0;
    int P = 1;
    if (y_planes && y_planes->ptr) {
        M2F_TRY(check_planes(y_planes, "ln_fwd y_planes", 4));
        M2F_REQUIRE(y_planes->rows >= M && y_planes->cols == D, M2F_ERR_ARG, "ln_fwd: y_planes extent mismatch");
        ypl = reinterpret_cast<__nv_bfloat16*>(y_planes->ptr);
        ypl_ld = y_planes->ld;
        ypl_stride = y_planes->plane_stride;
        P = y_planes->planes;
    }
    M2F_REQUIRE(y_f32 || ypl, M2F_ERR_ARG, "ln_fwd: no output given");
    const Drop drop = drop_make(drop_p, seed, step, site);
    auto st = reinterpret_cast<cudaStream_t>(stream);
    if (ln_v2_enabled(M, D) && (!skip || (reinterpret_cast<uintptr_t>(skip) % 16 == 0 && skip_ld % 4 == 0))) {
        const size_t smem = (size_t)LN2_STAGES * LN2_ROWS * D * 4 * (skip ? 2 : 1);
        int64_t blocks = (M + LN2_ROWS - 1) / LN2_ROWS;
        if (blocks > 148 * 2) blocks = 148 * 2;
        M2F_LN_DISPATCH(D, {
            M2F_CUDA(cudaFuncSetAttribute(ln_fwd_staged_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            M2F_CUDA(launch_chain(ln_fwd_staged_kernel<NCH>, dim3((unsigned)blocks), dim3(LN2_THREADS), smem, st, 1, x, x_ld, gamma, beta, skip,
                                  skip_ld, M, (int)D, y_f32, y_ld, ypl, ypl_ld, ypl_stride, P, drop, stats));
        });
        M2F_LAUNCH_CHECK();
        return M2F_OK;
    }
    static const int fwd_threads = [] {
        const char* e = std::getenv("M2F_LN_FWD_THREADS");
        const int v = e ? std::atoi(e) : LN_THREADS;
        return (v == 64 || v == 128 || v == 256) ? v : LN_THREADS;
    }();
    const int64_t fwd_blocks = std::min<int64_t>((M + fwd_threads / 32 - 1) / (fwd_threads / 32), 148 * 4 * (LN_THREADS / 32) / (fwd_threads / 32));
    M2F_LN_DISPATCH(D, M2F_CUDA(launch_chain(ln_fwd_kernel<NCH>, dim3((unsigned)std::max<int64_t>(fwd_blocks, 1)), dim3(fwd_threads), 0, st, 1, x, x_ld, gamma, beta, skip,
                                             skip_ld, M, (int)D, y_f32, y_ld, ypl, ypl_ld, ypl_stride, P, drop, stats)));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_ln_bwd(const float* g, int64_t g_ld, const float* x, int64_t x_ld, const float* gamma, const float* stats, int64_t M,
                          int64_t D, float* dx_f32, int64_t dx_ld, const m2f_planes_t* dx_planes, float drop_p, uint64_t seed, uint64_t step,
                          uint32_t site, float* dgamma, float* dbeta, float* dbias, void* stream) {
    M2F_TRY(ln_check(x, x_ld, M, D, "ln_bwd"));
    M2F_TRY(ln_check(g, g_ld, M, D, "ln_bwd"));
    M2F_REQUIRE(gamma && stats, M2F_ERR_ARG, "ln_bwd: null gamma/stats");
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(gamma) % 16) == 0, M2F_ERR_ALIGN, "ln_bwd: gamma must be 16-byte aligned");
    M2F_REQUIRE(!dx_f32 || ((dx_ld % 4) == 0 && (reinterpret_cast<uintptr_t>(dx_f32) % 16) == 0), M2F_ERR_ALIGN, "ln_bwd: dx alignment");
    M2F_REQUIRE(drop_p >= 0.f && drop_p < 1.f, M2F_ERR_ARG, "ln_bwd: dropout p must be in [0,1)");
    __nv_bfloat16* dpl = nullptr;
    int64_t dpl_ld = 0, dpl_stride = 0;
    int P = 1;
    if (dx_planes && dx_planes->ptr) {
        M2F_TRY(check_planes(dx_planes, "ln_bwd dx_planes", 4));
        M2F_REQUIRE(dx_planes->rows >= M && dx_planes->cols == D, M2F_ERR_ARG, "ln_bwd: dx_planes extent mismatch");
        dpl = reinterpret_cast<__nv_bfloat16*>(dx_planes->ptr);
        dpl_ld = dx_planes->ld;
        dpl_stride = dx_planes->plane_stride;
        P = dx_planes->planes;
    }
    const Drop drop = drop_make(drop_p, seed, step, site);
    auto st = reinterpret_cast<cudaStream_t>(stream);
    M2F_REQUIRE((!dgamma || (reinterpret_cast<uintptr_t>(dgamma) % 16) == 0) && (!dbeta || (reinterpret_cast<uintptr_t>(dbeta) % 16) == 0) &&
                    (!dbias || (reinterpret_cast<uintptr_t>(dbias) % 16) == 0), M2F_ERR_ALIGN, "ln_bwd: dgamma / dbeta / dbias must be 16-byte aligned");
    if (ln_v2_level(M, D) >= 2) {
        const size_t smem = (size_t)LN2_STAGES * LN2_ROWS * D * 4 * 2;
        int64_t blocks = (M + LN2_ROWS - 1) / LN2_ROWS;
        if (blocks > 148 * 2) blocks = 148 * 2;
        M2F_LN_DISPATCH(D, {
            M2F_CUDA(cudaFuncSetAttribute(ln_bwd_staged_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            M2F_CUDA(launch_chain(ln_bwd_staged_kernel<NCH>, dim3((unsigned)blocks), dim3(LN2_THREADS), smem, st, 1, g, g_ld, x, x_ld, gamma, stats,
                                  M, (int)D, dx_f32, dx_ld, dpl, dpl_ld, dpl_stride, P, drop, dgamma, dbeta, dbias));
        });
        M2F_LAUNCH_CHECK();
        return M2F_OK;
    }
    const int lnb_grid = (int)std::min<int64_t>((M + LNB_ROWS - 1) / LNB_ROWS, 148 * 4);
    const size_t lnb_smem = (size_t)3 * LNB_ROWS * D * sizeof(float);  // D <= 1024: <= 96 KB
    M2F_LN_DISPATCH(D, {
        static size_t set = 0;
        if (lnb_smem > 48 * 1024 && lnb_smem > set) {
            M2F_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lnb_smem));
            set = lnb_smem;
        }
        M2F_CUDA(launch_chain(ln_bwd_kernel<NCH>, dim3(lnb_grid), dim3(LNB_THREADS), lnb_smem, st, 1, g, g_ld, x, x_ld, gamma, stats, M, (int)D,
                              dx_f32, dx_ld, dpl, dpl_ld, dpl_stride, P, drop, dgamma, dbeta, dbias));
    });
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
