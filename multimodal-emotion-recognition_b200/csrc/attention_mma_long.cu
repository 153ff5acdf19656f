// Long-dialogue attention (64 < S <= 4096) on mma.sync: the forward -- the inference shape of BASELINE.json's config 5 (test() body,
// src/test.py:58-71; S = 256) in bf16 mode (1 plane) and fp32-parity mode (2 / 3 planes), and, with probability dropout
// (template DROP), the training forward at these lengths -- and the backward (attn_bwd_long_kernel, below).
// Same arithmetic as the short-dialogue fast path (torch/nn/functional.py:6607-6620: softmax(QK^T / sqrt(hd) + key
// mask) V) with the softmax taken online over key tiles of 64, so that a query strip never holds more than a 16 x 64
// score tile in registers.
// One CTA per (dialogue, head, block of QB query rows), one warp per 16 rows.  K / V tiles (64 keys, all planes) go
// through a ring of `nbuf` (two by default) shared-memory buffers filled by cp.async, one commit group per tile: the
// next tile streams in under the current one's plane products, and the footprint leaves room for a second CTA.  Tensor work is mma.sync.m16n8k16 via the helpers of attention_mma.cuh.
#include <math.h>
#include <stdlib.h>

#include "attention_mma.cuh"

namespace m2f {
namespace amma {
namespace {

constexpr int KT = 64;  // keys per online-softmax step
constexpr size_t SMEM_CAP = 227 * 1024;

__device__ __forceinline__ void wait_pending(int n) {  // cp.async groups that may still be in flight
    switch (n) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
    }
}

__device__ __forceinline__ float exp_for(int P, float x) { return P == 1 ? __expf(x) : expf(x); }  // bf16 mode: fast exp

// QB == KT whenever P > 1: strip_nt addresses the planes of both operands with one plane stride.
// PC: plane count known at compile time (1: the bf16 mode, plane loops fold away) or 0 (run-time planes).
// DROP: dropout on the probabilities (training): the row sums run over the unmasked exponentials, the P V products over the kept
// ones, and 1 / (1 - p) rides on the final 1 / l.
template <int HD, int QB, int PC, bool DROP>
__global__ void __launch_bounds__(QB * 2)
attn_fwd_long_kernel(View q, View k, View v, const uint8_t* __restrict__ key_pad, int S, int H, int planes, int nbuf, float scale, Drop drop,
                     OutView o) {
    const int P = PC ? PC : planes;
    pdl_wait();
    pdl_trigger();
    if (DROP) drop = drop_resolve(drop);
    extern __shared__ __align__(16) uint8_t smraw[];
    constexpr int LDS = HD + 8;
    const int nkt = (S + KT - 1) / KT;
    const int tile_elems = P * KT * LDS;
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smraw);
    __nv_bfloat16* sK = sQ + P * QB * LDS;   // nbuf tiles
    __nv_bfloat16* sV = sK + nbuf * tile_elems;
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int q0 = blockIdx.y * QB;
    const int64_t row0 = (int64_t)b * S;
    const int col0 = h * HD;
    auto issue = [&](int t) {
        const int buf = t % nbuf;
        load_tile<HD, KT>(sK + buf * tile_elems, k, P, row0 + t * KT, col0, S - t * KT);
        load_tile<HD, KT>(sV + buf * tile_elems, v, P, row0 + t * KT, col0, S - t * KT);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    load_tile<HD, QB>(sQ, q, P, row0 + q0, col0, S - q0);  // rides in the first tile's group
    for (int t = 0; t < nbuf && t < nkt; ++t) issue(t);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i0 = warp * 16;
    const bool active = q0 + i0 < S;  // idle strips still load and take part in the barriers
    const int qd = lane & 3;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;  // running row maxima / (per-thread partial) row sums
    float acc[HD / 8][4];
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll 1
    for (int t = 0; t < nkt; ++t) {
        wait_pending(min(nbuf, nkt - t) - 1);
        __syncthreads();
        if (active) {
            const __nv_bfloat16* tK = sK + (t % nbuf) * tile_elems;
            const __nv_bfloat16* tV = sV + (t % nbuf) * tile_elems;
            float s[KT / 8][4];
#pragma unroll
            for (int nt = 0; nt < KT / 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            strip_nt<HD, KT>(s, sQ, tK, P, i0, lane);
            const uint64_t kv = key_valid_bits(key_pad + row0 + t * KT, S - t * KT, lane);
            float t0 = -INFINITY, t1 = -INFINITY;
#pragma unroll
            for (int nt = 0; nt < KT / 8; ++nt) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const bool live = (kv >> (nt * 8 + 2 * qd + c)) & 1ull;
                    s[nt][c] = live ? s[nt][c] * scale : -INFINITY;
                    s[nt][2 + c] = live ? s[nt][2 + c] * scale : -INFINITY;
                    t0 = fmaxf(t0, s[nt][c]);
                    t1 = fmaxf(t1, s[nt][2 + c]);
                }
            }
            t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 1)); t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 2));
            t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 1)); t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 2));
            const float n0 = fmaxf(m0, t0), n1 = fmaxf(m1, t1);
            // a tile (or everything so far) may be fully masked: keep exp() away from (-inf) - (-inf)
            const float c0 = (m0 == -INFINITY) ? 0.f : exp_for(P, m0 - n0), c1 = (m1 == -INFINITY) ? 0.f : exp_for(P, m1 - n1);
            float p0 = 0.f, p1 = 0.f;
#pragma unroll
            for (int nt = 0; nt < KT / 8; ++nt) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    s[nt][c] = (s[nt][c] == -INFINITY) ? 0.f : exp_for(P, s[nt][c] - n0);
                    s[nt][2 + c] = (s[nt][2 + c] == -INFINITY) ? 0.f : exp_for(P, s[nt][2 + c] - n1);
                    p0 += s[nt][c];
                    p1 += s[nt][2 + c];
                }
            }
            l0 = l0 * c0 + p0;
            l1 = l1 * c1 + p1;
            m0 = n0;
            m1 = n1;
            if (DROP) {
                const uint32_t bits = drop_bits<KT>(drop, blockIdx.x, S, q0 + i0, lane, t * (KT / 8));
#pragma unroll
                for (int nt = 0; nt < KT / 8; ++nt)
#pragma unroll
                    for (int c = 0; c < 4; ++c) s[nt][c] *= drop_mul(bits, nt, c, 1.f);
            }
#pragma unroll
            for (int nt = 0; nt < HD / 8; ++nt) { acc[nt][0] *= c0; acc[nt][1] *= c0; acc[nt][2] *= c1; acc[nt][3] *= c1; }
            strip_regA_nn<HD, KT>(acc, s, tV, P, lane);
        }
        if (t + nbuf < nkt) {  // refill the buffer just consumed
            __syncthreads();
            issue(t + nbuf);
        }
    }
    if (!active) return;
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float keep_inv = DROP ? drop.scale : 1.f;
    const float r0 = keep_inv / l0, r1 = keep_inv / l1;  // every dialogue keeps >= 1 valid key
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) { acc[nt][0] *= r0; acc[nt][1] *= r0; acc[nt][2] *= r1; acc[nt][3] *= r1; }
    // this warp's 16 rows of sQ are dead (only this warp read them): stage O there, then coalesced rows to global
    stage_strip<HD, QB>(acc, 1.f, sQ, P, i0, lane);
    copy_out_strip<HD, QB>(sQ, o, P, row0 + q0, col0, i0, S - q0, lane, nullptr);
}

template <int HD, int QB, int PC, bool DROP>
int launch_long(const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, int P, float scale, const Drop& drop,
                const OutView& o, cudaStream_t st) {
    constexpr int LDS = HD + 8;
    const int nkt = (S + KT - 1) / KT;
    const size_t q_bytes = (size_t)P * QB * LDS * 2, tile_bytes = (size_t)2 * P * KT * LDS * 2;  // K + V of one tile
    int nbuf = (int)((SMEM_CAP - q_bytes) / tile_bytes);
    nbuf = nbuf > nkt ? nkt : nbuf;
    // two tiles in flight measured best at S = 256 (a smaller footprint puts two CTAs on an SM: 47.3 ms per C5 batch
    // against 49.5 with the whole dialogue resident); M2F_ATT_LONG_NBUF = 1..4 overrides for experiments
    static const int cap = [] { const char* e = getenv("M2F_ATT_LONG_NBUF"); const int v = e ? atoi(e) : 0; return v >= 1 && v <= 4 ? v : 2; }();
    nbuf = nbuf > cap ? cap : nbuf;
    M2F_REQUIRE(nbuf >= 1, M2F_ERR_UNSUPPORTED, "attn_fwd: head_dim %d with %d planes does not fit a key tile in shared memory", HD, P);
    const size_t smem = q_bytes + nbuf * tile_bytes;
    static size_t set = 0;
    if (smem > 48 * 1024 && smem > set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_fwd_long_kernel<HD, QB, PC, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        set = smem;
    }
    M2F_CUDA(launch_chain(attn_fwd_long_kernel<HD, QB, PC, DROP>, dim3((unsigned)BH, (unsigned)((S + QB - 1) / QB)), dim3(QB * 2), smem, st, 1, q, k,
                          v, pad, S, H, P, nbuf, scale, drop, o));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

template <int HD>
int launch_long_hd(const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, int P, float scale, const Drop& drop,
                   const OutView& o, cudaStream_t st) {
    if (drop.enabled)
        return P == 1 ? launch_long<HD, 128, 1, true>(q, k, v, pad, BH, S, H, P, scale, drop, o, st)
                      : launch_long<HD, KT, 0, true>(q, k, v, pad, BH, S, H, P, scale, drop, o, st);
    return P == 1 ? launch_long<HD, 128, 1, false>(q, k, v, pad, BH, S, H, P, scale, drop, o, st)
                  : launch_long<HD, KT, 0, false>(q, k, v, pad, BH, S, H, P, scale, drop, o, st);
}

// ------------------------------------------------------------------------------------------------ backward, 64 < S <= 4096
// dQ, dK, dV of softmax(Q K^T c + mask) (dropout) V for dialogues that do not fit the single-tile kernel of attention_mma_impl.cuh
// (torch autograd of torch/nn/functional.py:6607-6690 is what it has to match).  One CTA per (dialogue, head) walks T x T tiles of
// the S x S probability matrix three times, with nothing but the tile operands on chip and no scratch memory:
//   phase 1  per query tile, over all key tiles: row statistics, online -- lse_i = log sum_j exp(s_ij) and
//            delta_i = sum_j dP_ij P_ij (dP = (dO V^T) * dropout multiplier), kept in shared memory (2 floats per row);
//   phase 2  per query tile, over all key tiles: dS = P * (dP - delta) with P = exp(s - lse) recomputed, dQ += c dS K in registers;
//   phase 3  per key tile, over all query tiles: Pd = P * multiplier and dS through shared memory, dV += Pd^T dO, dK += c dS^T Q.
// Scores and dO V^T are recomputed in every phase (5 + 4 tile products against the 5 of a kernel that keeps S x S on chip): at these
// lengths the training step is GEMM-bound and this kernel trades flops for zero extra memory.  T = 64 (one / two planes) or 32
// (three planes: four operand tiles and two T x T tiles of all planes have to fit 227 KB at head size 128).
template <int HD, int T>
__global__ void __launch_bounds__(T * 2)
attn_bwd_long_kernel(View q, View k, View v, View d_o, const uint8_t* __restrict__ key_pad, int S, int H, int P, float scale, Drop drop, OutView dq,
                     OutView dk, OutView dv, float* dbq, float* dbk, float* dbv) {
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    extern __shared__ __align__(16) uint8_t smraw[];
    constexpr int LDS = HD + 8, LDP = T + 8;
    const int ntile = (S + T - 1) / T;
    const int hd_tile = P * T * LDS, pp_tile = P * T * LDP;
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smraw);
    __nv_bfloat16* sdO = sQ + hd_tile;
    __nv_bfloat16* sK = sdO + hd_tile;
    __nv_bfloat16* sV = sK + hd_tile;
    __nv_bfloat16* sPd = sV + hd_tile;   // [P][T queries][LDP keys]
    __nv_bfloat16* sdS = sPd + pp_tile;
    float* stats = reinterpret_cast<float*>(sdS + pp_tile);  // [ntile * T][2]: lse, delta
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int64_t row0 = (int64_t)b * S;
    const int col0 = h * HD;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i0 = warp * 16;  // this warp's strip inside a tile: query rows in phases 1 / 2, key rows for dK / dV in phase 3
    const int g = lane >> 2, qd = lane & 3;
    const float mscale = drop.enabled ? drop.scale : 1.f;

    // scores (scaled, dead keys at -inf) and dO V^T of this warp's query strip against the key tile in sK / sV
    auto strips = [&](float (&s)[T / 8][4], float (&dp)[T / 8][4], int kt) {
#pragma unroll
        for (int nt = 0; nt < T / 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
        }
        strip_nt<HD, T>(s, sQ, sK, P, i0, lane);
        strip_nt<HD, T>(dp, sdO, sV, P, i0, lane);
        const uint64_t kv = key_valid_bits(key_pad + row0 + kt * T, S - kt * T, lane);
#pragma unroll
        for (int nt = 0; nt < T / 8; ++nt)
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const bool live = (kv >> (nt * 8 + 2 * qd + c)) & 1ull;
                s[nt][c] = live ? s[nt][c] * scale : -INFINITY;
                s[nt][2 + c] = live ? s[nt][2 + c] * scale : -INFINITY;
            }
    };
    // does key tile kt hold a usable key?  The same answer in every thread (all read the same mask bytes).
    auto tile_live = [&](int kt) -> bool {
        uint64_t kv = key_valid_bits(key_pad + row0 + kt * T, S - kt * T, lane);
        if constexpr (T < 64) kv &= (1ull << T) - 1ull;
        return kv != 0ull;
    };
    auto load_q_do = [&](int qi) {
        load_tile<HD, T>(sQ, q, P, row0 + qi * T, col0, S - qi * T);
        load_tile<HD, T>(sdO, d_o, P, row0 + qi * T, col0, S - qi * T);
    };
    auto load_k_v = [&](int kt) {
        load_tile<HD, T>(sK, k, P, row0 + kt * T, col0, S - kt * T);
        load_tile<HD, T>(sV, v, P, row0 + kt * T, col0, S - kt * T);
    };

    // ---- phase 1: lse and delta of every query row
#pragma unroll 1
    for (int qi = 0; qi < ntile; ++qi) {
        __syncthreads();  // the previous tile's readers are done
        load_q_do(qi);
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f, a0 = 0.f, a1 = 0.f;
#pragma unroll 1
        for (int kt = 0; kt < ntile; ++kt) {
            if (!tile_live(kt)) continue;  // no usable key in the tile: it adds nothing to any row (block-uniform)
            __syncthreads();
            load_k_v(kt);
            cp_async_wait_all();
            __syncthreads();
            float s[T / 8][4], dp[T / 8][4];
            strips(s, dp, kt);
            const uint32_t bits = drop.enabled ? drop_bits<T>(drop, blockIdx.x, S, qi * T + i0, lane, kt * (T / 8)) : 0xffffffffu;
            float t0 = -INFINITY, t1 = -INFINITY;
#pragma unroll
            for (int nt = 0; nt < T / 8; ++nt) {
                t0 = fmaxf(t0, fmaxf(s[nt][0], s[nt][1]));
                t1 = fmaxf(t1, fmaxf(s[nt][2], s[nt][3]));
            }
            t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 1)); t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 2));
            t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 1)); t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 2));
            const float n0 = fmaxf(m0, t0), n1 = fmaxf(m1, t1);
            const float c0 = (m0 == -INFINITY) ? 0.f : expf(m0 - n0), c1 = (m1 == -INFINITY) ? 0.f : expf(m1 - n1);
            float p0 = 0.f, p1 = 0.f, w0 = 0.f, w1 = 0.f;
#pragma unroll
            for (int nt = 0; nt < T / 8; ++nt)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float sv = s[nt][c];
                    const float e = (sv == -INFINITY) ? 0.f : expf(sv - (c < 2 ? n0 : n1));
                    const float w = e * dp[nt][c] * drop_mul(bits, nt, c, mscale);
                    if (c < 2) { p0 += e; w0 += w; } else { p1 += e; w1 += w; }
                }
            l0 = l0 * c0 + p0; a0 = a0 * c0 + w0; m0 = n0;
            l1 = l1 * c1 + p1; a1 = a1 * c1 + w1; m1 = n1;
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        a0 += __shfl_xor_sync(0xffffffffu, a0, 1); a0 += __shfl_xor_sync(0xffffffffu, a0, 2);
        a1 += __shfl_xor_sync(0xffffffffu, a1, 1); a1 += __shfl_xor_sync(0xffffffffu, a1, 2);
        if (qd == 0) {  // every dialogue keeps >= 1 valid key: l > 0
            const int r = qi * T + i0 + g;
            stats[2 * r] = m0 + logf(l0);
            stats[2 * r + 1] = a0 / l0;
            stats[2 * (r + 8)] = m1 + logf(l1);
            stats[2 * (r + 8) + 1] = a1 / l1;
        }
    }

    // ---- phase 2: dQ
#pragma unroll 1
    for (int qi = 0; qi < ntile; ++qi) {
        __syncthreads();  // phase 1's statistics are visible; the previous tile's staging rows have been copied out
        load_q_do(qi);
        float acc[HD / 8][4];
#pragma unroll
        for (int nt = 0; nt < HD / 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
        const int r = qi * T + i0 + g;
        const float lse0 = stats[2 * r], del0 = stats[2 * r + 1], lse1 = stats[2 * (r + 8)], del1 = stats[2 * (r + 8) + 1];
#pragma unroll 1
        for (int kt = 0; kt < ntile; ++kt) {
            if (!tile_live(kt)) continue;
            __syncthreads();
            load_k_v(kt);
            cp_async_wait_all();
            __syncthreads();
            float s[T / 8][4], dp[T / 8][4];
            strips(s, dp, kt);
            const uint32_t bits = drop.enabled ? drop_bits<T>(drop, blockIdx.x, S, qi * T + i0, lane, kt * (T / 8)) : 0xffffffffu;
#pragma unroll
            for (int nt = 0; nt < T / 8; ++nt)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float sv = s[nt][c];
                    const float pr = (sv == -INFINITY) ? 0.f : expf(sv - (c < 2 ? lse0 : lse1));
                    s[nt][c] = pr * (dp[nt][c] * drop_mul(bits, nt, c, mscale) - (c < 2 ? del0 : del1));
                }
            strip_regA_nn<HD, T>(acc, s, sK, P, lane);  // dQ strip += dS (16 x T) K (T x HD)
        }
        // only this warp read its 16 rows of sQ: they stage the strip
        stage_strip<HD, T>(acc, scale, sQ, P, i0, lane);
        copy_out_strip<HD, T>(sQ, dq, P, row0 + qi * T, col0, i0, S - qi * T, lane, (S - qi * T - i0 > 0) ? dbq : nullptr);
    }

    // ---- phase 3: dK, dV
#pragma unroll 1
    for (int kt = 0; kt < ntile; ++kt) {
        __syncthreads();
        const bool live = tile_live(kt);  // a tile without a usable key gets zero gradients: the query loop is left out
        if (live) load_k_v(kt);
        float accv[HD / 8][4], acck[HD / 8][4];
#pragma unroll
        for (int nt = 0; nt < HD / 8; ++nt) {
            accv[nt][0] = accv[nt][1] = accv[nt][2] = accv[nt][3] = 0.f;
            acck[nt][0] = acck[nt][1] = acck[nt][2] = acck[nt][3] = 0.f;
        }
#pragma unroll 1
        for (int qi = 0; qi < (live ? ntile : 0); ++qi) {
            if (qi) __syncthreads();  // the previous query tile's dV / dK products have read sQ, sdO, sPd, sdS
            load_q_do(qi);
            cp_async_wait_all();
            __syncthreads();
            {
                float s[T / 8][4], dp[T / 8][4];
                strips(s, dp, kt);
                const uint32_t bits = drop.enabled ? drop_bits<T>(drop, blockIdx.x, S, qi * T + i0, lane, kt * (T / 8)) : 0xffffffffu;
                const int r = qi * T + i0 + g;
                const float lse0 = stats[2 * r], del0 = stats[2 * r + 1], lse1 = stats[2 * (r + 8)], del1 = stats[2 * (r + 8) + 1];
                const float live0 = r < S ? 1.f : 0.f, live1 = r + 8 < S ? 1.f : 0.f;  // rows beyond the dialogue contribute nothing
#pragma unroll
                for (int nt = 0; nt < T / 8; ++nt) {
                    float pdv[4], dsv[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float sv = s[nt][c];
                        const float pr = (sv == -INFINITY) ? 0.f : expf(sv - (c < 2 ? lse0 : lse1));
                        const float mul = drop_mul(bits, nt, c, mscale) * (c < 2 ? live0 : live1);
                        pdv[c] = pr * mul;
                        dsv[c] = pr * (dp[nt][c] * mul - (c < 2 ? del0 : del1)) * (c < 2 ? live0 : live1);
                    }
                    store_pair_planes(sPd + (i0 + g) * LDP + nt * 8 + 2 * qd, 8 * LDP, T * LDP, P, pdv[0], pdv[1], pdv[2], pdv[3]);
                    store_pair_planes(sdS + (i0 + g) * LDP + nt * 8 + 2 * qd, 8 * LDP, T * LDP, P, dsv[0], dsv[1], dsv[2], dsv[3]);
                }
            }
            __syncthreads();
            strip_tn<HD, T>(accv, sPd, sdO, P, i0, lane);  // dV (keys i0..) += Pd^T dO
            strip_tn<HD, T>(acck, sdS, sQ, P, i0, lane);   // dK (keys i0..) += dS^T Q
        }
        __syncthreads();  // every warp is done reading sK / sV: they become the staging tiles
        stage_strip<HD, T>(accv, 1.f, sV, P, i0, lane);
        copy_out_strip<HD, T>(sV, dv, P, row0 + kt * T, col0, i0, S - kt * T, lane, (S - kt * T - i0 > 0) ? dbv : nullptr);
        stage_strip<HD, T>(acck, scale, sK, P, i0, lane);
        copy_out_strip<HD, T>(sK, dk, P, row0 + kt * T, col0, i0, S - kt * T, lane, (S - kt * T - i0 > 0) ? dbk : nullptr);
    }
}

template <int HD, int T>
int launch_bwd_long(const View& q, const View& k, const View& v, const View& d_o, const uint8_t* pad, int64_t BH, int S, int H, int P, float scale,
                    const Drop& drop, const OutView& dq, const OutView& dk, const OutView& dv, float* dbq, float* dbk, float* dbv, cudaStream_t st) {
    constexpr int LDS = HD + 8, LDP = T + 8;
    const int ntile = (S + T - 1) / T;
    const size_t smem = ((size_t)4 * P * T * LDS + (size_t)2 * P * T * LDP) * 2 + (size_t)ntile * T * 2 * sizeof(float);
    M2F_REQUIRE(smem <= SMEM_CAP, M2F_ERR_UNSUPPORTED, "attn_bwd: head_dim %d with %d planes needs %zu bytes of shared memory", HD, P, smem);
    static size_t set = 0;
    if (smem > 48 * 1024 && smem > set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_bwd_long_kernel<HD, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        set = smem;
    }
    M2F_CUDA(launch_chain(attn_bwd_long_kernel<HD, T>, dim3((unsigned)BH), dim3(T * 2), smem, st, 1, q, k, v, d_o, pad, S, H, P, scale, drop, dq, dk,
                          dv, dbq, dbk, dbv));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

template <int HD>
int launch_bwd_long_hd(const View& q, const View& k, const View& v, const View& d_o, const uint8_t* pad, int64_t BH, int S, int H, int P,
                       float scale, const Drop& drop, const OutView& dq, const OutView& dk, const OutView& dv, float* dbq, float* dbk, float* dbv,
                       cudaStream_t st) {
    return P <= 2 ? launch_bwd_long<HD, 64>(q, k, v, d_o, pad, BH, S, H, P, scale, drop, dq, dk, dv, dbq, dbk, dbv, st)
                  : launch_bwd_long<HD, 32>(q, k, v, d_o, pad, BH, S, H, P, scale, drop, dq, dk, dv, dbq, dbk, dbv, st);
}

}  // namespace

// tile loops: no structural bound on S; 4096 keeps the backward's row statistics (8 bytes per row) a small part of shared memory
bool long_supported(int hd, int64_t S) { return (hd == 32 || hd == 64 || hd == 96 || hd == 128) && S > 64 && S <= 4096; }

int fwd_long(int hd, int P, const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, float scale,
             const Drop& drop, const OutView& o, cudaStream_t st) {
    switch (hd) {
        case 32: return launch_long_hd<32>(q, k, v, pad, BH, S, H, P, scale, drop, o, st);
        case 64: return launch_long_hd<64>(q, k, v, pad, BH, S, H, P, scale, drop, o, st);
        case 96: return launch_long_hd<96>(q, k, v, pad, BH, S, H, P, scale, drop, o, st);
        default: return launch_long_hd<128>(q, k, v, pad, BH, S, H, P, scale, drop, o, st);
    }
}

int bwd_long(int hd, int P, const View& q, const View& k, const View& v, const View& d_o, const uint8_t* pad, int64_t BH, int S, int H, float scale,
             const Drop& drop, const OutView& dq, const OutView& dk, const OutView& dv, float* dbq, float* dbk, float* dbv, cudaStream_t st) {
    switch (hd) {
        case 32: return launch_bwd_long_hd<32>(q, k, v, d_o, pad, BH, S, H, P, scale, drop, dq, dk, dv, dbq, dbk, dbv, st);
        case 64: return launch_bwd_long_hd<64>(q, k, v, d_o, pad, BH, S, H, P, scale, drop, dq, dk, dv, dbq, dbk, dbv, st);
        case 96: return launch_bwd_long_hd<96>(q, k, v, d_o, pad, BH, S, H, P, scale, drop, dq, dk, dv, dbq, dbk, dbv, st);
        default: return launch_bwd_long_hd<128>(q, k, v, d_o, pad, BH, S, H, P, scale, drop, dq, dk, dv, dbq, dbk, dbv, st);
    }
}

}  // namespace amma
}  // namespace m2f
