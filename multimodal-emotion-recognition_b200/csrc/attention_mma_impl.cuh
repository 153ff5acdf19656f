// Kernels of the short-dialogue attention fast path for ONE plane count P = M2F_ATT_P (included by attention_mma_p{1,2,3}.cu).
#pragma once
#include <math.h>

#include "attention_mma.cuh"

#ifndef M2F_ATT_P
#error "define M2F_ATT_P (1, 2 or 3) before including attention_mma_impl.cuh"
#endif

namespace m2f {
namespace amma {
namespace {
constexpr int P = M2F_ATT_P;


template <int HD, int SP>
struct Smem {
    static constexpr int LDS = HD + 8, LDP = SP + 8;
    static size_t fwd_bytes() { return (size_t)2 * P * SP * LDS * 2; }  // K and V; Q goes from global memory to A fragments
    static size_t bwd_bytes() { return (size_t)4 * P * SP * LDS * 2 + (size_t)2 * P * SP * LDP * 2; }
};

// scores (already scaled) -> probabilities in place; returns nothing.  Accumulator layout: thread (g, q) holds rows
// i0 + g (values [nt][0..1]) and i0 + g + 8 ([nt][2..3]), columns nt*8 + 2q, 2q + 1.
template <int SP>
__device__ __forceinline__ void softmax_rows(float (&s)[SP / 8][4], uint64_t kv, int lane) {
    const int q = lane & 3;
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < SP / 8; ++nt) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int j = nt * 8 + 2 * q + c;
            const bool dead = !((kv >> j) & 1ull);
            if (dead) { s[nt][c] = -INFINITY; s[nt][2 + c] = -INFINITY; }
            mx0 = fmaxf(mx0, s[nt][c]);
            mx1 = fmaxf(mx1, s[nt][2 + c]);
        }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    float sm0 = 0.f, sm1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < SP / 8; ++nt) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            s[nt][c] = (s[nt][c] == -INFINITY) ? 0.f : __expf(s[nt][c] - mx0);
            s[nt][2 + c] = (s[nt][2 + c] == -INFINITY) ? 0.f : __expf(s[nt][2 + c] - mx1);
            sm0 += s[nt][c];
            sm1 += s[nt][2 + c];
        }
    }
    sm0 += __shfl_xor_sync(0xffffffffu, sm0, 1); sm0 += __shfl_xor_sync(0xffffffffu, sm0, 2);
    sm1 += __shfl_xor_sync(0xffffffffu, sm1, 1); sm1 += __shfl_xor_sync(0xffffffffu, sm1, 2);
    const float i0 = 1.f / sm0, i1 = 1.f / sm1;  // every dialogue keeps >= 1 valid key
#pragma unroll
    for (int nt = 0; nt < SP / 8; ++nt) { s[nt][0] *= i0; s[nt][1] *= i0; s[nt][2] *= i1; s[nt][3] *= i1; }
}

template <int HD, int SP>
__global__ void __launch_bounds__(SP * 2)
attn_fwd_mma_kernel(View q, View k, View v, const uint8_t* __restrict__ key_pad, int S, int H, float scale, Drop drop, OutView o) {
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    extern __shared__ __align__(16) uint8_t smraw[];
    constexpr int LDS = HD + 8;
    __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smraw);
    __nv_bfloat16* sV = sK + P * SP * LDS;
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int64_t row0 = (int64_t)b * S;
    const int col0 = h * HD;
    load_tile<HD, SP>(sK, k, P, row0, col0, S);
    load_tile<HD, SP>(sV, v, P, row0, col0, S);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i0 = warp * 16;
    const bool active = i0 < S;  // idle strips (S <= 16 (warp - 1)) only take part in the barriers
    // Q is the A operand of one product and nothing else: its fragments come straight from global memory (under the K / V copies)
    // and the tile never occupies shared memory -- 60 -> 40 KB per CTA with two planes at hd 96, five CTAs per SM instead of three
    uint32_t qf[P][HD / 16][4];
    if (active) load_a_frags<HD, P>(qf, q, P, row0, col0, i0, S, lane);
    // nothing below depends on the tiles: the key mask (a global load) and the Philox words run under the copies
    const uint64_t kvalid = key_valid_bits(key_pad + row0, S, lane);
    const uint32_t dbits = (drop.enabled && active) ? drop_bits<SP>(drop, blockIdx.x, S, i0, lane) : 0xffffffffu;
    cp_async_wait_all();
    __syncthreads();
    const int g = lane >> 2, qd = lane & 3;
    (void)g; (void)qd;
    float acc[HD / 8][4];
    if (active) {
        float s[SP / 8][4];
#pragma unroll
        for (int nt = 0; nt < SP / 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        strip_nt_regA<HD, SP, P>(s, qf, sK, lane);
#pragma unroll
        for (int nt = 0; nt < SP / 8; ++nt) { s[nt][0] *= scale; s[nt][1] *= scale; s[nt][2] *= scale; s[nt][3] *= scale; }
        softmax_rows<SP>(s, kvalid, lane);
        if (drop.enabled) {
#pragma unroll
            for (int nt = 0; nt < SP / 8; ++nt)
#pragma unroll
                for (int c = 0; c < 4; ++c) s[nt][c] *= drop_mul(dbits, nt, c, drop.scale);
        }
#pragma unroll
        for (int nt = 0; nt < HD / 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
        strip_regA_nn_ct<HD, SP, P>(acc, s, sV, lane);
    }
    __syncthreads();  // every warp is done with K: its rows stage O (this warp's 16 rows), then coalesced rows go to global
    if (!active) return;
    stage_strip<HD, SP>(acc, 1.f, sK, P, i0, lane);
    copy_out_strip<HD, SP>(sK, o, P, row0, col0, i0, S, lane, nullptr);
}

// Backward.  Pd = P * M (M = dropout multiplier), dPd = dO V^T, dP = dPd * M, dS = P * (dP - rowsum(dP * P)),
// dQ = c dS K, dK = c dS^T Q, dV = Pd^T dO.  db* (nullable) += column sums of dQ / dK / dV (in-projection bias grads).
template <int HD, int SP>
__global__ void __launch_bounds__(SP * 2)
attn_bwd_mma_kernel(View q, View k, View v, View d_o, const uint8_t* __restrict__ key_pad, int S, int H, float scale, Drop drop,
                    OutView dq, OutView dk, OutView dv, float* dbq, float* dbk, float* dbv) {
    pdl_wait();
    pdl_trigger();
    drop = drop_resolve(drop);
    extern __shared__ __align__(16) uint8_t smraw[];
    constexpr int LDS = HD + 8, LDP = SP + 8;
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smraw);
    __nv_bfloat16* sK = sQ + P * SP * LDS;
    __nv_bfloat16* sV = sK + P * SP * LDS;
    __nv_bfloat16* sdO = sV + P * SP * LDS;
    __nv_bfloat16* sPd = sdO + P * SP * LDS;  // [P][SP queries][LDP keys]
    __nv_bfloat16* sdS = sPd + P * SP * LDP;
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int64_t row0 = (int64_t)b * S;
    const int col0 = h * HD;
    load_tile<HD, SP>(sQ, q, P, row0, col0, S);
    load_tile<HD, SP>(sK, k, P, row0, col0, S);
    load_tile<HD, SP>(sV, v, P, row0, col0, S);
    load_tile<HD, SP>(sdO, d_o, P, row0, col0, S);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i0 = warp * 16;  // blockDim = 32 * SP / 16: every warp owns one 16-row strip
    const int g = lane >> 2, qd = lane & 3;
    // nothing here depends on the tiles: the key mask (a global load) and the Philox words run under the copies
    const uint64_t kvalid = key_valid_bits(key_pad + row0, S, lane);
    const uint32_t bits = drop.enabled ? drop_bits<SP>(drop, blockIdx.x, S, i0, lane) : 0xffffffffu;
    cp_async_wait_all();
    __syncthreads();
    {
        float s[SP / 8][4], dp[SP / 8][4];
#pragma unroll
        for (int nt = 0; nt < SP / 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
        }
        strip_nt<HD, SP>(s, sQ, sK, P, i0, lane);
        strip_nt<HD, SP>(dp, sdO, sV, P, i0, lane);
#pragma unroll
        for (int nt = 0; nt < SP / 8; ++nt) { s[nt][0] *= scale; s[nt][1] *= scale; s[nt][2] *= scale; s[nt][3] *= scale; }
        softmax_rows<SP>(s, kvalid, lane);
        // dp <- dP = dPd * M; s stays P; pd = P * M written to shared; then dS
        const float mscale = drop.enabled ? drop.scale : 1.f;
        float del0 = 0.f, del1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < SP / 8; ++nt) {
            float pdv[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int i = i0 + g + (c >> 1) * 8;
                float m = drop_mul(bits, nt, c, mscale);
                if (i >= S) m = 0.f;  // rows beyond the dialogue contribute nothing to dK / dV
                dp[nt][c] *= m;
                pdv[c] = s[nt][c] * m;
                if (c < 2) del0 += dp[nt][c] * s[nt][c]; else del1 += dp[nt][c] * s[nt][c];
            }
            store_pair_planes(sPd + (i0 + g) * LDP + nt * 8 + 2 * qd, 8 * LDP, SP * LDP, P, pdv[0], pdv[1], pdv[2], pdv[3]);
        }
        del0 += __shfl_xor_sync(0xffffffffu, del0, 1); del0 += __shfl_xor_sync(0xffffffffu, del0, 2);
        del1 += __shfl_xor_sync(0xffffffffu, del1, 1); del1 += __shfl_xor_sync(0xffffffffu, del1, 2);
#pragma unroll
        for (int nt = 0; nt < SP / 8; ++nt) {
            dp[nt][0] = s[nt][0] * (dp[nt][0] - del0);
            dp[nt][1] = s[nt][1] * (dp[nt][1] - del0);
            dp[nt][2] = s[nt][2] * (dp[nt][2] - del1);
            dp[nt][3] = s[nt][3] * (dp[nt][3] - del1);
            store_pair_planes(sdS + (i0 + g) * LDP + nt * 8 + 2 * qd, 8 * LDP, SP * LDP, P, dp[nt][0], dp[nt][1], dp[nt][2], dp[nt][3]);
        }
    }
    // dQ strip = c * dS K   (K stored [key][d]: k = key, n = d); dS read back from shared (A = sdS rows i0.., k = key)
    __syncthreads();
    float acc[HD / 8][4];
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
    strip_smemA_nn<HD, SP>(acc, sdS, sK, P, i0, lane);
    // this warp's 16 keys j0 = i0: dV = Pd^T dO, dK = c dS^T Q
    float accv[HD / 8][4];
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) accv[nt][0] = accv[nt][1] = accv[nt][2] = accv[nt][3] = 0.f;
    strip_tn<HD, SP>(accv, sPd, sdO, P, i0, lane);
    __syncthreads();  // every warp is done reading sK / sV / sdO: they become the staging tiles of dQ / dV / dK
    stage_strip<HD, SP>(acc, scale, sK, P, i0, lane);
    copy_out_strip<HD, SP>(sK, dq, P, row0, col0, i0, S, lane, dbq);
    stage_strip<HD, SP>(accv, 1.f, sV, P, i0, lane);
    copy_out_strip<HD, SP>(sV, dv, P, row0, col0, i0, S, lane, dbv);
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
    strip_tn<HD, SP>(acc, sdS, sQ, P, i0, lane);
    stage_strip<HD, SP>(acc, scale, sdO, P, i0, lane);
    copy_out_strip<HD, SP>(sdO, dk, P, row0, col0, i0, S, lane, dbk);
}

template <int HD, int SP>
static int launch_fwd(const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, float scale,
                      const Drop& drop, const OutView& o, cudaStream_t st) {
    const size_t smem = Smem<HD, SP>::fwd_bytes();
    static size_t set = 0;
    if (smem > 48 * 1024 && smem > set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_fwd_mma_kernel<HD, SP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        set = smem;
    }
    M2F_CUDA(launch_chain(attn_fwd_mma_kernel<HD, SP>, dim3((unsigned)BH), dim3(SP * 2), smem, st, 1, q, k, v, pad, S, H, scale, drop, o));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
template <int HD, int SP>
static int launch_bwd(const View& q, const View& k, const View& v, const View& d_o, const uint8_t* pad, int64_t BH, int S, int H,
                      float scale, const Drop& drop, const OutView& dq, const OutView& dk, const OutView& dv, float* dbq, float* dbk,
                      float* dbv, cudaStream_t st) {
    const size_t smem = Smem<HD, SP>::bwd_bytes();
    static size_t set = 0;
    if (smem > 48 * 1024 && smem > set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_bwd_mma_kernel<HD, SP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        set = smem;
    }
    M2F_CUDA(launch_chain(attn_bwd_mma_kernel<HD, SP>, dim3((unsigned)BH), dim3(SP * 2), smem, st, 1, q, k, v, d_o, pad, S, H, scale, drop, dq, dk,
                          dv, dbq, dbk, dbv));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

#define M2F_ATT_DISPATCH(HDV, SPV, CALL)                                                     \
    do {                                                                                     \
        if (HDV == 16) { M2F_ATT_SP(16, SPV, CALL); }                                        \
        else if (HDV == 32) { M2F_ATT_SP(32, SPV, CALL); }                                   \
        else if (HDV == 64) { M2F_ATT_SP(64, SPV, CALL); }                                   \
        else if (HDV == 96) { M2F_ATT_SP(96, SPV, CALL); }                                   \
        else { M2F_ATT_SP(128, SPV, CALL); }                                                 \
    } while (0)
#define M2F_ATT_SP(HDC, SPV, CALL)                                                           \
    do {                                                                                     \
        constexpr int HD_ = HDC;                                                             \
        if (SPV == 16) { constexpr int SP_ = 16; return CALL; }                              \
        else if (SPV == 32) { constexpr int SP_ = 32; return CALL; }                         \
        else if (SPV == 48) { constexpr int SP_ = 48; return CALL; }                         \
        else { constexpr int SP_ = 64; return CALL; }                                        \
    } while (0)


}  // namespace

#define M2F_AMMA_CAT2(a, b) a##b
#define M2F_AMMA_CAT(a, b) M2F_AMMA_CAT2(a, b)

int M2F_AMMA_CAT(fwd_p, M2F_ATT_P)(AMMA_FWD_ARGS) {
    M2F_ATT_DISPATCH(hd, SPv, (launch_fwd<HD_, SP_>(q, k, v, pad, BH, S, H, scale, drop, o, st)));
}
int M2F_AMMA_CAT(bwd_p, M2F_ATT_P)(AMMA_BWD_ARGS) {
    M2F_ATT_DISPATCH(hd, SPv, (launch_bwd<HD_, SP_>(q, k, v, d_o, pad, BH, S, H, scale, drop, dq, dk, dv, dbq, dbk, dbv, st)));
}

}  // namespace amma
}  // namespace m2f
