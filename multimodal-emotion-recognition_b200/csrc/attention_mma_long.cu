// Long-dialogue attention forward (64 < S <= 256), no probability dropout: the inference shape of BASELINE.json's
// config 5 (test() body, src/test.py:58-71; S = 256), in bf16 mode (1 plane) and fp32-parity mode (2 / 3 planes).
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
template <int HD, int QB, int PC>
__global__ void __launch_bounds__(QB * 2)
attn_fwd_long_kernel(View q, View k, View v, const uint8_t* __restrict__ key_pad, int S, int H, int planes, int nbuf, float scale, OutView o) {
    const int P = PC ? PC : planes;
    pdl_wait();
    pdl_trigger();
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
    const float r0 = 1.f / l0, r1 = 1.f / l1;  // every dialogue keeps >= 1 valid key
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) { acc[nt][0] *= r0; acc[nt][1] *= r0; acc[nt][2] *= r1; acc[nt][3] *= r1; }
    // this warp's 16 rows of sQ are dead (only this warp read them): stage O there, then coalesced rows to global
    stage_strip<HD, QB>(acc, 1.f, sQ, P, i0, lane);
    copy_out_strip<HD, QB>(sQ, o, P, row0 + q0, col0, i0, S - q0, lane, nullptr);
}

template <int HD, int QB, int PC>
int launch_long(const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, int P, float scale, const OutView& o,
                cudaStream_t st) {
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
        M2F_CUDA(cudaFuncSetAttribute(attn_fwd_long_kernel<HD, QB, PC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        set = smem;
    }
    M2F_CUDA(launch_chain(attn_fwd_long_kernel<HD, QB, PC>, dim3((unsigned)BH, (unsigned)((S + QB - 1) / QB)), dim3(QB * 2), smem, st, 1, q, k, v, pad,
                          S, H, P, nbuf, scale, o));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

template <int HD>
int launch_long_hd(const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, int P, float scale, const OutView& o,
                   cudaStream_t st) {
    return P == 1 ? launch_long<HD, 128, 1>(q, k, v, pad, BH, S, H, P, scale, o, st) : launch_long<HD, KT, 0>(q, k, v, pad, BH, S, H, P, scale, o, st);
}

}  // namespace

bool long_supported(int hd, int64_t S) { return (hd == 32 || hd == 64 || hd == 96 || hd == 128) && S > 64 && S <= 256; }

int fwd_long(int hd, int P, const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, float scale,
             const OutView& o, cudaStream_t st) {
    switch (hd) {
        case 32: return launch_long_hd<32>(q, k, v, pad, BH, S, H, P, scale, o, st);
        case 64: return launch_long_hd<64>(q, k, v, pad, BH, S, H, P, scale, o, st);
        case 96: return launch_long_hd<96>(q, k, v, pad, BH, S, H, P, scale, o, st);
        default: return launch_long_hd<128>(q, k, v, pad, BH, S, H, P, scale, o, st);
    }
}

}  // namespace amma
}  // namespace m2f
