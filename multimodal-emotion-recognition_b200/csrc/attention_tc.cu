// Long-dialogue attention forward on the 5th-generation tensor cores (tcgen05 + TMEM + TMA): the inference shape of
// BASELINE.json's config 5 (test() body, src/test.py:58-71: S = 256, batch 512), bf16 mode (one plane), head_dim 96 (also 64 / 128).
// Same arithmetic as torch/nn/functional.py:6607-6620 -- softmax(Q K^T / sqrt(hd) + key-padding mask) V, no dropout.
//
// One CTA per (dialogue, head), one or two tiles of 128 query rows; 64 < S <= 256, so the whole key axis is ONE MMA tile:
//   TMA      Q [128 x hd], K [256 x hd] (K-major, 128-byte swizzle, one box per 64-column atom) and V [256 x hd] (read as the
//            MN-major B operand: 64-key x 64-column boxes) land in shared memory; rows past the dialogue belong to the next
//            dialogue (or are zero-filled past the tensor) and are masked out below.
//   MMA 1    S = Q K^T: one elected thread issues tcgen05.mma 128 x 256 x 16 over hd / 16 k-steps, accumulator = 256 TMEM columns.
//   softmax  16 warps, four threads per query row (TMEM lane = row; warps w, w + 4, w + 8, w + 12 own the same lanes and take 64 keys each):
//            tcgen05.ld the scores twice (row maximum, then exp2 / row sum -- TMEM reads are cheaper than 256 live registers),
//            unnormalised probabilities go back to shared memory as the bf16 K-major A operand of the second product
//            (16-byte chunk c of row r sits at chunk c ^ (r & 7) of its 128-byte line: the swizzle the MMA descriptor expects).
//   MMA 2    O = P V: 128 x 128 x 16 over 16 k-steps (hd padded to two 64-column atoms; columns >= hd are never read back).
//   epilogue tcgen05.ld O, scale by 1 / row sum, bf16, 16-byte stores of the row's hd columns.
// The mma.sync kernel of attention_mma_long.cu stays for 2 / 3 planes (fp32-parity mode) and other head sizes.
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "attention_mma.cuh"
#include "gemm.cuh"
#include "ptx.cuh"

namespace m2f {
namespace amma {
namespace {

constexpr int TC_QT = 128;     // query rows per tile (TMEM lanes)
constexpr int TC_KEYS = 256;   // key tile = the whole (padded) dialogue
constexpr int TC_SOFTMAX_WARPS = 16;          // four threads per query row, 64 keys each
constexpr int TC_SOFTMAX_THREADS = TC_SOFTMAX_WARPS * 32;
constexpr int TC_OUT_WARPS = 8;               // two threads per query row (alternate 32-column chunks): output tile TMEM -> global
constexpr int TC_OUT_THREADS = TC_OUT_WARPS * 32;
constexpr int TC_MMA_WARP = TC_SOFTMAX_WARPS + TC_OUT_WARPS;
constexpr int TC_THREADS = (TC_MMA_WARP + 1) * 32;
constexpr int TC_BAR_THREADS = TC_SOFTMAX_THREADS + 32;  // named barriers "probabilities ready": softmax warps + the MMA warp
// Q, K and V tiles are made of 32-column atoms (64-byte rows, 64-byte swizzle), hd / 32 of them: exactly the head's columns, so that
// hd = 96 costs 3/4 of what two 64-column atoms would -- the room that lets both query tiles of an item stay resident.
constexpr int TC_QATOM = TC_QT * 64;          // [128 rows][64 B]
constexpr int TC_KATOM = TC_KEYS * 64;        // [256 rows][64 B]
constexpr int TC_VATOM = 64 * 64;             // [64 keys][64 B], per 64-key block
constexpr int TC_SP = 4 * TC_QT * 128;        // probabilities of one query tile: four 64-key atoms of [128 rows][128 B], 128-byte swizzle
constexpr int TC_AUX = 3072;                  // barriers, TMEM slot, key mask, row-maximum exchange (first, so that the tiles stay 1024-aligned)
// aux | K | V | P | query tile 0 | query tile 1 (hd <= 96; with hd = 128 one slot is time-multiplexed).  The base is checked at run time.
__host__ __device__ constexpr int tc_smem_bytes(int hd) {
    return TC_AUX + (hd / 32) * (TC_KATOM + 4 * TC_VATOM) + TC_SP + (hd / 32) * TC_QATOM * (hd <= 96 ? 2 : 1);
}
static_assert(tc_smem_bytes(128) <= 227 * 1024 && tc_smem_bytes(96) <= 227 * 1024, "shared memory budget");
constexpr uint32_t TC_TMEM_COLS = 512;        // two score tiles of 256 columns; O_t reuses columns 0..127 of S_t, the row sums 128..131

struct TcParams {
    const uint8_t* key_pad;
    __nv_bfloat16* o;
    int64_t o_ld;
    int S, H, hd, items;  // items = dialogues x heads
    float scale_log2;     // log2(e) / sqrt(hd)
    unsigned long long* trace;  // debug (m2f_debug_set_attn_trace): %globaltimer stamps of CTA 0's ninth item, else nullptr
};

__device__ __forceinline__ unsigned long long tc_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)::"memory");
    return t;
}
// slot 0..15: MMA / TMA lane, 16..27: softmax thread 0, 28..31: first output thread
#define TC_STAMP(slot)                                                        \
    do {                                                                      \
        if (p.trace && blockIdx.x == 0 && it == 8) p.trace[(slot)] = tc_now(); \
    } while (0)

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void tmem_st_1(uint32_t taddr, uint32_t v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(v) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}

// Persistent: CTA c works on (dialogue, head) items c, c + gridDim.x, ...; both 128-row query tiles of an item share one copy of K, V.
// Three roles:
//   warp 20, lane 0: TMA loads and MMA issue.  Per item: S0 = Q0 K^T; the query slot then takes Q1; S1 = Q1 K^T; once the score
//                    products are complete the key tile and the query slot take the NEXT item's K and Q0; O0 = P0 V when the softmax
//                    warps have written P0, then O1 = P1 V; V of the next item is requested when that is complete.
//   warps 0-15:      softmax(S0) -> P0, softmax(S1) -> P1 (same buffer, once O0 is complete): scores from TMEM, probabilities to
//                    shared memory, the thread's partial row sum to a spare TMEM column of the tile.  Nothing else: they go
//                    straight on to the next item.
//   warps 16-19:     output: wait for O_t, tcgen05.ld it together with the four partial row sums, scale, store 32-byte pieces; then
//                    release the tile's TMEM columns for the next item's scores.
__global__ void __launch_bounds__(TC_THREADS, 1)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk, const __grid_constant__ CUtensorMap tv,
                   const TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int natoms = p.hd / 32;
    const bool two_q = p.hd <= 96;  // both query tiles resident; otherwise tile 1 takes tile 0's slot once S0 is complete
    const int q_bytes = natoms * TC_QATOM, k_bytes = natoms * TC_KATOM, v_bytes = natoms * 4 * TC_VATOM;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint8_t* sK = smem + TC_AUX;
    uint8_t* sV = sK + k_bytes;
    uint8_t* sP = sV + v_bytes;
    uint8_t* sQ = sP + TC_SP;
    uint8_t* sQ1 = two_q ? sQ + q_bytes : sQ;
    uint64_t* bar_q0k = bars;      // Q tile 0 and K landed
    uint64_t* bar_q1 = bars + 1;   // Q tile 1 landed
    uint64_t* bar_v = bars + 2;    // V landed
    uint64_t* bar_s = bars + 3;    // [2] scores of tile t complete in TMEM
    uint64_t* bar_o = bars + 5;    // [2] output of tile t complete in TMEM
    uint64_t* bar_free = bars + 7; // [2] output of tile t has been read: its TMEM columns may take the next item's scores
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);  // (aux: 128 + 16 + 32 + 2048 bytes <= TC_AUX)
    uint32_t* key_ok = tmem_slot + 4;                            // 8 words: bit j of word w = key 32 w + j is usable
    float* red_max = reinterpret_cast<float*>(key_ok + 8);       // [4][128] partial row maxima

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = p.S > TC_QT ? 2 : 1;

    if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) {
        printf("m2f: attn_fwd_tc_kernel needs a 1024-byte aligned shared-memory base\n");
        __trap();
    }
    if (warp == TC_MMA_WARP && lane == 0) {
        ptx::prefetch_tensormap(&tq);
        ptx::prefetch_tensormap(&tk);
        ptx::prefetch_tensormap(&tv);
        for (int i = 0; i < 7; ++i) ptx::mbar_init(bars + i, 1);
        ptx::mbar_init(&bar_free[0], TC_OUT_THREADS);
        ptx::mbar_init(&bar_free[1], TC_OUT_THREADS);
        ptx::fence_barrier_init();
    }
    if (warp == 0) ptx::tmem_alloc(tmem_slot, TC_TMEM_COLS);
    ptx::tcgen05_fence_before();
    __syncthreads();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_trigger();

    if (warp == TC_MMA_WARP) {
        auto load_q0k = [&](int item) {  // K, query tile 0 and, where it has a slot of its own, query tile 1
            const int row0 = (item / p.H) * p.S, col0 = (item % p.H) * p.hd;
            const bool with_q1 = two_q && ntiles == 2;
            ptx::mbar_expect_tx(bar_q0k, (uint32_t)(k_bytes + q_bytes * (with_q1 ? 2 : 1)));
            for (int a = 0; a < natoms; ++a) {
                ptx::tma_load_3d(sK + a * TC_KATOM, &tk, bar_q0k, col0 + a * 32, row0, 0);
                ptx::tma_load_3d(sQ + a * TC_QATOM, &tq, bar_q0k, col0 + a * 32, row0, 0);
                if (with_q1) ptx::tma_load_3d(sQ1 + a * TC_QATOM, &tq, bar_q0k, col0 + a * 32, row0 + TC_QT, 0);
            }
        };
        auto load_q1 = [&](int item) {  // only without a second slot
            const int row0 = (item / p.H) * p.S, col0 = (item % p.H) * p.hd;
            ptx::mbar_expect_tx(bar_q1, (uint32_t)q_bytes);
            for (int a = 0; a < natoms; ++a) ptx::tma_load_3d(sQ1 + a * TC_QATOM, &tq, bar_q1, col0 + a * 32, row0 + TC_QT, 0);
        };
        auto load_v = [&](int item) {
            const int row0 = (item / p.H) * p.S, col0 = (item % p.H) * p.hd;
            ptx::mbar_expect_tx(bar_v, (uint32_t)v_bytes);
            for (int kb = 0; kb < 4; ++kb)
                for (int a = 0; a < natoms; ++a)
                    ptx::tma_load_3d(sV + (kb * natoms + a) * TC_VATOM, &tv, bar_v, col0 + a * 32, row0 + kb * 64, 0);
        };
        if (lane == 0) {
            load_q0k(blockIdx.x);
            load_v(blockIdx.x);
        }
        constexpr uint32_t idesc_qk = ptx::instr_desc_bf16(TC_QT, TC_KEYS, false, false);
        const uint32_t idesc_pv = ptx::instr_desc_bf16(TC_QT, p.hd, false, true);  // N = hd: hd / 32 MN-major atoms
        // The issuing thread is one lane: every instruction between two MMAs is latency on the tensor core's queue, so the descriptors
        // are built once per atom and advanced by adding to their address field (bits 0..13, in 16-byte units).
        auto issue_qk = [&](int t) {  // S_t = Q_t K^T
            ptx::tcgen05_fence_after();
            const uint32_t q_base = ptx::smem_u32(t == 0 ? sQ : sQ1);
            const uint32_t d_tmem = tmem_base + (uint32_t)t * 256u;
            for (int a = 0; a < natoms; ++a) {  // 32-column atom = two k-steps of 16 columns (32 B apart)
                const uint64_t ad = ptx::smem_desc_sw64(q_base + a * TC_QATOM, 0, 512);
                const uint64_t bd = ptx::smem_desc_sw64(ptx::smem_u32(sK) + a * TC_KATOM, 0, 512);
                ptx::umma_bf16(d_tmem, ad, bd, idesc_qk, a > 0 ? 1u : 0u);
                ptx::umma_bf16(d_tmem, ad + 2, bd + 2, idesc_qk, 1u);
            }
            ptx::umma_commit(&bar_s[t]);
        };
        auto issue_pv = [&](int t) {  // O_t = P_t V
            ptx::tcgen05_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)t * 256u;
#pragma unroll
            for (int kb = 0; kb < 4; ++kb) {
                // P: 64-key atom kb, 16 keys = 32 B per k-step.  V: 64-key block kb, atoms TC_VATOM apart (LBO), 8-key groups 512 B
                // apart (SBO), 16 keys = 1024 B per k-step
                const uint64_t ad = ptx::smem_desc_sw128(ptx::smem_u32(sP) + kb * (TC_QT * 128), 0, 1024);
                const uint64_t bd = ptx::smem_desc_sw64(ptx::smem_u32(sV) + kb * natoms * TC_VATOM, TC_VATOM, 512);
                ptx::umma_bf16(d_tmem, ad, bd, idesc_pv, kb > 0 ? 1u : 0u);
                ptx::umma_bf16(d_tmem, ad + 2, bd + 64, idesc_pv, 1u);
                ptx::umma_bf16(d_tmem, ad + 4, bd + 128, idesc_pv, 1u);
                ptx::umma_bf16(d_tmem, ad + 6, bd + 192, idesc_pv, 1u);
            }
            ptx::umma_commit(&bar_o[t]);
        };
        int it = 0;
        bool qk0_issued = false;  // S0 of the coming item was already issued ahead of the previous item's last output product
        for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
            const uint32_t ph = (uint32_t)(it & 1);
            const int next = item + (int)gridDim.x;
            const bool has_next = next < p.items;
            if (lane == 0) {
                TC_STAMP(0);
                if (!qk0_issued) {
                    ptx::mbar_wait(bar_q0k, ph);
                    TC_STAMP(1);
                    if (it > 0) ptx::mbar_wait(&bar_free[0], ph ^ 1u);  // the previous item's output tile 0 has left TMEM
                    TC_STAMP(2);
                    issue_qk(0);
                    TC_STAMP(3);
                }
                if (ntiles == 2) {
                    if (!two_q) {  // the query slot is free once S0 is complete: it takes the second query tile
                        ptx::mbar_wait(&bar_s[0], ph);
                        load_q1(item);
                        ptx::mbar_wait(bar_q1, ph);
                    }
                    TC_STAMP(4);
                    if (it > 0) ptx::mbar_wait(&bar_free[1], ph ^ 1u);  // the previous item's output tile 1 has left TMEM
                    TC_STAMP(5);
                    issue_qk(1);
                    TC_STAMP(6);
                }
            }
            __syncwarp();
            // (K and the query tiles of the next item are requested by an output thread as soon as the score products are complete)
            // the probabilities of tile 0 are in shared memory once all softmax threads have passed this named barrier
            ptx::named_bar_sync(2, TC_BAR_THREADS);
            if (lane == 0) {
                TC_STAMP(8);
                ptx::mbar_wait(bar_v, ph);
                issue_pv(0);
                TC_STAMP(9);
            }
            __syncwarp();
            qk0_issued = false;
            if (ntiles == 2) {
                ptx::named_bar_sync(3, TC_BAR_THREADS);
                if (lane == 0) {
                    TC_STAMP(10);
                    if (has_next) {
                        // the next item's first score product goes AHEAD of this item's second output product: the softmax warps are
                        // about to wait for it, the output warps have no deadline.  Its operands were requested long ago and output
                        // tile 0 of this item was stored while the softmax warps worked on tile 1.
                        ptx::mbar_wait(bar_q0k, ph ^ 1u);
                        ptx::mbar_wait(&bar_free[0], ph);
                        issue_qk(0);
                    }
                    issue_pv(1);
                    TC_STAMP(11);
                }
                qk0_issued = has_next;
                __syncwarp();
            }
        }
    } else if (warp < TC_SOFTMAX_WARPS) {
        // -------------------------------------------------------------------- softmax: thread = (query row, quarter of the keys)
        const int lg = warp & 3, quarter = warp >> 2;
        const int row = lg * 32 + lane;
        const uint32_t sw = (uint32_t)(row & 7);
        // keys quarter * 64 .. + 63 = atom `quarter` of the probability buffer
        const uint32_t atom = ptx::smem_u32(sP) + (uint32_t)quarter * (TC_QT * 128) + (uint32_t)row * 128u;
        int it = 0;
        for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
            const uint32_t ph = (uint32_t)(it & 1);
            const int row0 = (item / p.H) * p.S;
            if (warp < 8) {  // usable keys: inside the dialogue and not padded; warp w covers keys 32 w .. 32 w + 31
                const int j = threadIdx.x;
                const bool ok = j < p.S && p.key_pad[(int64_t)row0 + j] == 0;
                const uint32_t word = __ballot_sync(0xffffffffu, ok);
                if (lane == 0) key_ok[warp] = word;
            }
            ptx::named_bar_sync(1, TC_SOFTMAX_THREADS);
            if (threadIdx.x == 0) TC_STAMP(16);
            const uint32_t w0 = key_ok[quarter * 2], w1 = key_ok[quarter * 2 + 1];
            const bool all_ok = (w0 & w1) == 0xffffffffu;  // warp-uniform: no masking work for fully valid key blocks
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                if (t < ntiles) {
                    ptx::mbar_wait(&bar_s[t], ph);
                    ptx::tcgen05_fence_after();
                    if (threadIdx.x == 0) TC_STAMP(17 + 5 * t);
                    const uint32_t t_tile = tmem_base + (uint32_t)t * 256u + ((uint32_t)(lg * 32) << 16);
                    const uint32_t t_row = t_tile + (uint32_t)(quarter * 64);
                    float mx = -INFINITY;
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32(t_row + (uint32_t)(c * 32), r);
                        ptx::tmem_ld_wait();
                        if (all_ok) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(r[j]));
                        } else {
                            const uint32_t w = c == 0 ? w0 : w1;
#pragma unroll
                            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, ((w >> j) & 1u) ? __uint_as_float(r[j]) : -INFINITY);
                        }
                    }
                    red_max[quarter * TC_QT + row] = mx;
                    ptx::named_bar_sync(1, TC_SOFTMAX_THREADS);
                    mx = fmaxf(fmaxf(red_max[row], red_max[TC_QT + row]), fmaxf(red_max[2 * TC_QT + row], red_max[3 * TC_QT + row]));
                    if (mx == -INFINITY) mx = 0.f;  // no usable key at all: every probability is 0 (such rows are never stored)
                    const float mscaled = mx * p.scale_log2;
                    ptx::named_bar_sync(1, TC_SOFTMAX_THREADS);  // every partial maximum has been read: the array is free for the next tile
                    if (threadIdx.x == 0) TC_STAMP(18 + 5 * t);
                    // the probability buffer is free once the product that read it last is complete
                    if (t == 1) ptx::mbar_wait(&bar_o[0], ph);
                    else if (it > 0) ptx::mbar_wait(&bar_o[ntiles - 1], ph ^ 1u);
                    if (threadIdx.x == 0) TC_STAMP(19 + 5 * t);
                    float sum = 0.f;
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32(t_row + (uint32_t)(c * 32), r);
                        ptx::tmem_ld_wait();
                        uint32_t pk[16];
                        if (all_ok) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const float e0 = ex2(fmaf(__uint_as_float(r[2 * j]), p.scale_log2, -mscaled));
                                const float e1 = ex2(fmaf(__uint_as_float(r[2 * j + 1]), p.scale_log2, -mscaled));
                                sum += e0 + e1;
                                pk[j] = pack2(e0, e1);
                            }
                        } else {
                            const uint32_t w = c == 0 ? w0 : w1;
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const float e0 = ((w >> (2 * j)) & 1u) ? ex2(fmaf(__uint_as_float(r[2 * j]), p.scale_log2, -mscaled)) : 0.f;
                                const float e1 = ((w >> (2 * j + 1)) & 1u) ? ex2(fmaf(__uint_as_float(r[2 * j + 1]), p.scale_log2, -mscaled)) : 0.f;
                                sum += e0 + e1;
                                pk[j] = pack2(e0, e1);
                            }
                        }
                        // 32 keys = 16-byte chunks c * 4 .. + 3 of the row's 128-byte line
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint32_t chunk = ((uint32_t)(c * 4 + q)) ^ sw;
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(atom + (chunk << 4)), "r"(pk[4 * q]), "r"(pk[4 * q + 1]),
                                         "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                                         : "memory");
                        }
                    }
                    // this thread's partial row sum -> column 128 + quarter of the tile (the output product writes columns 0..127 only);
                    // those columns hold scores of the third quarter: every softmax thread must have finished reading first
                    ptx::named_bar_sync(1, TC_SOFTMAX_THREADS);
                    tmem_st_1(t_tile + 128u + (uint32_t)quarter, __float_as_uint(sum));
                    tmem_st_wait();
                    if (threadIdx.x == 0) TC_STAMP(20 + 5 * t);
                    ptx::fence_proxy_async();      // the probabilities are read by the tensor core through the async proxy
                    ptx::tcgen05_fence_before();   // ... and this tile's score columns are about to be overwritten by its output
                    ptx::named_bar_arrive(2 + t, TC_BAR_THREADS);  // releases the MMA warp (which syncs on it); no waiting on this side
                    if (threadIdx.x == 0) TC_STAMP(21 + 5 * t);
                }
            }
        }
    } else {
        // -------------------------------------------------------------------- output: thread = query row
        const int lg = warp & 3, part = (warp - TC_SOFTMAX_WARPS) >> 2;  // part 0: chunks 0, 2; part 1: chunks 1, 3
        const int row = lg * 32 + lane;
        int it = 0;
        for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
            const uint32_t ph = (uint32_t)(it & 1);
            const int row0 = (item / p.H) * p.S, col0 = (item % p.H) * p.hd;
            const int next = item + (int)gridDim.x;
            if (threadIdx.x == TC_SOFTMAX_THREADS + 5 * 32 && next < p.items) {  // (a warp of the half with fewer chunks)
                // K and the query slots are free once the item's score products are complete: the second output warp's first thread
                // requests the next item's (nine TMA instructions that the MMA lane does not have to spend 0.8 us on)
                ptx::mbar_wait(&bar_s[ntiles - 1], ph);
                TC_STAMP(7);
                const int nrow0 = (next / p.H) * p.S, ncol0 = (next % p.H) * p.hd;
                const bool with_q1 = two_q && ntiles == 2;
                ptx::mbar_expect_tx(bar_q0k, (uint32_t)(k_bytes + q_bytes * (with_q1 ? 2 : 1)));
                for (int a = 0; a < natoms; ++a) {
                    ptx::tma_load_3d(sK + a * TC_KATOM, &tk, bar_q0k, ncol0 + a * 32, nrow0, 0);
                    ptx::tma_load_3d(sQ + a * TC_QATOM, &tq, bar_q0k, ncol0 + a * 32, nrow0, 0);
                    if (with_q1) ptx::tma_load_3d(sQ1 + a * TC_QATOM, &tq, bar_q0k, ncol0 + a * 32, nrow0 + TC_QT, 0);
                }
            }
            __syncwarp();
            for (int t = 0; t < ntiles; ++t) {
                ptx::mbar_wait(&bar_o[t], ph);
                ptx::tcgen05_fence_after();
                if (threadIdx.x == TC_SOFTMAX_THREADS) TC_STAMP(28 + 2 * t);
                if (threadIdx.x == TC_SOFTMAX_THREADS + 4 * 32) {
                    if (t == ntiles - 1 && next < p.items) {  // V is free once the item's last output product is complete: the first
                        // output thread requests the next item's V (the MMA lane does not have to come back for it)
                        const int nrow0 = (next / p.H) * p.S, ncol0 = (next % p.H) * p.hd;
                        ptx::mbar_expect_tx(bar_v, (uint32_t)v_bytes);
                        for (int kb = 0; kb < 4; ++kb)
                            for (int a = 0; a < natoms; ++a)
                                ptx::tma_load_3d(sV + (kb * natoms + a) * TC_VATOM, &tv, bar_v, ncol0 + a * 32, nrow0 + kb * 64, 0);
                    }
                }
                const uint32_t t_tile = tmem_base + (uint32_t)t * 256u + ((uint32_t)(lg * 32) << 16);
                uint32_t sums[4];
                tmem_ld_4(t_tile + 128u, sums);
                ptx::tmem_ld_wait();
                const float iv = 1.f / ((__uint_as_float(sums[0]) + __uint_as_float(sums[1])) + (__uint_as_float(sums[2]) + __uint_as_float(sums[3])));
                const int qrow = t * TC_QT + row;
                __nv_bfloat16* orow = p.o + ((int64_t)row0 + qrow) * p.o_ld + col0;
                for (int cc = part * 32; cc < p.hd; cc += 64) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32(t_tile + (uint32_t)cc, r);
                    ptx::tmem_ld_wait();
                    if (qrow < p.S) {
#pragma unroll
                        for (int q = 0; q < 2; ++q) {  // 2 x 32 bytes: whole sectors (16-byte stores cost a transaction per half sector)
                            uint32_t w[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                w[i] = pack2(__uint_as_float(r[16 * q + 2 * i]) * iv, __uint_as_float(r[16 * q + 2 * i + 1]) * iv);
                            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(orow + cc + 16 * q), "r"(w[0]), "r"(w[1]),
                                         "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                                         : "memory");
                        }
                    }
                }
                ptx::tcgen05_fence_before();
                ptx::mbar_arrive(&bar_free[t]);  // this thread has read its row of output tile t (and its sums)
                if (threadIdx.x == TC_SOFTMAX_THREADS) TC_STAMP(29 + 2 * t);
            }
        }
    }
    ptx::tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) {
        ptx::tcgen05_fence_after();
        ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
    }
}

}  // namespace

// hd in {64, 96, 128}; one plane; enough rows for the 256-row key box; M2F_ATT_TC=0 keeps the mma.sync kernel (A/B, debugging)
static unsigned long long* g_tc_trace = nullptr;

bool tc_supported(int hd, int64_t S, int planes, int64_t rows) {
    static const bool off = [] { const char* e = getenv("M2F_ATT_TC"); return e && e[0] == '0'; }();
    return !off && planes == 1 && (hd == 64 || hd == 96 || hd == 128) && S > 64 && S <= TC_KEYS && rows >= TC_KEYS;
}

int fwd_long_tc(int hd, const m2f_planes_t* q, const m2f_planes_t* k, const m2f_planes_t* v, const uint8_t* pad, int64_t B, int S, int H,
                const OutView& o, cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    M2F_TRY(make_planes_tmap_sw64(&tq, q, TC_QT));
    M2F_TRY(make_planes_tmap_sw64(&tk, k, TC_KEYS));
    M2F_TRY(make_planes_tmap_sw64(&tv, v, 64));
    static bool attr_set = false;
    if (!attr_set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes(128)));
        attr_set = true;
    }
    TcParams p;
    p.key_pad = pad;
    p.o = o.ptr;
    p.o_ld = o.ld;
    p.S = S;
    p.H = H;
    p.hd = hd;
    p.items = (int)(B * H);
    p.scale_log2 = 1.4426950408889634f / sqrtf((float)hd);
    p.trace = g_tc_trace;
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        M2F_CUDA(cudaGetDevice(&dev));
        M2F_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const unsigned grid = (unsigned)std::min<int64_t>(B * H, sms);  // persistent: one CTA per SM walks the (dialogue, head) items
    M2F_CUDA(launch_chain(attn_fwd_tc_kernel, dim3(grid), dim3(TC_THREADS), (size_t)tc_smem_bytes(hd), st, 1, tq, tk, tv, p));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

}  // namespace amma
}  // namespace m2f

// profiling aid (tools/atc_trace.py): 32 x u64 of %globaltimer stamps written by CTA 0 during its second item; nullptr turns it off
extern "C" void m2f_debug_set_attn_trace(unsigned long long* buf) { m2f::amma::g_tc_trace = buf; }
