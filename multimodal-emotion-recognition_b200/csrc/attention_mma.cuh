// Fast path of K2 for short dialogues: S <= 64 utterances, head_dim in {16, 32, 64, 96, 128}.
// One CTA per (dialogue, head); Q, K, V (and dO) of that head live in shared memory as bf16 planes; the S x S
// score / probability matrices never leave registers (forward) or shared memory (backward).  The small matrix
// products run on the warp-level tensor path (mma.sync m16n8k16, bf16 in, fp32 accumulate): a 33 x 33 x 96 problem
// is far below a tcgen05 tile (M = 64/128 rows + TMEM round trip) and the kernel is bound by its global loads/stores.
// fp32-parity mode: every product runs over the significant bf16 plane pairs (hi/mid/lo), like the GEMM.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace m2f {
namespace amma {

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D += A(16x16, row) * B(16x8, col)
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float x, float y) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    return *reinterpret_cast<const uint32_t*>(&h);
}
// bf16x2 plane `pl` of the pair (x, y): hi, then the residuals
__device__ __forceinline__ void split2(float x, float y, int P, uint32_t (&out)[3]) {
    __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    out[0] = *reinterpret_cast<const uint32_t*>(&h);
    if (P > 1) {
        x -= __low2float(h); y -= __high2float(h);
        h = __floats2bfloat162_rn(x, y);
        out[1] = *reinterpret_cast<const uint32_t*>(&h);
        if (P > 2) {
            x -= __low2float(h); y -= __high2float(h);
            h = __floats2bfloat162_rn(x, y);
            out[2] = *reinterpret_cast<const uint32_t*>(&h);
        }
    }
}

// (x, y) pairs of two rows -> bf16x2 planes at d (row A) and d + row_step (row B); planes are plane_step apart
__device__ __forceinline__ void store_pair_planes(__nv_bfloat16* d, int row_step, int plane_step, int P, float x0, float y0, float x1, float y1) {
    uint32_t h0[3], h1[3];
    split2(x0, y0, P, h0);
    split2(x1, y1, P, h1);
    *reinterpret_cast<uint32_t*>(d) = h0[0];
    *reinterpret_cast<uint32_t*>(d + row_step) = h1[0];
    if (P > 1) {
        *reinterpret_cast<uint32_t*>(d + plane_step) = h0[1];
        *reinterpret_cast<uint32_t*>(d + plane_step + row_step) = h1[1];
        if (P > 2) {
            *reinterpret_cast<uint32_t*>(d + 2 * plane_step) = h0[2];
            *reinterpret_cast<uint32_t*>(d + 2 * plane_step + row_step) = h1[2];
        }
    }
}

// plane products (a-plane, b-plane), least significant first
__device__ __forceinline__ int nprod(int P) { return P == 1 ? 1 : (P == 2 ? 3 : 6); }
__device__ __forceinline__ void prod(int P, int i, int& pa, int& pb) {
    if (P == 1) { pa = 0; pb = 0; return; }
    if (P == 2) { pa = i == 0 ? 1 : 0; pb = i == 1 ? 1 : 0; return; }
    pa = (0x046 >> (2 * i)) & 3;  // 2,1,0,1,0,0
    pb = (0x124 >> (2 * i)) & 3;  // 0,1,2,0,1,0
}

struct View {  // bf16 planes view of [rows, H*hd]
    const __nv_bfloat16* ptr;
    int64_t ld, stride;
};
struct OutView {
    __nv_bfloat16* ptr;
    int64_t ld, stride;
};

#define AMMA_FWD_ARGS                                                                                                         \
    int hd, int SPv, const ::m2f::amma::View &q, const ::m2f::amma::View &k, const ::m2f::amma::View &v, const uint8_t *pad, \
        int64_t BH, int S, int H, float scale, const ::m2f::Drop &drop, const ::m2f::amma::OutView &o, cudaStream_t st
#define AMMA_BWD_ARGS                                                                                                            \
    int hd, int SPv, const ::m2f::amma::View &q, const ::m2f::amma::View &k, const ::m2f::amma::View &v,                        \
        const ::m2f::amma::View &d_o, const uint8_t *pad, int64_t BH, int S, int H, float scale, const ::m2f::Drop &drop,       \
        const ::m2f::amma::OutView &dq, const ::m2f::amma::OutView &dk, const ::m2f::amma::OutView &dv, float *dbq, float *dbk, \
        float *dbv, cudaStream_t st

// global [S, HD] head slice (all planes) -> shared [P][SP][HD + 8], rows >= S zeroed.  Asynchronous 16-byte copies
// (cp.async, zero-fill for the padding rows): every copy of the CTA is in flight at once; the caller waits with
// cp_async_wait_all() + __syncthreads().
template <int HD, int SP>
__device__ __forceinline__ void load_tile(__nv_bfloat16* dst, const View& t, int P, int64_t row0, int col0, int S) {
    constexpr int LDS = HD + 8, CH = HD / 8;
    if (blockDim.x % CH == 0) {
        // thread = (row of the pass, 16-byte chunk of the row): one division up front, then pointer increments -- these kernels are
        // instruction-bound and the index arithmetic of the general loop below was a fifth of their instructions
        const int rpp = blockDim.x / CH;  // rows per pass
        const int r0 = threadIdx.x / CH, c = (threadIdx.x - r0 * CH) * 8;
        const __nv_bfloat16* base = t.ptr + row0 * t.ld + col0 + c;
        for (int pl = 0; pl < P; ++pl) {
            const __nv_bfloat16* src = base + pl * t.stride + (int64_t)r0 * t.ld;
            uint32_t d = ptx::smem_u32(dst + (pl * SP + r0) * LDS + c);
            for (int r = r0; r < SP; r += rpp) {
                const bool live = r < S;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(live ? src : base), "r"(live ? 16 : 0) : "memory");
                src += (int64_t)rpp * t.ld;
                d += (uint32_t)(rpp * LDS * 2);
            }
        }
        return;
    }
    for (int i = threadIdx.x; i < P * SP * CH; i += blockDim.x) {
        const int pl = i / (SP * CH), rem = i - pl * (SP * CH);
        const int r = rem / CH, c = (rem - r * CH) * 8;
        const int rr = r < S ? r : 0;  // keep the address valid; src-size 0 writes zeros
        const __nv_bfloat16* src = t.ptr + pl * t.stride + (row0 + rr) * t.ld + col0 + c;
        const uint32_t d = ptx::smem_u32(dst + (pl * SP + r) * LDS + c);
        const int nbytes = r < S ? 16 : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(nbytes) : "memory");
    }
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// acc[nt][4] (16 x SP strip) += A(16 x HD from sA planes, rows i0..) * B^T (sB planes stored [n][k], k contiguous)
template <int HD, int SP>
__device__ __forceinline__ void strip_nt(float (&acc)[SP / 8][4], const __nv_bfloat16* sA, const __nv_bfloat16* sB, int P, int i0, int lane) {
    constexpr int LDS = HD + 8;
    const int a_row = i0 + (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 8;
    const int b_row = (lane & 7) + (lane >> 4) * 8, b_col = ((lane >> 3) & 1) * 8;
    const int np = nprod(P);
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
#pragma unroll 1
        for (int pi = 0; pi < np; ++pi) {
            int pa, pb;
            prod(P, pi, pa, pb);
            uint32_t a[4];
            ldsm_x4(ptx::smem_u32(sA + (pa * SP + a_row) * LDS + kk * 16 + a_col), a);
#pragma unroll
            for (int n2 = 0; n2 < SP / 16; ++n2) {
                uint32_t b[4];
                ldsm_x4(ptx::smem_u32(sB + (pb * SP + n2 * 16 + b_row) * LDS + kk * 16 + b_col), b);
                mma16816(acc[2 * n2], a, b[0], b[1]);
                mma16816(acc[2 * n2 + 1], a, b[2], b[3]);
            }
        }
    }
}

// A fragments (mma.sync m16n8k16, row-major A) of a 16-row strip read straight from global memory: qf[pl][kk] holds, for the 16 columns
// of k-step kk, {row g cols 2q..2q+1, row g+8 same, row g cols 8+2q.., row g+8 same} of plane pl -- 4-byte loads, the four lanes of a
// quad cover 16 contiguous bytes of a row.  Rows >= S read as zero.  Used where a tile is the A operand of one product only (Q in the
// forward): it then never occupies shared memory.
template <int HD, int PMAX>
__device__ __forceinline__ void load_a_frags(uint32_t (&qf)[PMAX][HD / 16][4], const View& t, int P, int64_t row0, int col0, int i0, int S, int lane) {
    const int g = lane >> 2, qd = lane & 3;
    const bool live0 = i0 + g < S, live1 = i0 + g + 8 < S;
#pragma unroll
    for (int pl = 0; pl < PMAX; ++pl) {
        if (pl < P) {
            const __nv_bfloat16* r0 = t.ptr + pl * t.stride + (row0 + i0 + g) * t.ld + col0 + 2 * qd;
            const __nv_bfloat16* r1 = r0 + 8 * t.ld;
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk) {
                qf[pl][kk][0] = live0 ? __ldg(reinterpret_cast<const uint32_t*>(r0 + kk * 16)) : 0u;
                qf[pl][kk][1] = live1 ? __ldg(reinterpret_cast<const uint32_t*>(r1 + kk * 16)) : 0u;
                qf[pl][kk][2] = live0 ? __ldg(reinterpret_cast<const uint32_t*>(r0 + kk * 16 + 8)) : 0u;
                qf[pl][kk][3] = live1 ? __ldg(reinterpret_cast<const uint32_t*>(r1 + kk * 16 + 8)) : 0u;
            }
        }
    }
}

// acc[nt][4] (16 x SP strip) += A (register fragments of load_a_frags) * B^T (sB planes stored [n][k], k contiguous)
template <int HD, int SP, int PMAX>
__device__ __forceinline__ void strip_nt_regA(float (&acc)[SP / 8][4], const uint32_t (&qf)[PMAX][HD / 16][4], const __nv_bfloat16* sB, int lane) {
    constexpr int LDS = HD + 8;
    const int b_row = (lane & 7) + (lane >> 4) * 8, b_col = ((lane >> 3) & 1) * 8;
    constexpr int NP = PMAX == 1 ? 1 : (PMAX == 2 ? 3 : 6);
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
#pragma unroll
        for (int pi = 0; pi < NP; ++pi) {
            // plane products, least significant first (prod())
            constexpr int PA2[3] = {1, 0, 0}, PB2[3] = {0, 1, 0};
            constexpr int PA3[6] = {2, 1, 0, 1, 0, 0}, PB3[6] = {0, 1, 2, 0, 1, 0};
            const int pa = PMAX == 1 ? 0 : (PMAX == 2 ? PA2[pi % 3] : PA3[pi]), pb = PMAX == 1 ? 0 : (PMAX == 2 ? PB2[pi % 3] : PB3[pi]);
#pragma unroll
            for (int n2 = 0; n2 < SP / 16; ++n2) {
                uint32_t b[4];
                ldsm_x4(ptx::smem_u32(sB + (pb * SP + n2 * 16 + b_row) * LDS + kk * 16 + b_col), b);
                mma16816(acc[2 * n2], qf[pa][kk], b[0], b[1]);
                mma16816(acc[2 * n2 + 1], qf[pa][kk], b[2], b[3]);
            }
        }
    }
}

// out[nt][4] (16 x HD strip) += A(16 x SP, fp32 values in accumulator layout `pv`, split into planes on the fly)
//                               * B (sB planes stored [k][n], n contiguous: ldmatrix.trans)
template <int HD, int SP>
__device__ __forceinline__ void strip_regA_nn(float (&out)[HD / 8][4], const float (&pv)[SP / 8][4], const __nv_bfloat16* sB, int P, int lane) {
    constexpr int LDS = HD + 8;
    const int b_row = (lane & 7) + ((lane >> 3) & 1) * 8, b_col = (lane >> 4) * 8;
    const int np = nprod(P);
#pragma unroll
    for (int kt = 0; kt < SP / 16; ++kt) {
        uint32_t ap[4][3];
        split2(pv[2 * kt][0], pv[2 * kt][1], P, ap[0]);
        split2(pv[2 * kt][2], pv[2 * kt][3], P, ap[1]);
        split2(pv[2 * kt + 1][0], pv[2 * kt + 1][1], P, ap[2]);
        split2(pv[2 * kt + 1][2], pv[2 * kt + 1][3], P, ap[3]);
#pragma unroll 1
        for (int pi = 0; pi < np; ++pi) {
            int pa, pb;
            prod(P, pi, pa, pb);
            uint32_t a[4];
            a[0] = pa == 0 ? ap[0][0] : (pa == 1 ? ap[0][1] : ap[0][2]);
            a[1] = pa == 0 ? ap[1][0] : (pa == 1 ? ap[1][1] : ap[1][2]);
            a[2] = pa == 0 ? ap[2][0] : (pa == 1 ? ap[2][1] : ap[2][2]);
            a[3] = pa == 0 ? ap[3][0] : (pa == 1 ? ap[3][1] : ap[3][2]);
#pragma unroll
            for (int n2 = 0; n2 < HD / 16; ++n2) {
                uint32_t b[4];
                ldsm_x4_t(ptx::smem_u32(sB + (pb * SP + kt * 16 + b_row) * LDS + n2 * 16 + b_col), b);
                mma16816(out[2 * n2], a, b[0], b[1]);
                mma16816(out[2 * n2 + 1], a, b[2], b[3]);
            }
        }
    }
}

// The same with the plane count known at compile time: the product loop unrolls and the plane selects fold away.
template <int HD, int SP, int PMAX>
__device__ __forceinline__ void strip_regA_nn_ct(float (&out)[HD / 8][4], const float (&pv)[SP / 8][4], const __nv_bfloat16* sB, int lane) {
    constexpr int LDS = HD + 8;
    const int b_row = (lane & 7) + ((lane >> 3) & 1) * 8, b_col = (lane >> 4) * 8;
    constexpr int NP = PMAX == 1 ? 1 : (PMAX == 2 ? 3 : 6);
#pragma unroll
    for (int kt = 0; kt < SP / 16; ++kt) {
        uint32_t ap[4][3];
        split2(pv[2 * kt][0], pv[2 * kt][1], PMAX, ap[0]);
        split2(pv[2 * kt][2], pv[2 * kt][3], PMAX, ap[1]);
        split2(pv[2 * kt + 1][0], pv[2 * kt + 1][1], PMAX, ap[2]);
        split2(pv[2 * kt + 1][2], pv[2 * kt + 1][3], PMAX, ap[3]);
#pragma unroll
        for (int pi = 0; pi < NP; ++pi) {
            constexpr int PA2[3] = {1, 0, 0}, PB2[3] = {0, 1, 0};
            constexpr int PA3[6] = {2, 1, 0, 1, 0, 0}, PB3[6] = {0, 1, 2, 0, 1, 0};
            const int pa = PMAX == 1 ? 0 : (PMAX == 2 ? PA2[pi % 3] : PA3[pi]), pb = PMAX == 1 ? 0 : (PMAX == 2 ? PB2[pi % 3] : PB3[pi]);
            const uint32_t a[4] = {ap[0][pa], ap[1][pa], ap[2][pa], ap[3][pa]};
#pragma unroll
            for (int n2 = 0; n2 < HD / 16; ++n2) {
                uint32_t b[4];
                ldsm_x4_t(ptx::smem_u32(sB + (pb * SP + kt * 16 + b_row) * LDS + n2 * 16 + b_col), b);
                mma16816(out[2 * n2], a, b[0], b[1]);
                mma16816(out[2 * n2 + 1], a, b[2], b[3]);
            }
        }
    }
}

// out[nt][4] (16 x HD strip, rows i0..) += A (sA planes stored [m][k], leading dim SP + 8) * B (sB planes stored [k][n])
template <int HD, int SP>
__device__ __forceinline__ void strip_smemA_nn(float (&out)[HD / 8][4], const __nv_bfloat16* sA, const __nv_bfloat16* sB, int P, int i0, int lane) {
    constexpr int LDA = SP + 8, LDB = HD + 8;
    const int a_row = i0 + (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 8;
    const int b_row = (lane & 7) + ((lane >> 3) & 1) * 8, b_col = (lane >> 4) * 8;
    const int np = nprod(P);
#pragma unroll
    for (int kt = 0; kt < SP / 16; ++kt) {
#pragma unroll 1
        for (int pi = 0; pi < np; ++pi) {
            int pa, pb;
            prod(P, pi, pa, pb);
            uint32_t a[4];
            ldsm_x4(ptx::smem_u32(sA + (pa * SP + a_row) * LDA + kt * 16 + a_col), a);
#pragma unroll
            for (int n2 = 0; n2 < HD / 16; ++n2) {
                uint32_t b[4];
                ldsm_x4_t(ptx::smem_u32(sB + (pb * SP + kt * 16 + b_row) * LDB + n2 * 16 + b_col), b);
                mma16816(out[2 * n2], a, b[0], b[1]);
                mma16816(out[2 * n2 + 1], a, b[2], b[3]);
            }
        }
    }
}

// out[nt][4] (16 x HD strip, rows = keys j0..) += A^T * B with A stored [k][m] in sA (planes, leading dim LDA,
// rows k = queries) and B stored [k][n] in sB (planes, leading dim HD + 8): both through ldmatrix.trans
template <int HD, int SP>
__device__ __forceinline__ void strip_tn(float (&out)[HD / 8][4], const __nv_bfloat16* sA, const __nv_bfloat16* sB, int P, int j0, int lane) {
    constexpr int LDA = SP + 8, LDB = HD + 8;
    const int a_row = (lane & 7) + (lane >> 4) * 8, a_col = j0 + ((lane >> 3) & 1) * 8;
    const int b_row = (lane & 7) + ((lane >> 3) & 1) * 8, b_col = (lane >> 4) * 8;
    const int np = nprod(P);
#pragma unroll
    for (int kt = 0; kt < SP / 16; ++kt) {
#pragma unroll 1
        for (int pi = 0; pi < np; ++pi) {
            int pa, pb;
            prod(P, pi, pa, pb);
            uint32_t a[4];
            ldsm_x4_t(ptx::smem_u32(sA + (pa * SP + kt * 16 + a_row) * LDA + a_col), a);
#pragma unroll
            for (int n2 = 0; n2 < HD / 16; ++n2) {
                uint32_t b[4];
                ldsm_x4_t(ptx::smem_u32(sB + (pb * SP + kt * 16 + b_row) * LDB + n2 * 16 + b_col), b);
                mma16816(out[2 * n2], a, b[0], b[1]);
                mma16816(out[2 * n2 + 1], a, b[2], b[3]);
            }
        }
    }
}

// 64-bit mask of usable keys (bit j set: j < S and not padded), built by one warp with two ballots
__device__ __forceinline__ uint64_t key_valid_bits(const uint8_t* pad, int S, int lane) {
    const bool v0 = lane < S && !pad[lane];
    const bool v1 = lane + 32 < S && !pad[lane + 32];
    const uint32_t lo = __ballot_sync(0xffffffffu, v0), hi = __ballot_sync(0xffffffffu, v1);
    return ((uint64_t)hi << 32) | lo;
}

// Dropout keep-bits of this thread's probabilities.  Element (query i, key j) of (dialogue, head) `bh` is element
// (bh*S + i)*S8 + j of the site (S8 = S rounded up to 8), so an 8-wide tile row is exactly one Philox call (eight 16-bit
// lanes) and the two columns 2q, 2q+1 a thread owns are the two halves of word q.
// Bit (nt*2 + rowsel)*2 + c  <->  accumulator element [nt][rowsel*2 + c].  Rolled: one Philox body.
// `j8_0`: first 8-key tile of the strip (key tile offset / 8) for the kernels that walk the keys in tiles.
template <int SP>
__device__ __forceinline__ uint32_t drop_bits(const Drop& d, uint64_t bh, int S, int i0, int lane, int j8_0 = 0) {
    const int g = lane >> 2, q = lane & 3;
    const uint64_t s8o = (uint64_t)((S + 7) >> 3);  // S8 / 8
    uint32_t bits = 0;
    // The four lanes of a quad own the four words of the same calls (same rows, same 8-key tiles): each lane makes one call out of
    // four and a 4 x 4 transpose inside the quad (two butterfly stages, four shuffles) hands every lane its word of each call.
#pragma unroll 1
    for (int m = 0; m < (SP / 8) / 2; ++m) {
        const int e_own = 4 * m + q;
        uint32_t a[4];
        philox4(d.ph, (bh * S + (i0 + g + (e_own & 1) * 8)) * s8o + (uint64_t)(j8_0 + (e_own >> 1)), a);
        {   // stage 1: partner lane ^ 1 -- even lanes hand over words 1, 3 and receive the partner's words 0, 2 into them
            const bool odd = (q & 1) != 0;
            const uint32_t x = __shfl_xor_sync(0xffffffffu, odd ? a[0] : a[1], 1), y = __shfl_xor_sync(0xffffffffu, odd ? a[2] : a[3], 1);
            if (odd) { a[0] = x; a[2] = y; } else { a[1] = x; a[3] = y; }
        }
        {   // stage 2: partner lane ^ 2 -- lanes 0, 1 hand over words 2, 3 and receive the partner's words 0, 1 into them
            const bool hi = (q & 2) != 0;
            const uint32_t x = __shfl_xor_sync(0xffffffffu, hi ? a[0] : a[2], 2), y = __shfl_xor_sync(0xffffffffu, hi ? a[1] : a[3], 2);
            if (hi) { a[0] = x; a[1] = y; } else { a[2] = x; a[3] = y; }
        }
        // a[j] is now word q of call 4 m + j
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t w = a[j];
            bits |= (((w & 0xFFFFu) >= d.thresh ? 1u : 0u) | ((w >> 16) >= d.thresh ? 2u : 0u)) << (2 * (4 * m + j));
        }
    }
    return bits;
}
__device__ __forceinline__ float drop_mul(uint32_t bits, int nt, int c, float scale) {
    return ((bits >> ((nt * 2 + (c >> 1)) * 2 + (c & 1))) & 1u) ? scale : 0.f;
}

// Accumulator strip (16 rows x HD) -> bf16 planes in this warp's own 16 rows of a shared tile [P][SP][HD + 8]
template <int HD, int SP>
__device__ __forceinline__ void stage_strip(const float (&acc)[HD / 8][4], float mul, __nv_bfloat16* tile, int P, int i0, int lane) {
    constexpr int LDS = HD + 8;
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) {
        uint32_t h0[3], h1[3];
        split2(acc[nt][0] * mul, acc[nt][1] * mul, P, h0);
        split2(acc[nt][2] * mul, acc[nt][3] * mul, P, h1);
        __nv_bfloat16* d0 = tile + (i0 + g) * LDS + nt * 8 + 2 * q;
        *reinterpret_cast<uint32_t*>(d0) = h0[0];
        *reinterpret_cast<uint32_t*>(d0 + 8 * LDS) = h1[0];
        if (P > 1) {
            *reinterpret_cast<uint32_t*>(d0 + SP * LDS) = h0[1];
            *reinterpret_cast<uint32_t*>(d0 + SP * LDS + 8 * LDS) = h1[1];
            if (P > 2) {
                *reinterpret_cast<uint32_t*>(d0 + 2 * SP * LDS) = h0[2];
                *reinterpret_cast<uint32_t*>(d0 + 2 * SP * LDS + 8 * LDS) = h1[2];
            }
        }
    }
}

// Staged rows i0 .. i0+15 (valid < S) of all planes -> global, 16 bytes per lane, full rows coalesced; optional column
// sums over the valid rows (planes added) into dbias[col0 + c] -- the in-projection bias gradient.
template <int HD, int SP>
__device__ __forceinline__ void copy_out_strip(const __nv_bfloat16* tile, const OutView& o, int P, int64_t row0, int col0, int i0, int S, int lane,
                                               float* dbias) {
    constexpr int LDS = HD + 8, CH = HD / 8;
    __syncwarp();
    const int nrow = min(16, S - i0);
    for (int pl = 0; pl < P; ++pl)
        for (int idx = lane; idx < nrow * CH; idx += 32) {
            const int r = idx / CH, c = (idx - r * CH) * 8;
            const uint4 v = *reinterpret_cast<const uint4*>(tile + (pl * SP + i0 + r) * LDS + c);
            *reinterpret_cast<uint4*>(o.ptr + pl * o.stride + (row0 + i0 + r) * o.ld + col0 + c) = v;
        }
    if (dbias) {
        for (int c = lane; c < HD; c += 32) {
            float s = 0.f;
            for (int pl = 0; pl < P; ++pl)
                for (int r = 0; r < nrow; ++r) s += __bfloat162float(tile[(pl * SP + i0 + r) * LDS + c]);
            atomicAdd(dbias + col0 + c, s);
        }
    }
    __syncwarp();
}

}  // namespace amma
}  // namespace m2f
