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

// global [S, HD] head slice (all planes) -> shared [P][SP][HD + 8], rows >= S zeroed.  16-byte chunks.
template <int HD, int SP>
__device__ __forceinline__ void load_tile(__nv_bfloat16* dst, const View& t, int P, int64_t row0, int col0, int S) {
    constexpr int LDS = HD + 8, CH = HD / 8;
    for (int i = threadIdx.x; i < P * SP * CH; i += blockDim.x) {
        const int pl = i / (SP * CH), rem = i - pl * (SP * CH);
        const int r = rem / CH, c = (rem - r * CH) * 8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (r < S) v = *reinterpret_cast<const uint4*>(t.ptr + pl * t.stride + (row0 + r) * t.ld + col0 + c);
        *reinterpret_cast<uint4*>(dst + (pl * SP + r) * LDS + c) = v;
    }
}

// acc[nt][4] (16 x SP strip) += A(16 x HD from sA planes, rows i0..) * B^T (sB planes stored [n][k], k contiguous)
template <int HD, int SP>
__device__ __forceinline__ void strip_nt(float (&acc)[SP / 8][4], const __nv_bfloat16* sA, const __nv_bfloat16* sB, int P, int i0, int lane) {
    constexpr int LDS = HD + 8;
    const int a_row = i0 + (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 8;
    const int b_row = (lane & 7) + (lane >> 4) * 8, b_col = ((lane >> 3) & 1) * 8;
    const int np = nprod(P);
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
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

// store a 16 x HD accumulator strip (rows r0.., valid rows < S) as planes; optional column sums into dbias[col]
template <int HD>
__device__ __forceinline__ void store_strip(const float (&acc)[HD / 8][4], float mul, const OutView& o, int P, int64_t row0, int col0, int r0, int S,
                                            int lane, float* dbias) {
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int nt = 0; nt < HD / 8; ++nt) {
        const int col = col0 + nt * 8 + 2 * q;
        float v[4] = {acc[nt][0] * mul, acc[nt][1] * mul, acc[nt][2] * mul, acc[nt][3] * mul};
        const bool ok0 = r0 + g < S, ok1 = r0 + g + 8 < S;
        if (dbias) {
            float s0 = (ok0 ? v[0] : 0.f) + (ok1 ? v[2] : 0.f), s1 = (ok0 ? v[1] : 0.f) + (ok1 ? v[3] : 0.f);
#pragma unroll
            for (int o2 = 4; o2 < 32; o2 <<= 1) {
                s0 += __shfl_xor_sync(0xffffffffu, s0, o2);
                s1 += __shfl_xor_sync(0xffffffffu, s1, o2);
            }
            if (g == 0) {
                atomicAdd(dbias + col, s0);
                atomicAdd(dbias + col + 1, s1);
            }
        }
        uint32_t h0[3], h1[3];
        split2(v[0], v[1], P, h0);
        split2(v[2], v[3], P, h1);
        for (int pl = 0; pl < P; ++pl) {
            __nv_bfloat16* base = o.ptr + pl * o.stride + col;
            if (ok0) *reinterpret_cast<uint32_t*>(base + (row0 + r0 + g) * o.ld) = pl == 0 ? h0[0] : (pl == 1 ? h0[1] : h0[2]);
            if (ok1) *reinterpret_cast<uint32_t*>(base + (row0 + r0 + g + 8) * o.ld) = pl == 0 ? h1[0] : (pl == 1 ? h1[1] : h1[2]);
        }
    }
}

}  // namespace amma
}  // namespace m2f
