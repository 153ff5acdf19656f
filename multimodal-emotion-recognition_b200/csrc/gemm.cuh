// Internal interface of the tcgen05 GEMM (gemm_sm100.cu), shared with the model-level executor (plan.cu).
#pragma once
#include "common.cuh"

namespace m2f {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // one 128-byte swizzle atom of bf16
constexpr int UMMA_K = 16;
constexpr int MAX_SEGS = 4;
constexpr int MAX_PROD = 6;
constexpr int GEMM_THREADS = 192;
constexpr int EPI_STG_LD = 36;                    // floats per staged accumulator row (32 + 4: conflict-free float4 rows)
constexpr int EPI_STG_FLOATS = 32 * EPI_STG_LD;   // one epilogue warp's staging tile

struct GemmParams {
    int M, N;
    int nsegs;
    int seg_a_k0[MAX_SEGS], seg_b_k0[MAX_SEGS], seg_nkb[MAX_SEGS];
    int nprod;
    int prod_a[MAX_PROD], prod_b[MAX_PROD];
    int total_kb;
    // epilogue
    const float* bias;
    int relu;
    int accumulate;
    const __nv_bfloat16* gate;
    int64_t gate_ld;
    float gate_scale;
    Drop drop;
    int64_t drop_ld;
    const float* resid;
    int64_t resid_ld;
    float* out_f32;
    int64_t out_f32_ld;
    float* colsum;
    int vec4;
    int dbg;
    unsigned long long* trace;  // debug: 8 timestamps per CTA (nullptr in production)
    __nv_bfloat16* out_pl;
    int64_t out_pl_ld, out_pl_stride;
    int out_pl_n;
};

// host helpers (gemm_sm100.cu)
int make_planes_tmap(CUtensorMap* map, const m2f_planes_t* t, int box_rows);
constexpr int GEMM_TILE_CTA2 = 0x1000;
int gemm_block_n(int64_t M, int64_t N);   // tile code: BLOCK_N | GEMM_TILE_CTA2
int gemm_b_box_rows(int tile);            // rows of B one CTA stages per k-block (K-major tensor-map box)
int gemm_fill_params(GemmParams& p, const m2f_planes_t* A, int a_mn, const m2f_planes_t* B, int b_mn, const m2f_kseg_t* segs,
                     int nsegs, int64_t M, int64_t N, const m2f_epilogue_t* e);
int gemm_launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tile, int a_mn, int b_mn,
                cudaStream_t stream);

}  // namespace m2f
