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
    int b_static;  // B is not written by any kernel of the step (weights): its first tiles may be fetched before the PDL wait
    int epi_tma;  // 1: the [M,N]-shaped epilogue operands have tensor maps (EpiMaps) -- TMA epilogue; 0: pointer epilogue
};

// Tensor maps of the [M,N]-shaped epilogue operands, box = 32 rows x 32 columns (one epilogue warp's chunk):
// fp32 maps use the 128-byte swizzle (32 x 4 B rows), bf16 maps the 64-byte swizzle.  Unused ones stay zeroed.
struct EpiMaps {
    CUtensorMap out_f32;  // 2-D fp32 [M, N]: stored, and loaded first when accumulating
    CUtensorMap out_pl;   // 3-D bf16 [planes, M, N]
    CUtensorMap gate;     // 2-D bf16 [M, N]
    CUtensorMap resid;    // 2-D fp32 [M, N]
};
// per epilogue warp; <= 1 output plane: 2 x out f32 4K | 2 x out plane 2K | gate 2K | resid 4K | accumulate 4K (22 K);
// 2-3 planes: out f32 4K | out planes 3 x 2K | gate 2K | resid 4K | accumulate 4K (20 K)
constexpr int EPI2_WARP_BYTES = 23 * 1024;

// host helpers (gemm_sm100.cu)
int make_planes_tmap(CUtensorMap* map, const m2f_planes_t* t, int box_rows);
int make_planes_tmap_sw64(CUtensorMap* map, const m2f_planes_t* t, int box_rows);  // 32-column boxes, 64-byte swizzle
constexpr int GEMM_TILE_CTA2 = 0x1000;
constexpr int GEMM_TILE_PERSIST = 0x2000;  // very many row tiles: the persistent pair kernel (gemm_persist.cu) where its epilogue applies
bool gemm_persist_eligible(const GemmParams& p, int a_mn, int b_mn);
int gemm_launch_persist(const CUtensorMap& ta, const CUtensorMap& tb, const EpiMaps& em, const GemmParams& p, cudaStream_t stream);
int gemm_block_n(int64_t M, int64_t N, int64_t K);   // tile code: BLOCK_N | GEMM_TILE_CTA2
int gemm_b_box_rows(int tile);            // rows of B one CTA stages per k-block (K-major tensor-map box)
int gemm_fill_params(GemmParams& p, const m2f_planes_t* A, int a_mn, const m2f_planes_t* B, int b_mn, const m2f_kseg_t* segs,
                     int nsegs, int64_t M, int64_t N, const m2f_epilogue_t* e);
// decides whether the TMA epilogue applies (alignment of every operand present) and encodes its maps; sets p.epi_tma
int gemm_make_epi_maps(GemmParams& p, EpiMaps* maps);
int gemm_launch(const CUtensorMap& ta, const CUtensorMap& tb, const EpiMaps& em, const GemmParams& p, int tile, int a_mn, int b_mn,
                cudaStream_t stream);

}  // namespace m2f
