// K2: masked multi-head attention, one CTA per (dialogue, head); the whole problem lives in shared memory.
//   P = softmax(q k^T / sqrt(hd), -inf on padded KEYS);  O = dropout(P) v
// Reference arithmetic: F.multi_head_attention_forward, torch/nn/functional.py:6607-6620 (key_padding_mask ->
// additive -inf mask), :6630-6660 (need_weights=True branch taken by the fusion layers, src/model.py:14) and
// :6676-6690 (scaled_dot_product_attention branch taken by the encoder layers); dropout is applied to P after
// the softmax (:6643-6645).  Padded QUERY rows are computed like any other row.
// Backward recomputes P from q, k (no S x S tensor is ever written to HBM).
#include <math.h>

#include "attention_mma.cuh"
#include <algorithm>

#include "common.cuh"

namespace m2f {

constexpr int ATT_THREADS = 128;
constexpr int ATT_WARPS = ATT_THREADS / 32;
constexpr int ATTB_THREADS = 256;
constexpr int ATTB_WARPS = ATTB_THREADS / 32;
constexpr int ATT_MAX_KPL = 8;  // keys per lane => S <= 256

struct PlanesView {
    const __nv_bfloat16* ptr;
    int64_t ld, stride;
    int P;
};
struct PlanesOut {
    __nv_bfloat16* ptr;
    int64_t ld, stride;
    int P;
};

// load a [S, hd] head slice (rows b*S.., cols h*hd..) into padded fp32 smem [S][hd_pad]
__device__ __forceinline__ void load_head(float* dst, int hd_pad, const PlanesView& t, int64_t row0, int col0, int S, int hd,
                                          float scale) {
    if ((hd & 3) == 0 && (col0 & 3) == 0) {
        const int hd4 = hd >> 2;
        for (int i = threadIdx.x; i < S * hd4; i += blockDim.x) {
            const int r = i / hd4, c = (i - r * hd4) * 4;
            float v[4];
            load_planes4(t.ptr, t.stride, t.P, (row0 + r) * t.ld + col0 + c, v);
            float* d = dst + r * hd_pad + c;
            d[0] = v[0] * scale; d[1] = v[1] * scale; d[2] = v[2] * scale; d[3] = v[3] * scale;
        }
    } else {
        for (int i = threadIdx.x; i < S * hd; i += blockDim.x) {
            const int r = i / hd, c = i - r * hd;
            dst[r * hd_pad + c] = load_planes(t.ptr, t.stride, t.P, (row0 + r) * t.ld + col0 + c) * scale;
        }
    }
}

// ------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_kernel(PlanesView q, PlanesView k, PlanesView v, const uint8_t* __restrict__ key_pad, int S, int H, int hd, float scale,
                Drop drop, PlanesOut o) {
    drop = drop_resolve(drop);
    extern __shared__ float sm[];
    const int hd_pad = hd | 1;  // odd row stride: conflict-free column walks
    float* sK = sm;
    float* sV = sK + S * hd_pad;
    float* sQ = sV + S * hd_pad;            // [ATT_WARPS][hd_pad]
    float* sP = sQ + ATT_WARPS * hd_pad;    // [ATT_WARPS][S]
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int64_t row0 = (int64_t)b * S;
    const int col0 = h * hd;
    load_head(sK, hd_pad, k, row0, col0, S, hd, 1.f);
    load_head(sV, hd_pad, v, row0, col0, S, hd, 1.f);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* myQ = sQ + warp * hd_pad;
    float* myP = sP + warp * S;
    const uint8_t* pad = key_pad + row0;
    for (int i = warp; i < S; i += ATT_WARPS) {
        for (int d = lane; d < hd; d += 32) myQ[d] = load_planes(q.ptr, q.stride, q.P, (row0 + i) * q.ld + col0 + d) * scale;
        __syncwarp();
        float sc[ATT_MAX_KPL];
        float mx = -INFINITY;
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            const int j = lane + 32 * t;
            float s = -INFINITY;
            if (j < S && !pad[j]) {
                s = 0.f;
                const float* kr = sK + j * hd_pad;
                for (int d = 0; d < hd; ++d) s = fmaf(myQ[d], kr[d], s);
            }
            sc[t] = s;
            mx = fmaxf(mx, s);
        }
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            sc[t] = (sc[t] == -INFINITY) ? 0.f : __expf(sc[t] - mx);
            sum += sc[t];
        }
        sum = warp_sum(sum);
        const float inv = 1.f / sum;  // every dialogue keeps >= 1 valid key
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            const int j = lane + 32 * t;
            if (j < S) {
                float p = sc[t] * inv;
                if (drop.enabled) p *= drop_mult1(drop, ((uint64_t)blockIdx.x * S + i) * ((S + 7) & ~7) + j);
                myP[j] = p;
            }
        }
        __syncwarp();
        for (int d = lane; d < hd; d += 32) {
            float acc = 0.f;
            for (int j = 0; j < S; ++j) acc = fmaf(myP[j], sV[j * hd_pad + d], acc);
            store_planes(o.ptr, o.stride, o.P, (row0 + i) * o.ld + col0 + d, acc);
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------ backward
// dV = Pd^T dO;  dPd = dO V^T;  dP = dPd * M;  dS = P * (dP - rowsum(dP * P));  dQ = c dS K;  dK = c dS^T Q
// with Pd = P * M, M = dropout mask / (1 - p), c = 1 / sqrt(hd).
__global__ void __launch_bounds__(ATTB_THREADS)
attn_bwd_kernel(PlanesView q, PlanesView k, PlanesView v, PlanesView d_o, const uint8_t* __restrict__ key_pad, int S, int H, int hd,
                float scale, Drop drop, PlanesOut dq, PlanesOut dk, PlanesOut dv) {
    drop = drop_resolve(drop);
    extern __shared__ float sm[];
    const int hd_pad = hd | 1;
    const int s_pad = S | 1;
    float* sQ = sm;                      // holds q * scale
    float* sK = sQ + S * hd_pad;
    float* sV = sK + S * hd_pad;
    float* sdO = sV + S * hd_pad;
    float* sPd = sdO + S * hd_pad;       // [S][s_pad]  dropped probabilities
    float* sdS = sPd + S * s_pad;        // [S][s_pad]
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int64_t row0 = (int64_t)b * S;
    const int col0 = h * hd;
    load_head(sQ, hd_pad, q, row0, col0, S, hd, scale);
    load_head(sK, hd_pad, k, row0, col0, S, hd, 1.f);
    load_head(sV, hd_pad, v, row0, col0, S, hd, 1.f);
    load_head(sdO, hd_pad, d_o, row0, col0, S, hd, 1.f);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint8_t* pad = key_pad + row0;
    for (int i = warp; i < S; i += ATTB_WARPS) {
        const float* qr = sQ + i * hd_pad;
        const float* gr = sdO + i * hd_pad;
        float sc[ATT_MAX_KPL], dp[ATT_MAX_KPL];
        float mx = -INFINITY;
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            const int j = lane + 32 * t;
            float s = -INFINITY, g = 0.f;
            if (j < S && !pad[j]) {
                s = 0.f;
                const float* kr = sK + j * hd_pad;
                const float* vr = sV + j * hd_pad;
                for (int d = 0; d < hd; ++d) {
                    s = fmaf(qr[d], kr[d], s);
                    g = fmaf(gr[d], vr[d], g);
                }
            }
            sc[t] = s;
            dp[t] = g;
            mx = fmaxf(mx, s);
        }
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            sc[t] = (sc[t] == -INFINITY) ? 0.f : __expf(sc[t] - mx);
            sum += sc[t];
        }
        const float inv = 1.f / warp_sum(sum);
        float delta = 0.f;
        float mlt[ATT_MAX_KPL];
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            const int j = lane + 32 * t;
            sc[t] *= inv;  // P
            mlt[t] = 1.f;
            if (drop.enabled && j < S) mlt[t] = drop_mult1(drop, ((uint64_t)blockIdx.x * S + i) * ((S + 7) & ~7) + j);
            dp[t] *= mlt[t];  // dP
            delta += dp[t] * sc[t];
        }
        delta = warp_sum(delta);
#pragma unroll
        for (int t = 0; t < ATT_MAX_KPL; ++t) {
            const int j = lane + 32 * t;
            if (j < S) {
                sPd[i * s_pad + j] = sc[t] * mlt[t];
                sdS[i * s_pad + j] = sc[t] * (dp[t] - delta);
            }
        }
    }
    __syncthreads();
    // outputs: each warp owns rows r = warp, warp + W, ...; lanes own head-dim columns
    for (int r = warp; r < S; r += ATTB_WARPS) {
        for (int d = lane; d < hd; d += 32) {
            float aq = 0.f, ak = 0.f, av = 0.f;
            for (int t = 0; t < S; ++t) {
                aq = fmaf(sdS[r * s_pad + t], sK[t * hd_pad + d], aq);   // dQ[r] = sum_j dS[r][j] K[j]
                ak = fmaf(sdS[t * s_pad + r], sQ[t * hd_pad + d], ak);   // dK[r] = sum_i dS[i][r] (c q[i])
                av = fmaf(sPd[t * s_pad + r], sdO[t * hd_pad + d], av);  // dV[r] = sum_i Pd[i][r] dO[i]
            }
            store_planes(dq.ptr, dq.stride, dq.P, (row0 + r) * dq.ld + col0 + d, aq * scale);
            store_planes(dk.ptr, dk.stride, dk.P, (row0 + r) * dk.ld + col0 + d, ak);
            store_planes(dv.ptr, dv.stride, dv.P, (row0 + r) * dv.ld + col0 + d, av);
        }
    }
}

// ------------------------------------------------------------------------------------ tensor-path kernels (S <= 64)
// attention_mma_p{1,2,3}.cu instantiate attention_mma_impl.cuh for one plane count each (compile time, code size)
namespace amma {
int fwd_p1(AMMA_FWD_ARGS); int fwd_p2(AMMA_FWD_ARGS); int fwd_p3(AMMA_FWD_ARGS);
int bwd_p1(AMMA_BWD_ARGS); int bwd_p2(AMMA_BWD_ARGS); int bwd_p3(AMMA_BWD_ARGS);
// attention_mma_long.cu: 64 < S <= LONG_MAX_S on mma.sync -- forward (online softmax over key tiles; with or without probability
// dropout) and backward (tiled, statistics recomputed on chip)
constexpr int LONG_MAX_S = 4096;
bool long_supported(int hd, int64_t S);
int fwd_long(int hd, int P, const View& q, const View& k, const View& v, const uint8_t* pad, int64_t BH, int S, int H, float scale,
             const Drop& drop, const OutView& o, cudaStream_t st);
int bwd_long(int hd, int P, const View& q, const View& k, const View& v, const View& d_o, const uint8_t* pad, int64_t BH, int S, int H, float scale,
             const Drop& drop, const OutView& dq, const OutView& dk, const OutView& dv, float* dbq, float* dbk, float* dbv, cudaStream_t st);
// attention_tc.cu: the same on tcgen05 / TMEM / TMA for one plane and head sizes 64 / 96 / 128
bool tc_supported(int hd, int64_t S, int planes, int64_t rows);
int fwd_long_tc(int hd, const m2f_planes_t* q, const m2f_planes_t* k, const m2f_planes_t* v, const uint8_t* pad, int64_t B, int S, int H,
                const OutView& o, cudaStream_t st);
static bool supported(int hd, int64_t S) { return S <= 64 && (hd == 16 || hd == 32 || hd == 64 || hd == 96 || hd == 128); }
// 16-byte row chunks: base, leading dimension, plane stride and the head's first column must be multiples of 8 elements
static bool aligned16(const m2f_planes_t* t) {
    return (reinterpret_cast<uintptr_t>(t->ptr) % 16) == 0 && (t->ld % 8) == 0 && (t->planes == 1 || (t->plane_stride % 8) == 0);
}
}  // namespace amma

static int att_view(const m2f_planes_t* t, const char* name, int64_t rows, int64_t cols, PlanesView* out) {
    M2F_TRY(check_planes(t, name, 1));
    M2F_REQUIRE(t->rows >= rows && t->cols == cols, M2F_ERR_ARG, "%s: extent [%lld,%lld] does not cover [%lld,%lld]", name,
                (long long)t->rows, (long long)t->cols, (long long)rows, (long long)cols);
    out->ptr = reinterpret_cast<const __nv_bfloat16*>(t->ptr);
    out->ld = t->ld;
    out->stride = t->plane_stride;
    out->P = t->planes;
    return M2F_OK;
}
static PlanesOut as_out(const PlanesView& v) { return PlanesOut{const_cast<__nv_bfloat16*>(v.ptr), v.ld, v.stride, v.P}; }

// vectorised loads need 8-byte aligned rows
static bool vec_ok(const m2f_planes_t* t) {
    return (t->ld % 4) == 0 && (reinterpret_cast<uintptr_t>(t->ptr) % 8) == 0 && (t->planes == 1 || (t->plane_stride % 4) == 0);
}

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_attn_fwd(const m2f_planes_t* q, const m2f_planes_t* k, const m2f_planes_t* v, const uint8_t* key_pad, int64_t B,
                            int64_t S, int32_t H, float drop_p, uint64_t seed, uint64_t step, uint32_t site, const m2f_planes_t* o,
                            void* stream) {
    M2F_REQUIRE(q && k && v && o && key_pad && B > 0 && S > 0 && H > 0, M2F_ERR_ARG, "attn_fwd: bad argument");
    M2F_REQUIRE(q->cols % H == 0, M2F_ERR_ARG, "attn_fwd: embed dim %lld not divisible by %d heads", (long long)q->cols, H);
    M2F_REQUIRE(drop_p >= 0.f && drop_p < 1.f, M2F_ERR_ARG, "attn_fwd: dropout p must be in [0,1)");
    const int64_t E = q->cols;
    const int hd = (int)(E / H);
    PlanesView vq, vk, vv, vo;
    M2F_TRY(att_view(q, "attn_fwd q", B * S, E, &vq));
    M2F_TRY(att_view(k, "attn_fwd k", B * S, E, &vk));
    M2F_TRY(att_view(v, "attn_fwd v", B * S, E, &vv));
    M2F_TRY(att_view(o, "attn_fwd o", B * S, E, &vo));
    M2F_REQUIRE(vec_ok(q) && vec_ok(k) && vec_ok(v), M2F_ERR_ALIGN, "attn_fwd: q/k/v need ld %% 4 == 0 and 8-byte aligned bases");
    M2F_REQUIRE(q->planes == k->planes && q->planes == v->planes && q->planes == o->planes, M2F_ERR_ARG, "attn_fwd: plane counts differ");
    if (amma::supported(hd, S) && amma::aligned16(q) && amma::aligned16(k) && amma::aligned16(v) && amma::aligned16(o)) {
        // short dialogues: tensor-path kernel, whole (dialogue, head) problem on chip
        const amma::View aq{vq.ptr, vq.ld, vq.stride}, ak{vk.ptr, vk.ld, vk.stride}, av{vv.ptr, vv.ld, vv.stride};
        const amma::OutView ao{const_cast<__nv_bfloat16*>(vo.ptr), vo.ld, vo.stride};
        const int SPv = (int)((S + 15) / 16) * 16;
        const Drop dropm = drop_make(drop_p, seed, step, site);
        const float sc = 1.f / sqrtf((float)hd);
        auto st = reinterpret_cast<cudaStream_t>(stream);
        auto fn = q->planes == 1 ? amma::fwd_p1 : (q->planes == 2 ? amma::fwd_p2 : amma::fwd_p3);
        return fn(hd, SPv, aq, ak, av, key_pad, B * H, (int)S, H, sc, dropm, ao, st);
    }
    if (amma::long_supported(hd, S) && amma::aligned16(q) && amma::aligned16(k) && amma::aligned16(v) && amma::aligned16(o)) {
        // long dialogues: K/V tiles through shared memory, online softmax over 64-key tiles (inference: the tcgen05 kernel where it applies)
        const amma::View aq{vq.ptr, vq.ld, vq.stride}, ak{vk.ptr, vk.ld, vk.stride}, av{vv.ptr, vv.ld, vv.stride};
        const amma::OutView ao{const_cast<__nv_bfloat16*>(vo.ptr), vo.ld, vo.stride};
        if (drop_p == 0.f && amma::tc_supported(hd, S, q->planes, std::min(std::min(q->rows, k->rows), v->rows)) &&
            (reinterpret_cast<uintptr_t>(o->ptr) % 32) == 0 && (o->ld % 16) == 0)  // 32-byte output stores
            return amma::fwd_long_tc(hd, q, k, v, key_pad, B, (int)S, H, ao, reinterpret_cast<cudaStream_t>(stream));
        return amma::fwd_long(hd, q->planes, aq, ak, av, key_pad, B * H, (int)S, H, 1.f / sqrtf((float)hd), drop_make(drop_p, seed, step, site), ao,
                              reinterpret_cast<cudaStream_t>(stream));
    }
    M2F_REQUIRE(S <= 32 * ATT_MAX_KPL, M2F_ERR_UNSUPPORTED,
                "attn_fwd: S = %lld exceeds the single-CTA limit %d of the generic kernel (head sizes 32 / 64 / 96 / 128 take S <= %d)", (long long)S,
                32 * ATT_MAX_KPL, amma::LONG_MAX_S);
    const int hd_pad = hd | 1;
    const size_t smem = sizeof(float) * ((size_t)2 * S * hd_pad + ATT_WARPS * hd_pad + ATT_WARPS * S);
    M2F_REQUIRE(smem <= 227 * 1024, M2F_ERR_UNSUPPORTED, "attn_fwd: S = %lld, head_dim = %d needs %zu bytes of shared memory (> 227 KB)",
                (long long)S, hd, smem);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    const Drop drop = drop_make(drop_p, seed, step, site);
    attn_fwd_kernel<<<(unsigned)(B * H), ATT_THREADS, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
        vq, vk, vv, key_pad, (int)S, H, hd, 1.f / sqrtf((float)hd), drop, as_out(vo));
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_attn_bwd(const m2f_planes_t* q, const m2f_planes_t* k, const m2f_planes_t* v, const m2f_planes_t* d_o,
                            const uint8_t* key_pad, int64_t B, int64_t S, int32_t H, float drop_p, uint64_t seed, uint64_t step,
                            uint32_t site, const m2f_planes_t* dq, const m2f_planes_t* dk, const m2f_planes_t* dv, float* dbq, float* dbk,
                            float* dbv, void* stream) {
    M2F_REQUIRE(q && k && v && d_o && dq && dk && dv && key_pad && B > 0 && S > 0 && H > 0, M2F_ERR_ARG, "attn_bwd: bad argument");
    M2F_REQUIRE(q->cols % H == 0, M2F_ERR_ARG, "attn_bwd: embed dim %lld not divisible by %d heads", (long long)q->cols, H);
    M2F_REQUIRE(drop_p >= 0.f && drop_p < 1.f, M2F_ERR_ARG, "attn_bwd: dropout p must be in [0,1)");
    const int64_t E = q->cols;
    const int hd = (int)(E / H);
    PlanesView vq, vk, vv, vdo, vdq, vdk, vdv;
    M2F_TRY(att_view(q, "attn_bwd q", B * S, E, &vq));
    M2F_TRY(att_view(k, "attn_bwd k", B * S, E, &vk));
    M2F_TRY(att_view(v, "attn_bwd v", B * S, E, &vv));
    M2F_TRY(att_view(d_o, "attn_bwd d_o", B * S, E, &vdo));
    M2F_TRY(att_view(dq, "attn_bwd dq", B * S, E, &vdq));
    M2F_TRY(att_view(dk, "attn_bwd dk", B * S, E, &vdk));
    M2F_TRY(att_view(dv, "attn_bwd dv", B * S, E, &vdv));
    M2F_REQUIRE(vec_ok(q) && vec_ok(k) && vec_ok(v) && vec_ok(d_o), M2F_ERR_ALIGN,
                "attn_bwd: q/k/v/d_o need ld %% 4 == 0 and 8-byte aligned bases");
    M2F_REQUIRE(q->planes == k->planes && q->planes == v->planes && q->planes == d_o->planes && q->planes == dq->planes &&
                    q->planes == dk->planes && q->planes == dv->planes, M2F_ERR_ARG, "attn_bwd: plane counts differ");
    auto out_ok = [](const m2f_planes_t* t) { return amma::aligned16(t); };
    if (amma::supported(hd, S) && amma::aligned16(q) && amma::aligned16(k) && amma::aligned16(v) && amma::aligned16(d_o) && out_ok(dq) &&
        out_ok(dk) && out_ok(dv)) {
        const amma::View aq{vq.ptr, vq.ld, vq.stride}, ak{vk.ptr, vk.ld, vk.stride}, av{vv.ptr, vv.ld, vv.stride}, ado{vdo.ptr, vdo.ld, vdo.stride};
        const amma::OutView odq{const_cast<__nv_bfloat16*>(vdq.ptr), vdq.ld, vdq.stride}, odk{const_cast<__nv_bfloat16*>(vdk.ptr), vdk.ld, vdk.stride},
            odv{const_cast<__nv_bfloat16*>(vdv.ptr), vdv.ld, vdv.stride};
        const int SPv = (int)((S + 15) / 16) * 16;
        const size_t need = (size_t)4 * q->planes * SPv * (hd + 8) * 2 + (size_t)2 * q->planes * SPv * (SPv + 8) * 2;
        if (need <= 227 * 1024) {
            const Drop dropm = drop_make(drop_p, seed, step, site);
            const float sc = 1.f / sqrtf((float)hd);
            auto st = reinterpret_cast<cudaStream_t>(stream);
            auto fn = q->planes == 1 ? amma::bwd_p1 : (q->planes == 2 ? amma::bwd_p2 : amma::bwd_p3);
            return fn(hd, SPv, aq, ak, av, ado, key_pad, B * H, (int)S, H, sc, dropm, odq, odk, odv, dbq, dbk, dbv, st);
        }
    }
    if (amma::long_supported(hd, S) && amma::aligned16(q) && amma::aligned16(k) && amma::aligned16(v) && amma::aligned16(d_o) && out_ok(dq) &&
        out_ok(dk) && out_ok(dv)) {
        // long dialogues: tiled backward, row statistics recomputed on chip (attention_mma_long.cu)
        const amma::View aq{vq.ptr, vq.ld, vq.stride}, ak{vk.ptr, vk.ld, vk.stride}, av{vv.ptr, vv.ld, vv.stride}, ado{vdo.ptr, vdo.ld, vdo.stride};
        const amma::OutView odq{const_cast<__nv_bfloat16*>(vdq.ptr), vdq.ld, vdq.stride}, odk{const_cast<__nv_bfloat16*>(vdk.ptr), vdk.ld, vdk.stride},
            odv{const_cast<__nv_bfloat16*>(vdv.ptr), vdv.ld, vdv.stride};
        return amma::bwd_long(hd, q->planes, aq, ak, av, ado, key_pad, B * H, (int)S, H, 1.f / sqrtf((float)hd), drop_make(drop_p, seed, step, site),
                              odq, odk, odv, dbq, dbk, dbv, reinterpret_cast<cudaStream_t>(stream));
    }
    M2F_REQUIRE(S <= 32 * ATT_MAX_KPL, M2F_ERR_UNSUPPORTED,
                "attn_bwd: S = %lld exceeds the single-CTA limit %d of the generic kernel (head sizes 32 / 64 / 96 / 128 take S <= %d)", (long long)S,
                32 * ATT_MAX_KPL, amma::LONG_MAX_S);
    const int hd_pad = hd | 1, s_pad = (int)S | 1;
    const size_t smem = sizeof(float) * ((size_t)4 * S * hd_pad + (size_t)2 * S * s_pad);
    M2F_REQUIRE(smem <= 227 * 1024, M2F_ERR_UNSUPPORTED,
                "attn_bwd: S = %lld, head_dim = %d needs %zu bytes of shared memory (> 227 KB) in the generic kernel; head sizes 32 / 64 / 96 / 128 "
                "train at any S <= %d", (long long)S, hd, smem, amma::LONG_MAX_S);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        M2F_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    const Drop drop = drop_make(drop_p, seed, step, site);
    attn_bwd_kernel<<<(unsigned)(B * H), ATTB_THREADS, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
        vq, vk, vv, vdo, key_pad, (int)S, H, hd, 1.f / sqrtf((float)hd), drop, as_out(vdq), as_out(vdk), as_out(vdv));
    M2F_LAUNCH_CHECK();
    // generic path: the in-projection bias gradients are separate column-sum launches
    if (dbq) M2F_TRY(m2f_colsum_planes(dq, dbq, stream));
    if (dbk) M2F_TRY(m2f_colsum_planes(dk, dbk, stream));
    if (dbv) M2F_TRY(m2f_colsum_planes(dv, dbv, stream));
    return M2F_OK;
}
