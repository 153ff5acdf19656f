// K4: padding-masked, label-smoothed, optionally class-weighted cross entropy, forward + backward.
// Reference: torch.nn.CrossEntropyLoss(weight?, ignore_index=-1, label_smoothing=0.1) applied to
// outputs.permute(0,2,1) (src/train.py:48-50,229).  With w_c = 1 when unweighted, V = {rows with y != -1}:
//   L = [ (1-eps) * sum_V -w_y lp_y + (eps/C) * sum_V sum_c -w_c lp_c ] / sum_V w_y        
//   dL/dz_c (row in V) = [ (1-eps) w_y (p_c - [c == y]) + (eps/C) (p_c sum_c' w_c' - w_c) ] / sum_V w_y
#include "common.cuh"

namespace m2f {

constexpr int CE_MAX_C = 32;
constexpr int CE_THREADS = 256;

// CMAX: compile-time bound on C (8 or 32) so that the row lives in registers: loops are unrolled over CMAX with c < C tests
// (a loop over a runtime C indexed lp[] dynamically and put it in local memory).
template <int CMAX>
__device__ __forceinline__ void ce_row(const float* z, int C, float lp[CMAX]) {
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < CMAX; ++c) {
        lp[c] = c < C ? z[c] : -INFINITY;
        mx = fmaxf(mx, lp[c]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CMAX; ++c) s += c < C ? expf(lp[c] - mx) : 0.f;
    const float lse = mx + logf(s);
#pragma unroll
    for (int c = 0; c < CMAX; ++c) lp[c] -= lse;
}

// acc[1] += sum_V w_y; acc[2] += sum_V -w_y lp_y; acc[3] += sum_V sum_c -w_c lp_c
template <int CMAX>
__global__ void __launch_bounds__(CE_THREADS)
ce_reduce_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target, const float* __restrict__ w, int64_t M,
                 int C, float* acc) {
    __shared__ float red[3][CE_THREADS / 32];
    float a_den = 0.f, a_nll = 0.f, a_sm = 0.f;
    for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = target[m];
        if (y < 0 || y >= C) continue;  // ignore_index = -1
        float lp[CMAX];
        ce_row<CMAX>(logits + m * ld, C, lp);
        const float wy = w ? w[y] : 1.f;
        a_den += wy;
        float s = 0.f, lpy = 0.f;
#pragma unroll
        for (int c = 0; c < CMAX; ++c) {
            if (c < C) {
                s -= (w ? w[c] : 1.f) * lp[c];
                if (c == y) lpy = lp[c];
            }
        }
        a_nll -= wy * lpy;
        a_sm += s;
    }
    a_den = warp_sum(a_den);
    a_nll = warp_sum(a_nll);
    a_sm = warp_sum(a_sm);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        red[0][warp] = a_den;
        red[1][warp] = a_nll;
        red[2][warp] = a_sm;
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        float s = 0.f;
        for (int i = 0; i < CE_THREADS / 32; ++i) s += red[threadIdx.x][i];
        atomicAdd(acc + 1 + threadIdx.x, s);
    }
}

template <int CMAX>
__global__ void __launch_bounds__(CE_THREADS)
ce_grad_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target, const float* __restrict__ w, int64_t M,
               int C, float eps, float grad_scale, float* acc, float* __restrict__ dl, int64_t dl_ld, __nv_bfloat16* __restrict__ dpl,
               int64_t dpl_ld, int64_t dpl_stride, int dpl_cols, int P) {
    const float den = acc[1];
    if (blockIdx.x == 0 && threadIdx.x == 0) acc[0] = ((1.f - eps) * acc[2] + (eps / (float)C) * acc[3]) / den;
    if (!dl && !dpl) return;
    float wsum = 0.f;
    for (int c = 0; c < C; ++c) wsum += w ? w[c] : 1.f;
    const float k = grad_scale / den;
    for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = target[m];
        const bool valid = y >= 0 && y < C;
        float lp[CMAX];
        if (valid) ce_row<CMAX>(logits + m * ld, C, lp);
        const float wy = valid ? (w ? w[y] : 1.f) : 0.f;
#pragma unroll
        for (int c = 0; c < CMAX; ++c) {
            if (c < C) {
                float gval = 0.f;
                if (valid) {
                    const float p = expf(lp[c]);
                    gval = k * ((1.f - eps) * wy * (p - (c == y ? 1.f : 0.f)) + (eps / (float)C) * (p * wsum - (w ? w[c] : 1.f)));
                }
                if (dl) dl[m * dl_ld + c] = gval;
                if (dpl) store_planes(dpl, dpl_stride, P, m * dpl_ld + c, gval);
            }
        }
        if (dpl)
            for (int c = C; c < dpl_cols; ++c) store_planes(dpl, dpl_stride, P, m * dpl_ld + c, 0.f);  // zero the TMA padding columns
    }
}

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_ce_fwd_bwd(const float* logits, int64_t ld, const int64_t* target, const float* class_w, int64_t M, int32_t C, float eps,
                              float grad_scale, float* loss_out, float* dlogits, int64_t dl_ld, const m2f_planes_t* dl_planes, void* stream) {
    M2F_REQUIRE(logits && target && loss_out && M > 0, M2F_ERR_ARG, "ce: bad argument");
    M2F_REQUIRE(C >= 1 && C <= CE_MAX_C, M2F_ERR_UNSUPPORTED, "ce: C = %d unsupported (1..%d)", C, CE_MAX_C);
    M2F_REQUIRE(ld >= C && (!dlogits || dl_ld >= C), M2F_ERR_ARG, "ce: row stride smaller than C");
    __nv_bfloat16* dpl = nullptr;
    int64_t dpl_ld = 0, dpl_stride = 0;
    int P = 1, dpl_cols = 0;
    if (dl_planes && dl_planes->ptr) {
        M2F_TRY(check_planes(dl_planes, "ce dl_planes", 1));
        M2F_REQUIRE(dl_planes->rows >= M && dl_planes->cols >= C, M2F_ERR_ARG, "ce: dl_planes extent mismatch");
        dpl = reinterpret_cast<__nv_bfloat16*>(dl_planes->ptr);
        dpl_ld = dl_planes->ld;
        dpl_stride = dl_planes->plane_stride;
        P = dl_planes->planes;
        dpl_cols = (int)dl_planes->cols;
    }
    auto st = reinterpret_cast<cudaStream_t>(stream);
    M2F_CUDA(cudaMemsetAsync(loss_out, 0, 4 * sizeof(float), st));
    int64_t g = (M + CE_THREADS - 1) / CE_THREADS;
    if (g > 148 * 4) g = 148 * 4;
    if (C <= 8) {
        ce_reduce_kernel<8><<<(unsigned)g, CE_THREADS, 0, st>>>(logits, ld, target, class_w, M, C, loss_out);
        M2F_LAUNCH_CHECK();
        ce_grad_kernel<8><<<(unsigned)g, CE_THREADS, 0, st>>>(logits, ld, target, class_w, M, C, eps, grad_scale, loss_out, dlogits, dl_ld, dpl,
                                                              dpl_ld, dpl_stride, dpl_cols, P);
    } else {
        ce_reduce_kernel<CE_MAX_C><<<(unsigned)g, CE_THREADS, 0, st>>>(logits, ld, target, class_w, M, C, loss_out);
        M2F_LAUNCH_CHECK();
        ce_grad_kernel<CE_MAX_C><<<(unsigned)g, CE_THREADS, 0, st>>>(logits, ld, target, class_w, M, C, eps, grad_scale, loss_out, dlogits, dl_ld,
                                                                     dpl, dpl_ld, dpl_stride, dpl_cols, P);
    }
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
