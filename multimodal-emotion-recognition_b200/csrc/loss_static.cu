// K4s: cross entropy forward + backward for callers that may not allocate (code under CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  Same arithmetic as m2f_ce_fwd_bwd (loss.cu; torch.nn.CrossEntropyLoss(weight?,
// ignore_index=-1, label_smoothing) of src/train.py:48-50,229), but the caller passes NO scratch: the three running sums
// and one log-sum-exp per row (so the gradient pass does not redo the row reductions) live in a file-scope device array.
//
// The array is part of the library image: every GPU that loads libm2f_b200.so holds CE_STATIC_BYTES (about 1 MB) until the
// process exits.  Every call on a device shares it, and it keeps the previous call's contents: the sums are cleared at the
// start of each call (cudaMemsetAsync on the stream, no allocation), the per-row values are rewritten for every valid row
// before they are read.  Two calls that overlap on one GPU (different streams, or a graph replayed concurrently with
// itself) race on the array -- serialise them.
//
// Host code never passes the array's NAME to a kernel (on the host that is a stand-in symbol, not a device address): the
// device address is obtained once per device, with that device current, through cudaGetSymbolAddress, and checked.
#include <mutex>

#include "common.cuh"

namespace m2f {

constexpr int CES_MAX_C = 32;
constexpr int CES_THREADS = 256;
constexpr int64_t CES_MAX_ROWS = M2F_CE_STATIC_MAX_ROWS;  // 262144 rows: the C5 batch (512 x 256) twice over
constexpr int CES_ACC_FLOATS = 64;                        // sums in the first 256 bytes, per-row values behind them
constexpr int CES_MAX_DEVICES = 64;

static __device__ __align__(256) float g_ce_static_scratch[CES_ACC_FLOATS + CES_MAX_ROWS];

// scratch[1] += sum_V w_y; scratch[2] += sum_V -w_y lp_y; scratch[3] += sum_V sum_c -w_c lp_c; lse[m] for rows in V
__global__ void __launch_bounds__(CES_THREADS)
ces_reduce_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target, const float* __restrict__ w, int64_t M,
                  int C, float* __restrict__ scratch) {
    __shared__ float red[3][CES_THREADS / 32];
    float* __restrict__ lse_out = scratch + CES_ACC_FLOATS;
    float a_den = 0.f, a_nll = 0.f, a_sm = 0.f;
    for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = target[m];
        if (y < 0 || y >= C) continue;  // ignore_index = -1
        const float* z = logits + m * ld;
        float mx = -INFINITY;
        for (int c = 0; c < C; ++c) mx = fmaxf(mx, z[c]);
        float s = 0.f;
        for (int c = 0; c < C; ++c) s += expf(z[c] - mx);
        const float lse = mx + logf(s);
        lse_out[m] = lse;
        const float wy = w ? w[y] : 1.f;
        a_den += wy;
        a_nll -= wy * (z[y] - lse);
        float t = 0.f;
        for (int c = 0; c < C; ++c) t -= (w ? w[c] : 1.f) * (z[c] - lse);
        a_sm += t;
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
        for (int i = 0; i < CES_THREADS / 32; ++i) s += red[threadIdx.x][i];
        atomicAdd(scratch + 1 + threadIdx.x, s);
    }
}

__global__ void __launch_bounds__(CES_THREADS)
ces_grad_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target, const float* __restrict__ w, int64_t M,
                int C, float eps, float grad_scale, const float* __restrict__ scratch, float* __restrict__ loss_out, float* __restrict__ dl,
                int64_t dl_ld, __nv_bfloat16* __restrict__ dpl, int64_t dpl_ld, int64_t dpl_stride, int dpl_cols, int P) {
    const float* __restrict__ lse_in = scratch + CES_ACC_FLOATS;
    const float den = scratch[1];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        loss_out[0] = ((1.f - eps) * scratch[2] + (eps / (float)C) * scratch[3]) / den;
        loss_out[1] = den;
    }
    if (!dl && !dpl) return;
    float wsum = 0.f;
    for (int c = 0; c < C; ++c) wsum += w ? w[c] : 1.f;
    const float k = grad_scale / den;
    for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = target[m];
        const bool valid = y >= 0 && y < C;
        const float lse = valid ? lse_in[m] : 0.f;
        const float wy = valid ? (w ? w[y] : 1.f) : 0.f;
        for (int c = 0; c < C; ++c) {
            float gval = 0.f;
            if (valid) {
                const float p = expf(logits[m * ld + c] - lse);
                gval = k * ((1.f - eps) * wy * (p - (c == y ? 1.f : 0.f)) + (eps / (float)C) * (p * wsum - (w ? w[c] : 1.f)));
            }
            if (dl) dl[m * dl_ld + c] = gval;
            if (dpl) store_planes(dpl, dpl_stride, P, m * dpl_ld + c, gval);
        }
        if (dpl)
            for (int c = C; c < dpl_cols; ++c) store_planes(dpl, dpl_stride, P, m * dpl_ld + c, 0.f);  // TMA padding columns
    }
}

// Device address of the scratch array on the CURRENT device (resolved once per device, under a lock).
static int ces_scratch_ptr(float** out) {
    static std::mutex mu;
    static float* ptrs[CES_MAX_DEVICES] = {};
    int dev = -1;
    M2F_CUDA(cudaGetDevice(&dev));
    M2F_REQUIRE(dev >= 0 && dev < CES_MAX_DEVICES, M2F_ERR_UNSUPPORTED, "ce_static: device ordinal %d outside 0..%d", dev, CES_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        float* p = nullptr;
        M2F_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ce_static_scratch));
        M2F_REQUIRE(p != nullptr, M2F_ERR_CUDA, "ce_static: cudaGetSymbolAddress returned a null address on device %d", dev);
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return M2F_OK;
}

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_ce_fwd_bwd_static(const float* logits, int64_t ld, const int64_t* target, const float* class_w, int64_t M, int32_t C,
                                     float eps, float grad_scale, float* loss_out, float* dlogits, int64_t dl_ld,
                                     const m2f_planes_t* dl_planes, void* stream) {
    M2F_REQUIRE(logits && target && loss_out && M > 0, M2F_ERR_ARG, "ce_static: bad argument");
    M2F_REQUIRE(C >= 1 && C <= CES_MAX_C, M2F_ERR_UNSUPPORTED, "ce_static: C = %d unsupported (1..%d)", C, CES_MAX_C);
    M2F_REQUIRE(M <= CES_MAX_ROWS, M2F_ERR_UNSUPPORTED,
                "ce_static: M = %lld rows exceed the static scratch (M2F_CE_STATIC_MAX_ROWS = %lld); use m2f_ce_fwd_bwd", (long long)M,
                (long long)CES_MAX_ROWS);
    M2F_REQUIRE(ld >= C && (!dlogits || dl_ld >= C), M2F_ERR_ARG, "ce_static: row stride smaller than C");
    __nv_bfloat16* dpl = nullptr;
    int64_t dpl_ld = 0, dpl_stride = 0;
    int P = 1, dpl_cols = 0;
    if (dl_planes && dl_planes->ptr) {
        M2F_TRY(check_planes(dl_planes, "ce_static dl_planes", 1));
        M2F_REQUIRE(dl_planes->rows >= M && dl_planes->cols >= C, M2F_ERR_ARG, "ce_static: dl_planes extent mismatch");
        dpl = reinterpret_cast<__nv_bfloat16*>(dl_planes->ptr);
        dpl_ld = dl_planes->ld;
        dpl_stride = dl_planes->plane_stride;
        P = dl_planes->planes;
        dpl_cols = (int)dl_planes->cols;
    }
    float* scratch = nullptr;
    M2F_TRY(ces_scratch_ptr(&scratch));
    auto st = reinterpret_cast<cudaStream_t>(stream);
    M2F_CUDA(cudaMemsetAsync(scratch, 0, CES_ACC_FLOATS * sizeof(float), st));  // the sums start at zero; rows are rewritten
    int64_t g = (M + CES_THREADS - 1) / CES_THREADS;
    if (g > 148 * 4) g = 148 * 4;
    ces_reduce_kernel<<<(unsigned)g, CES_THREADS, 0, st>>>(logits, ld, target, class_w, M, C, scratch);
    M2F_LAUNCH_CHECK();
    ces_grad_kernel<<<(unsigned)g, CES_THREADS, 0, st>>>(logits, ld, target, class_w, M, C, eps, grad_scale, scratch, loss_out, dlogits, dl_ld,
                                                         dpl, dpl_ld, dpl_stride, dpl_cols, P);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
