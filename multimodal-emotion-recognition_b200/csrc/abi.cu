// Error plumbing, version / device query and the small element-wise kernels of the C ABI.
#include <stdarg.h>

#include <atomic>
#include <stdio.h>

#include <cstdlib>
#include "common.cuh"

namespace m2f {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
void note_launches(long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
static thread_local const unsigned long long* tl_rng_source = nullptr;
const unsigned long long* rng_source() { return tl_rng_source; }
bool pdl_enabled() {
    static const bool on = [] { const char* e = std::getenv("M2F_PDL"); return !(e && e[0] == '0'); }();
    return on;
}

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return M2F_ERR_CUDA;
}

int check_planes(const m2f_planes_t* t, const char* name, int vec) {
    M2F_REQUIRE(t && t->ptr, M2F_ERR_ARG, "%s: null planes view", name);
    M2F_REQUIRE(t->rows > 0 && t->cols > 0 && t->planes >= 1 && t->planes <= 3, M2F_ERR_ARG, "%s: bad extent/planes", name);
    M2F_REQUIRE(t->ld >= t->cols && (t->ld % vec) == 0 && (t->planes == 1 || (t->plane_stride % vec) == 0), M2F_ERR_ALIGN,
                "%s: ld/plane stride must be multiples of %d", name, vec);
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(t->ptr) % (2 * vec)) == 0, M2F_ERR_ALIGN, "%s: base must be %d-byte aligned", name, 2 * vec);
    return M2F_OK;
}

// ---------------------------------------------------------------- fp32 -> planes
__global__ void split_planes_kernel(const float* __restrict__ src, int64_t src_ld, __nv_bfloat16* dst, int64_t dst_ld,
                                    int64_t plane_stride, int P, int64_t rows, int64_t cols4) {
    const int64_t total = rows * cols4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols4, c = (i - r * cols4) * 4;
        const float4 q = *reinterpret_cast<const float4*>(src + r * src_ld + c);
        const float v[4] = {q.x, q.y, q.z, q.w};
        store_planes4(dst, plane_stride, P, r * dst_ld + c, v);
    }
}
__global__ void split_planes_scalar_kernel(const float* __restrict__ src, int64_t src_ld, __nv_bfloat16* dst, int64_t dst_ld,
                                           int64_t plane_stride, int P, int64_t rows, int64_t cols) {
    const int64_t total = rows * cols;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols, c = i - r * cols;
        store_planes(dst, plane_stride, P, r * dst_ld + c, src[r * src_ld + c]);
    }
}

__global__ void relu_planes_kernel(const __nv_bfloat16* __restrict__ src, int64_t src_ld, int64_t src_ps, int P,
                                   __nv_bfloat16* dst, int64_t dst_ld, int64_t dst_ps, int64_t rows, int64_t cols4) {
    const int64_t total = rows * cols4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols4, c = (i - r * cols4) * 4;
        float v[4];
        load_planes4(src, src_ps, P, r * src_ld + c, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = fmaxf(v[j], 0.f);
        store_planes4(dst, dst_ps, P, r * dst_ld + c, v);
    }
}

// out[n] += sum_m src[m, n]: block = 32 x 8 threads, one column per x-lane, rows strided over y and blockIdx.y
__global__ void colsum_planes_kernel(const __nv_bfloat16* __restrict__ src, int64_t ld, int64_t ps, int P, int64_t rows,
                                     int64_t cols, float* out) {
    __shared__ float red[8][33];
    const int64_t n = blockIdx.x * 32 + threadIdx.x;
    float acc = 0.f;
    if (n < cols)
        for (int64_t m = blockIdx.y * 8 + threadIdx.y; m < rows; m += (int64_t)gridDim.y * 8) acc += load_planes(src, ps, P, m * ld + n);
    red[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && n < cols) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += red[j][threadIdx.x];
        atomicAdd(out + n, s);
    }
}

__global__ void dropout_mask_kernel(Drop d, int64_t n, uint8_t* mask) {
    d = drop_resolve(d);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        mask[i] = drop_mult1(d, (uint64_t)i) > 0.f ? 1 : 0;
}

static int grid_for(int64_t total, int threads) {
    int64_t g = (total + threads - 1) / threads;
    if (g > 148 * 16) g = 148 * 16;
    return g < 1 ? 1 : (int)g;
}

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_abi_version(void) { return M2F_ABI_VERSION; }
extern "C" const char* m2f_last_error(void) { return g_err; }

extern "C" int m2f_set_rng_source(const uint64_t* seed_step) {
    m2f::tl_rng_source = reinterpret_cast<const unsigned long long*>(seed_step);
    return M2F_OK;
}
extern "C" int64_t m2f_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

extern "C" int m2f_check_device(void) {
    int dev = 0;
    M2F_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    M2F_CUDA(cudaGetDeviceProperties(&prop, dev));
    M2F_REQUIRE(prop.major == 10, M2F_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library contains sm_100a code only", dev, prop.major,
                prop.minor);
    return M2F_OK;
}

extern "C" int m2f_split_planes(const float* src, int64_t src_ld, const m2f_planes_t* dst, void* stream) {
    M2F_REQUIRE(src != nullptr, M2F_ERR_ARG, "split_planes: null src");
    M2F_TRY(check_planes(dst, "split_planes dst", 1));
    auto st = reinterpret_cast<cudaStream_t>(stream);
    auto* d = reinterpret_cast<__nv_bfloat16*>(dst->ptr);
    const bool vec = (dst->cols % 4) == 0 && (src_ld % 4) == 0 && (dst->ld % 4) == 0 && (dst->planes == 1 || (dst->plane_stride % 4) == 0) &&
                     (reinterpret_cast<uintptr_t>(src) % 16) == 0 && (reinterpret_cast<uintptr_t>(d) % 8) == 0;
    if (vec) {
        const int64_t total = dst->rows * (dst->cols / 4);
        split_planes_kernel<<<grid_for(total, 256), 256, 0, st>>>(src, src_ld, d, dst->ld, dst->plane_stride, dst->planes, dst->rows,
                                                                   dst->cols / 4);
    } else {
        const int64_t total = dst->rows * dst->cols;
        split_planes_scalar_kernel<<<grid_for(total, 256), 256, 0, st>>>(src, src_ld, d, dst->ld, dst->plane_stride, dst->planes,
                                                                          dst->rows, dst->cols);
    }
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_relu_planes(const m2f_planes_t* src, const m2f_planes_t* dst, void* stream) {
    M2F_TRY(check_planes(src, "relu_planes src", 4));
    M2F_TRY(check_planes(dst, "relu_planes dst", 4));
    M2F_REQUIRE(src->rows == dst->rows && src->cols == dst->cols && src->planes == dst->planes && (src->cols % 4) == 0, M2F_ERR_ARG,
                "relu_planes: extents differ or cols not a multiple of 4");
    const int64_t total = src->rows * (src->cols / 4);
    relu_planes_kernel<<<grid_for(total, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(src->ptr), src->ld, src->plane_stride, src->planes,
        reinterpret_cast<__nv_bfloat16*>(dst->ptr), dst->ld, dst->plane_stride, src->rows, src->cols / 4);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_colsum_planes(const m2f_planes_t* src, float* out, void* stream) {
    M2F_TRY(check_planes(src, "colsum_planes src", 1));
    M2F_REQUIRE(out != nullptr, M2F_ERR_ARG, "colsum_planes: null out");
    int64_t gy = (src->rows + 63) / 64;
    if (gy > 64) gy = 64;
    dim3 grid((unsigned)((src->cols + 31) / 32), (unsigned)gy), block(32, 8);
    colsum_planes_kernel<<<grid, block, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(src->ptr), src->ld, src->plane_stride, src->planes, src->rows, src->cols, out);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int m2f_dropout_mask(float p, uint64_t seed, uint64_t step, uint32_t site, int64_t n, uint8_t* mask, void* stream) {
    M2F_REQUIRE(mask != nullptr && n > 0 && p >= 0.f && p < 1.f, M2F_ERR_ARG, "dropout_mask: bad argument");
    Drop d = drop_make(p, seed, step, site);
    d.scale = 1.f;  // p == 0 must still report "kept"
    if (p == 0.f) d.thresh = 0;
    dropout_mask_kernel<<<grid_for(n, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(d, n, mask);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
