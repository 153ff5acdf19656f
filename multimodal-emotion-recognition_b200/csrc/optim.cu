// Fused Adam over the flat parameter arena.  torch.optim.Adam semantics as the reference
// configures it (src/train.py:56: lr, weight_decay as L2 folded into the gradient, betas (0.9, 0.999), eps 1e-8,
// amsgrad off): g += wd * p; m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2;
// p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps).
#include <math.h>

#include "common.cuh"

namespace m2f {

__global__ void __launch_bounds__(256)
adam_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v, int64_t n4, float b1,
            float b2, float eps, float wd, float step_size, float inv_sqrt_bc2, float grad_scale) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 pp = p[i], gg = g[i], mm = m[i], vv = v[i];
        float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float gr = ga[k] * grad_scale + wd * pa[k];
            ma[k] = b1 * ma[k] + (1.f - b1) * gr;
            va[k] = b2 * va[k] + (1.f - b2) * gr * gr;
            const float denom = sqrtf(va[k]) * inv_sqrt_bc2 + eps;
            pa[k] -= step_size * (ma[k] / denom);
        }
        p[i] = pp; m[i] = mm; v[i] = vv;
    }
}

}  // namespace m2f

using namespace m2f;

extern "C" int m2f_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                             float beta2, float eps, float weight_decay, int64_t t, float grad_scale, void* stream) {
    M2F_REQUIRE(params && grads && exp_avg && exp_avg_sq && n > 0 && t >= 1, M2F_ERR_ARG, "adam_step: bad argument");
    M2F_REQUIRE((n % 4) == 0, M2F_ERR_ALIGN, "adam_step: n must be a multiple of 4 (the arena is padded to 64 floats)");
    M2F_REQUIRE(((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) | reinterpret_cast<uintptr_t>(exp_avg) |
                  reinterpret_cast<uintptr_t>(exp_avg_sq)) % 16) == 0, M2F_ERR_ALIGN, "adam_step: pointers must be 16-byte aligned");
    const double bc1 = 1.0 - pow((double)beta1, (double)t), bc2 = 1.0 - pow((double)beta2, (double)t);
    const float step_size = (float)((double)lr / bc1), inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    const int64_t n4 = n / 4;
    int64_t g = (n4 + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    adam_kernel<<<(unsigned)g, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<float4*>(params), reinterpret_cast<const float4*>(grads), reinterpret_cast<float4*>(exp_avg),
        reinterpret_cast<float4*>(exp_avg_sq), n4, beta1, beta2, eps, weight_decay, step_size, inv_sqrt_bc2, grad_scale);
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}
