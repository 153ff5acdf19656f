// Shared device/host helpers for the M2FNet sm_100a kernels.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/m2f_b200.h"

namespace m2f {

// ------------------------------------------------------------------ error plumbing (host)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define M2F_CUDA(expr)                                                              \
    do {                                                                            \
        cudaError_t _e = (expr);                                                    \
        if (_e != cudaSuccess) return ::m2f::cuda_fail(_e, #expr, __FILE__, __LINE__); \
    } while (0)

#define M2F_REQUIRE(cond, code, ...)              \
    do {                                          \
        if (!(cond)) {                            \
            ::m2f::set_error(__VA_ARGS__);        \
            return (code);                        \
        }                                         \
    } while (0)

#define M2F_TRY(expr)               \
    do {                            \
        int _r = (expr);            \
        if (_r != M2F_OK) return _r; \
    } while (0)

// after a kernel launch: catches bad launch configurations without synchronising
void note_launch();
void note_launches(long long n);  // kernels replayed by a CUDA graph launch
#define M2F_LAUNCH_CHECK()                  \
    do {                                    \
        ::m2f::note_launch();               \
        M2F_CUDA(cudaPeekAtLastError());    \
    } while (0)

int check_planes(const m2f_planes_t* t, const char* name, int vec);

// ------------------------------------------------------------------ programmatic dependent launch
// Kernels of the step's dependency chains are launched with programmatic stream serialization: the next kernel's
// CTAs become resident and run their prologue while the previous kernel drains; `pdl_wait()` (first thing before any
// global-memory access) blocks until the previous kernel has completed and its writes are visible, `pdl_trigger()`
// lets the kernel after this one be scheduled.  Both are no-ops for a normal launch.  M2F_PDL=0 turns the attribute off.
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <class... KArgs, class... Args>
inline cudaError_t launch_chain(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int n = 0;
    if (pdl_enabled()) {
        at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    if (cluster_x > 1) {
        at[n].id = cudaLaunchAttributeClusterDimension;
        at[n].val.clusterDim.x = (unsigned)cluster_x;
        at[n].val.clusterDim.y = 1;
        at[n].val.clusterDim.z = 1;
        ++n;
    }
    cfg.attrs = at;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// ------------------------------------------------------------------ Philox4x32-7
// Counter layout: (idx8_lo, idx8_hi, site, step_lo) ; key: (seed_lo, seed_hi ^ step_hi).
// One call yields 128 bits = eight 16-bit lanes: the keep/drop decisions of elements 8*idx8 .. 8*idx8+7 of a site (lane k of
// element 8*idx8 + k is bits 16*(k&1).. of word k>>1).  Seven rounds are the fewest with which Philox4x32 passes BigCrush
// (Salmon et al., SC'11; ten is that paper's safety margin): the dropout-carrying GEMM epilogues and LayerNorm kernels are
// instruction-issue bound and the generator was most of their instructions (80 per 4 elements -> 56 per 8).
struct Philox {
    uint32_t k0, k1, site, step;
};

__host__ __device__ __forceinline__ Philox philox_make(uint64_t seed, uint64_t step, uint32_t site) {
    Philox p;
    p.k0 = (uint32_t)seed;
    p.k1 = (uint32_t)(seed >> 32) ^ (uint32_t)(step >> 32);
    p.site = site;
    p.step = (uint32_t)step;
    return p;
}

__host__ __device__ __forceinline__ void philox4(const Philox& p, uint64_t idx8, uint32_t out[4]) {
    uint32_t c0 = (uint32_t)idx8, c1 = (uint32_t)(idx8 >> 32), c2 = p.site, c3 = p.step;
    uint32_t k0 = p.k0, k1 = p.k1;
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1;
        c3 = (uint32_t)p0;
        c0 = n0;
        c2 = n2;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// keep iff lane >= thresh where thresh = floor(p * 2^16)  (P(keep) = 1 - p to within 2^-16)
__host__ __device__ __forceinline__ uint32_t drop_threshold(float p) {
    double t = (double)p * 65536.0;
    if (t <= 0.0) return 0u;
    if (t >= 65535.0) return 0xFFFFu;
    return (uint32_t)t;
}

// Dropout descriptor passed by value into kernels.  p == 0 => disabled.
struct Drop {
    Philox ph;
    uint32_t thresh;
    float scale;  // 1 / (1 - p)
    int enabled;
    const unsigned long long* rng;  // device {seed, step} that replaces the by-value pair (m2f_set_rng_source), or NULL
};

// m2f_set_rng_source: set per host thread; every dropout launch made while it is non-NULL reads (seed, step) from there
const unsigned long long* rng_source();

__host__ __device__ __forceinline__ Drop drop_make_raw(float p, uint64_t seed, uint64_t step, uint32_t site) {
    Drop d;
    d.ph = philox_make(seed, step, site);
    d.thresh = drop_threshold(p);
    d.scale = p < 1.f ? 1.f / (1.f - p) : 0.f;
    d.enabled = p > 0.f;
    d.rng = nullptr;
    return d;
}
static inline Drop drop_make(float p, uint64_t seed, uint64_t step, uint32_t site) {
    Drop d = drop_make_raw(p, seed, step, site);
    d.rng = rng_source();
    return d;
}
// first thing a kernel does with its descriptor: pick up the device-resident (seed, step) if there is one
__device__ __forceinline__ Drop drop_resolve(Drop d) {
    if (d.enabled && d.rng) {
        const unsigned long long seed = __ldg(d.rng), step = __ldg(d.rng + 1);
        d.ph = philox_make(seed, step, d.ph.site);
    }
    return d;
}

// multiplier (0 or scale) for the single element `idx` (slow path: one Philox per element)
__device__ __forceinline__ float drop_mult1(const Drop& d, uint64_t idx) {
    uint32_t r[4];
    philox4(d.ph, idx >> 3, r);
    const uint32_t k = (uint32_t)idx & 7u;
    const uint32_t w = (k & 4u) ? ((k & 2u) ? r[3] : r[2]) : ((k & 2u) ? r[1] : r[0]);
    return ((w >> (16u * (k & 1u))) & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
}

// multipliers for the aligned group of four elements 4*idx4 .. 4*idx4+3 (half of a Philox call: the neighbouring group
// recomputes it -- kernels whose threads own 8 consecutive elements use drop_mult8)
__device__ __forceinline__ void drop_mult4(const Drop& d, uint64_t idx4, float m[4]) {
    uint32_t r[4];
    philox4(d.ph, idx4 >> 1, r);
    const bool hi = (idx4 & 1u) != 0;
    const uint32_t a = hi ? r[2] : r[0], b = hi ? r[3] : r[1];
    m[0] = (a & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
    m[1] = (a >> 16) >= d.thresh ? d.scale : 0.f;
    m[2] = (b & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
    m[3] = (b >> 16) >= d.thresh ? d.scale : 0.f;
}

// Two aligned 4-groups of this lane, idx4_a and idx4_b, in a layout where the neighbouring lane (lane ^ 1) owns idx4_a ^ 1 and
// idx4_b ^ 1 and both groups' parity equals the lane's: the lane pair makes the two Philox calls once between them (even
// lane: a's, odd lane: b's) and swaps the halves the other needs -- half a call per group.  Every lane of the warp must call.
__device__ __forceinline__ void drop_mult4_pair(const Drop& d, uint64_t idx4_a, uint64_t idx4_b, int lane, float ma[4], float mb[4]) {
    const bool odd = (lane & 1) != 0;
    uint32_t r[4];
    philox4(d.ph, (odd ? idx4_b : idx4_a) >> 1, r);
    const uint32_t g0 = __shfl_xor_sync(0xffffffffu, odd ? r[0] : r[2], 1), g1 = __shfl_xor_sync(0xffffffffu, odd ? r[1] : r[3], 1);
    const uint32_t a0 = odd ? g0 : r[0], a1 = odd ? g1 : r[1], b0 = odd ? r[2] : g0, b1 = odd ? r[3] : g1;
    ma[0] = (a0 & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
    ma[1] = (a0 >> 16) >= d.thresh ? d.scale : 0.f;
    ma[2] = (a1 & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
    ma[3] = (a1 >> 16) >= d.thresh ? d.scale : 0.f;
    mb[0] = (b0 & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
    mb[1] = (b0 >> 16) >= d.thresh ? d.scale : 0.f;
    mb[2] = (b1 & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
    mb[3] = (b1 >> 16) >= d.thresh ? d.scale : 0.f;
}

// multipliers for the aligned group of eight elements 8*idx8 .. 8*idx8+7: one Philox call
__device__ __forceinline__ void drop_mult8(const Drop& d, uint64_t idx8, float m[8]) {
    uint32_t r[4];
    philox4(d.ph, idx8, r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        m[2 * i] = (r[i] & 0xFFFFu) >= d.thresh ? d.scale : 0.f;
        m[2 * i + 1] = (r[i] >> 16) >= d.thresh ? d.scale : 0.f;
    }
}

// ------------------------------------------------------------------ bf16 planes
// A value v is stored as P bf16 planes hi, mid, lo with hi + mid + lo ~= v (P = 1: plain bf16,
// P = 3: ~24 mantissa bits, the fp32-accurate mode).  Plane p lives at base + p * plane_stride.
__device__ __forceinline__ void split_planes(float v, int P, __nv_bfloat16 out[3]) {
    out[0] = __float2bfloat16_rn(v);
    if (P > 1) {
        float r = v - __bfloat162float(out[0]);
        out[1] = __float2bfloat16_rn(r);
        if (P > 2) {
            r -= __bfloat162float(out[1]);
            out[2] = __float2bfloat16_rn(r);
        }
    }
}

__device__ __forceinline__ void store_planes(__nv_bfloat16* base, int64_t plane_stride, int P, int64_t idx, float v) {
    __nv_bfloat16 h[3];
    split_planes(v, P, h);
    base[idx] = h[0];
    if (P > 1) base[plane_stride + idx] = h[1];
    if (P > 2) base[2 * plane_stride + idx] = h[2];
}

__device__ __forceinline__ float load_planes(const __nv_bfloat16* base, int64_t plane_stride, int P, int64_t idx) {
    float v = __bfloat162float(base[idx]);
    if (P > 1) v += __bfloat162float(base[plane_stride + idx]);
    if (P > 2) v += __bfloat162float(base[2 * plane_stride + idx]);
    return v;
}

// four consecutive elements (idx multiple of 4, 8-byte aligned rows)
__device__ __forceinline__ void load_planes4(const __nv_bfloat16* base, int64_t plane_stride, int P, int64_t idx, float v[4]) {
    v[0] = v[1] = v[2] = v[3] = 0.f;
    for (int p = P - 1; p >= 0; --p) {  // small terms first
        const uint2 raw = *reinterpret_cast<const uint2*>(base + p * plane_stride + idx);
        const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
        const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
        v[0] += __low2float(a); v[1] += __high2float(a);
        v[2] += __low2float(b); v[3] += __high2float(b);
    }
}

__device__ __forceinline__ void store_planes4(__nv_bfloat16* base, int64_t plane_stride, int P, int64_t idx, const float v[4]) {
    // registers only: the plane loop is unrolled with static indices (a rolled loop over a runtime P put the residuals in
    // local memory), each plane packs the rounded residuals two by two and subtracts them for the next one
    float r[4] = {v[0], v[1], v[2], v[3]};
#pragma unroll
    for (int p = 0; p < 3; ++p) {
        if (p < P) {
            const __nv_bfloat162 a = __floats2bfloat162_rn(r[0], r[1]);
            const __nv_bfloat162 b = __floats2bfloat162_rn(r[2], r[3]);
            uint2 raw;
            raw.x = *reinterpret_cast<const uint32_t*>(&a);
            raw.y = *reinterpret_cast<const uint32_t*>(&b);
            *reinterpret_cast<uint2*>(base + p * plane_stride + idx) = raw;
            if (p + 1 < P) {
                r[0] -= __low2float(a); r[1] -= __high2float(a);
                r[2] -= __low2float(b); r[3] -= __high2float(b);
            }
        }
    }
}

// ------------------------------------------------------------------ streaming accesses
// L2 eviction policy "evict first" for data that is touched once per step (optimizer state, master weights, consumed gradients):
// the optimizer pass streams 2.8 GB through a 126 MB L2 that, when the pass runs beside the backward (data parallel), also holds
// the operands the GEMMs re-read (activations of a layer, weight tiles shared by a column of CTAs).
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ float4 ldg_stream4(const float* ptr, uint64_t pol) {
    float4 v;
    asm volatile("ld.global.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(ptr), "l"(pol));
    return v;
}
__device__ __forceinline__ void stg_stream4(float* ptr, const float4 v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(ptr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}

// ------------------------------------------------------------------ warp helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace m2f
