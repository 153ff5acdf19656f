// Gradient exchange of the data-parallel step as one kernel over NVLink 5 / NVSwitch peer memory.
//
// The reference has no multi-GPU path (src/train.py:20 pins cuda:0); the exchange the sharded hot path needs is the
// all-reduce of the gradient arena between the backward (src/train.py:230) and optimizer.step() (src/train.py:231).
//
// Every rank's gradient arena lives in a *symmetric* allocation: the same buffer is mapped on every GPU (peer_base[r])
// and, where the fabric supports it, once more through an NVSwitch multicast object (mc_base).  A range [lo, hi) is
// reduced in two shots by a few CTAs per GPU:
//   1. barrier: block b of every rank has started  => every rank's producers of the range have completed
//   2. rank r owns slice r of the range: `multimem.ld_reduce.add.v4.f32` returns the sum over all GPUs computed *inside
//      the switch* (one load, one NVLink response), the sum is scaled (1/world) and `multimem.st` writes it back to every
//      GPU through the same multicast mapping.  Without multicast: world peer loads + world peer stores per element.
//   3. barrier: block b of every rank has stored  => when the kernel has completed on a rank, all of its range is final.
//      Within a sequence of calls on one stream only the last call needs it (M2F_XCHG_NO_EXIT_BARRIER on the others).
// SMs pull and push range / world bytes per GPU; the switch does the fan-in / fan-out, and the NVLink traffic of the range
// (every GPU serves the whole range and receives the whole reduced range) is what bounds the kernel: 8 CTAs saturate it, the
// default is at most 16, so it runs beside the backward's GEMMs at the cost of a few of their CTA slots (NCCL's ring /
// tree kernels take many more).
//
// Barrier words: uint32 [M2F_XCHG_MAX_CTAS][M2F_XCHG_MAX_RANKS] at signal_off_bytes inside the symmetric buffer, zero
// before the first call.  Slot (b, s) on rank d is written by rank s only: put = CAS 0 -> 1 (release, system scope, spins
// while the previous signal has not been consumed), wait = CAS 1 -> 0 (acquire) on the local copy; the protocol resets
// itself, so calls can follow each other (and be replayed from CUDA graphs) without host involvement.
#include <stdio.h>

#include "common.cuh"

namespace m2f {

struct PeerPtrs {
    float* base[M2F_XCHG_MAX_RANKS];
    uint32_t* sig[M2F_XCHG_MAX_RANKS];
};

__device__ __forceinline__ uint64_t gtimer() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* a, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(a), "r"(cmp), "r"(val) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* a, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(a), "r"(cmp), "r"(val) : "memory");
    return old;
}

// A peer that never arrives (crashed rank, mismatched call sequence) must not hang the GPU for good: give up after 2 minutes
// (long enough for the start-up skew of 8 processes sharing a few host cores).
constexpr uint64_t XCHG_TIMEOUT_NS = 120ull * 1000 * 1000 * 1000;

// All threads of the block call this.  Threads [0, world) each handshake with one peer.
__device__ __forceinline__ void block_barrier_all_ranks(const PeerPtrs& pp, int rank, int world) {
    __syncthreads();  // the block's earlier global writes happen-before the release below (cumulativity through bar.sync)
    const int t = threadIdx.x;
    if (t < world) {
        const int slot_base = blockIdx.x * M2F_XCHG_MAX_RANKS;
        const uint64_t t0 = gtimer();
        uint32_t* theirs = pp.sig[t] + slot_base + rank;  // my slot on peer t
        while (cas_release_sys(theirs, 0u, 1u) != 0u) {
            if (gtimer() - t0 > XCHG_TIMEOUT_NS) {
                printf("m2f exchange: rank %d block %d timed out signalling rank %d\n", rank, (int)blockIdx.x, t);
                asm volatile("trap;");
            }
        }
        uint32_t* mine = pp.sig[rank] + slot_base + t;  // peer t's slot on me
        while (cas_acquire_sys(mine, 1u, 0u) != 1u) {
            if (gtimer() - t0 > XCHG_TIMEOUT_NS) {
                printf("m2f exchange: rank %d block %d timed out waiting for rank %d\n", rank, (int)blockIdx.x, t);
                asm volatile("trap;");
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ float4 mc_ld_reduce_v4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st_v4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float mc_ld_reduce_f32(const float* mc) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st_f32(float* mc, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ float4 peer_ld_v4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void peer_st_v4(float* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float peer_ld_f32(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void peer_st_f32(float* p, float v) { asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }

// Debug trace (always on, one thread per launch): ring of {t_enter, t_after_first_barrier, t_data_done, t_exit (globaltimer ns),
// lo, floats} so the exchange can be placed on the step's timeline (tools/dp_trace.py, M2F_TIMELINE).
__device__ unsigned long long g_xchg_trace[M2F_XCHG_TRACE_RECORDS * 6];
__device__ unsigned int g_xchg_trace_n;

// float offsets: [lo, lo + head) scalar, then n4 float4, then `tail` scalars; head / tail < 4 (arena ranges that do not
// start or end on 16 bytes, e.g. the 7-wide classifier bias at the end of the 1-D region)
// THREADS x UNROLL x 16 B of loads are in flight per CTA (512 x 8: 64 KB)
template <bool MC, int XCHG_THREADS, int XCHG_UNROLL>
__global__ void __launch_bounds__(XCHG_THREADS)
grad_allreduce_kernel(float* __restrict__ mc, PeerPtrs pp, int rank, int world, int64_t lo, int head, int64_t n4, int tail, float scale, int flags) {
    const bool tracer = blockIdx.x == 0 && threadIdx.x == 0;
    unsigned long long* tr = nullptr;
    if (tracer) {
        tr = g_xchg_trace + (size_t)(atomicAdd(&g_xchg_trace_n, 1u) % M2F_XCHG_TRACE_RECORDS) * 6;
        tr[0] = gtimer();
        tr[4] = (unsigned long long)lo;
        tr[5] = (unsigned long long)(head + 4 * n4 + tail);
    }
    block_barrier_all_ranks(pp, rank, world);
    if (tracer) tr[1] = gtimer();

    // the slice of the vector body this rank reduces, split over the blocks in contiguous runs of XCHG_UNROLL vectors per thread
    const int64_t per_rank = (n4 + world - 1) / world;
    const int64_t v_lo = min(n4, per_rank * rank), v_hi = min(n4, v_lo + per_rank);
    const int64_t body = lo + head;  // float offset of vector 0
    const int64_t stride = (int64_t)gridDim.x * XCHG_THREADS;
    for (int64_t i0 = v_lo + blockIdx.x * (int64_t)XCHG_THREADS + threadIdx.x; i0 < v_hi; i0 += stride * XCHG_UNROLL) {
        float4 acc[XCHG_UNROLL];
#pragma unroll
        for (int u = 0; u < XCHG_UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < v_hi) {
                if (MC) {
                    acc[u] = mc_ld_reduce_v4(mc + body + 4 * i);
                } else {
                    acc[u] = peer_ld_v4(pp.base[0] + body + 4 * i);
                    for (int r = 1; r < world; ++r) {
                        const float4 x = peer_ld_v4(pp.base[r] + body + 4 * i);
                        acc[u].x += x.x; acc[u].y += x.y; acc[u].z += x.z; acc[u].w += x.w;
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < XCHG_UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < v_hi) {
                float4 v = acc[u];
                v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
                if (MC) {
                    mc_st_v4(mc + body + 4 * i, v);
                } else {
                    for (int r = 0; r < world; ++r) peer_st_v4(pp.base[r] + body + 4 * i, v);
                }
            }
        }
    }
    // unaligned ends: a handful of scalars, done by the last rank's first block
    if (rank == world - 1 && blockIdx.x == 0 && (int)threadIdx.x < head + tail) {
        const int k = threadIdx.x;
        const int64_t off = (k < head) ? lo + k : body + 4 * n4 + (k - head);
        float v;
        if (MC) {
            v = mc_ld_reduce_f32(mc + off) * scale;
            mc_st_f32(mc + off, v);
        } else {
            v = 0.f;
            for (int r = 0; r < world; ++r) v += peer_ld_f32(pp.base[r] + off);
            v *= scale;
            for (int r = 0; r < world; ++r) peer_st_f32(pp.base[r] + off, v);
        }
    }
    __threadfence_system();
    if (tracer) tr[2] = gtimer();
    // Exit barrier: "every rank has stored its slice into my copy".  A sequence of exchanges on one stream needs it on the last
    // call only: a rank's stores of call k are complete before its kernel of call k+1 starts, so the last call's handshake
    // vouches for all of them (the ranges are disjoint and nobody reads them in between).
    if (!(flags & M2F_XCHG_NO_EXIT_BARRIER)) block_barrier_all_ranks(pp, rank, world);
    if (tracer) tr[3] = gtimer();
}

// ------------------------------------------------------------------------------------------ bf16 wire format
// The exchange is NVLink-bound for the whole backward (345 MB of fp32 gradients per step against a 1.5 ms backward,
// profiles/r2_exchange_sweep_2gpu.txt), so the lever is the byte count: with M2F_XCHG_WIRE_BF16 a range travels as bf16.
// Three launches on the caller's stream: (1) this rank's fp32 range -> its bf16 staging copy in the symmetric buffer (local
// HBM only); (2) the two-shot all-reduce on the staging copies -- `multimem.ld_reduce.add.acc::f32.v4.bf16x2`: the switch sums
// the ranks' bf16 values in fp32 and returns eight rounded sums per 16 bytes, `multimem.st` broadcasts them -- with the entry
// AND the closing handshake (the next launch reads every rank's slices); (3) staging -> this rank's fp32 range.  Half the NVLink
// bytes for 12 bytes per element of extra local traffic.  The reduced gradient carries one bf16 rounding of each rank's
// contribution and one of the sum (relative 2^-9 per element, unbiased): well inside the 2e-2 budget of the bf16 training
// mode, NOT for the fp32-parity mode (the host side only enables it for bf16 precision modes).
__global__ void __launch_bounds__(256)
wire_pack_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n8) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 a = *reinterpret_cast<const float4*>(src + 8 * i), b = *reinterpret_cast<const float4*>(src + 8 * i + 4);
        const __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
        const __nv_bfloat162 p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
        uint4 o;
        o.x = *reinterpret_cast<const uint32_t*>(&p0); o.y = *reinterpret_cast<const uint32_t*>(&p1);
        o.z = *reinterpret_cast<const uint32_t*>(&p2); o.w = *reinterpret_cast<const uint32_t*>(&p3);
        *reinterpret_cast<uint4*>(dst + 8 * i) = o;
    }
}
__global__ void __launch_bounds__(256)
wire_unpack_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int64_t n8) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const uint4 v = *reinterpret_cast<const uint4*>(src + 8 * i);
        const __nv_bfloat162 p0 = *reinterpret_cast<const __nv_bfloat162*>(&v.x), p1 = *reinterpret_cast<const __nv_bfloat162*>(&v.y);
        const __nv_bfloat162 p2 = *reinterpret_cast<const __nv_bfloat162*>(&v.z), p3 = *reinterpret_cast<const __nv_bfloat162*>(&v.w);
        *reinterpret_cast<float4*>(dst + 8 * i) = make_float4(__low2float(p0), __high2float(p0), __low2float(p1), __high2float(p1));
        *reinterpret_cast<float4*>(dst + 8 * i + 4) = make_float4(__low2float(p2), __high2float(p2), __low2float(p3), __high2float(p3));
    }
}
__device__ __forceinline__ uint4 mc_ld_reduce_bf16x8(const __nv_bfloat16* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t scale_bf16x2(uint32_t w, float scale) {
    const __nv_bfloat162 p = *reinterpret_cast<const __nv_bfloat162*>(&w);
    const __nv_bfloat162 q = __floats2bfloat162_rn(__low2float(p) * scale, __high2float(p) * scale);
    return *reinterpret_cast<const uint32_t*>(&q);
}

// the all-reduce proper on the bf16 staging copies; wire = multicast address of staging element `lo`, n8 vectors of 8
template <int XCHG_THREADS, int XCHG_UNROLL>
__global__ void __launch_bounds__(XCHG_THREADS)
grad_allreduce_bf16_kernel(__nv_bfloat16* __restrict__ mc_wire, PeerPtrs pp, int rank, int world, int64_t lo, int64_t n8, float scale) {
    const bool tracer = blockIdx.x == 0 && threadIdx.x == 0;
    unsigned long long* tr = nullptr;
    if (tracer) {
        tr = g_xchg_trace + (size_t)(atomicAdd(&g_xchg_trace_n, 1u) % M2F_XCHG_TRACE_RECORDS) * 6;
        tr[0] = gtimer();
        tr[4] = (unsigned long long)lo;
        tr[5] = (unsigned long long)(8 * n8);
    }
    block_barrier_all_ranks(pp, rank, world);  // every rank's staging copy of the range is complete (its pack kernel has finished)
    if (tracer) tr[1] = gtimer();
    const int64_t per_rank = (n8 + world - 1) / world;
    const int64_t v_lo = min(n8, per_rank * rank), v_hi = min(n8, v_lo + per_rank);
    const int64_t stride = (int64_t)gridDim.x * XCHG_THREADS;
    for (int64_t i0 = v_lo + blockIdx.x * (int64_t)XCHG_THREADS + threadIdx.x; i0 < v_hi; i0 += stride * XCHG_UNROLL) {
        uint4 acc[XCHG_UNROLL];
#pragma unroll
        for (int u = 0; u < XCHG_UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < v_hi) acc[u] = mc_ld_reduce_bf16x8(mc_wire + 8 * i);
        }
#pragma unroll
        for (int u = 0; u < XCHG_UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < v_hi) {
                uint4 v = acc[u];
                v.x = scale_bf16x2(v.x, scale); v.y = scale_bf16x2(v.y, scale); v.z = scale_bf16x2(v.z, scale); v.w = scale_bf16x2(v.w, scale);
                mc_st_v4(reinterpret_cast<float*>(mc_wire + 8 * i),
                         make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w)));
            }
        }
    }
    __threadfence_system();
    if (tracer) tr[2] = gtimer();
    block_barrier_all_ranks(pp, rank, world);  // every rank has stored its slice into my staging copy: the unpack launch may read it
    if (tracer) tr[3] = gtimer();
}

}  // namespace m2f

using namespace m2f;

// launch shape (tools/exchange_sweep.py): every rank must use the same one
static int g_xchg_threads = 512, g_xchg_unroll = 8;

extern "C" int m2f_exchange_tune(int32_t threads, int32_t unroll) {
    const bool ok = (threads == 512 && (unroll == 8 || unroll == 4)) || (threads == 256 && (unroll == 8 || unroll == 16)) ||
                    (threads == 1024 && unroll == 4) || (threads == 128 && unroll == 16);
    M2F_REQUIRE(ok, M2F_ERR_UNSUPPORTED, "exchange_tune: (threads, unroll) must be one of (512,8) (512,4) (256,8) (256,16) (1024,4) (128,16)");
    g_xchg_threads = threads;
    g_xchg_unroll = unroll;
    return M2F_OK;
}

extern "C" int m2f_grad_allreduce_peer(const m2f_peer_group_t* g, int64_t lo, int64_t hi, float scale, int32_t ctas, int32_t flags,
                                       void* stream) {
    M2F_REQUIRE(g != nullptr, M2F_ERR_ARG, "grad_allreduce_peer: group is NULL");
    M2F_REQUIRE(g->world >= 1 && g->world <= M2F_XCHG_MAX_RANKS && g->rank >= 0 && g->rank < g->world, M2F_ERR_ARG,
                "grad_allreduce_peer: rank %d / world %d outside [1, %d]", g->rank, g->world, M2F_XCHG_MAX_RANKS);
    M2F_REQUIRE(lo >= 0 && hi >= lo, M2F_ERR_ARG, "grad_allreduce_peer: bad range [%lld, %lld)", (long long)lo, (long long)hi);
    M2F_REQUIRE(g->signal_off_bytes >= hi * 4 && (g->signal_off_bytes % 16) == 0, M2F_ERR_ARG,
                "grad_allreduce_peer: the signal area (offset %lld) must lie behind the reduced range and be 16-byte aligned",
                (long long)g->signal_off_bytes);
    if (hi == lo) return M2F_OK;
    PeerPtrs pp;
    for (int r = 0; r < M2F_XCHG_MAX_RANKS; ++r) {
        pp.base[r] = nullptr;
        pp.sig[r] = nullptr;
    }
    for (int r = 0; r < g->world; ++r) {
        M2F_REQUIRE(g->peer_base[r] != nullptr, M2F_ERR_ARG, "grad_allreduce_peer: peer_base[%d] is NULL", r);
        M2F_REQUIRE((reinterpret_cast<uintptr_t>(g->peer_base[r]) % 16) == 0, M2F_ERR_ALIGN, "grad_allreduce_peer: peer_base[%d] must be 16-byte aligned", r);
        pp.base[r] = reinterpret_cast<float*>(g->peer_base[r]);
        pp.sig[r] = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(g->peer_base[r]) + g->signal_off_bytes);
    }
    M2F_REQUIRE((reinterpret_cast<uintptr_t>(g->mc_base) % 16) == 0, M2F_ERR_ALIGN, "grad_allreduce_peer: mc_base must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    float* mc = reinterpret_cast<float*>(g->mc_base);
    // bf16 wire: needs the multicast mapping, a staging area, more than one rank and a range made of whole 16-byte bf16
    // vectors (every rank evaluates the same condition: it depends on the arguments only); anything else travels as fp32
    if ((flags & M2F_XCHG_WIRE_BF16) && mc && g->world > 1 && g->wire_off_bytes > 0 && (lo & 7) == 0 && ((hi - lo) & 7) == 0) {
        M2F_REQUIRE(g->wire_off_bytes >= g->signal_off_bytes + M2F_XCHG_SIGNAL_BYTES && (g->wire_off_bytes % 16) == 0, M2F_ERR_ARG,
                    "grad_allreduce_peer: the bf16 staging area (offset %lld) must lie behind the signal words and be 16-byte aligned",
                    (long long)g->wire_off_bytes);
        const int64_t n8 = (hi - lo) / 8;
        float* mine = pp.base[g->rank];
        __nv_bfloat16* my_wire = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<uint8_t*>(mine) + g->wire_off_bytes) + lo;
        __nv_bfloat16* mc_wire = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<uint8_t*>(mc) + g->wire_off_bytes) + lo;
        int64_t cg = (n8 + 255) / 256;
        if (cg > 148 * 2) cg = 148 * 2;
        wire_pack_kernel<<<(unsigned)cg, 256, 0, st>>>(mine + lo, my_wire, n8);
        M2F_LAUNCH_CHECK();
        const int64_t per_rank_w = (n8 + g->world - 1) / g->world;
        const int64_t per_iter_w = (int64_t)512 * 8;
        int grid_w;
        if (ctas > 0) grid_w = ctas;
        else {
            const int64_t want = (per_rank_w + 2 * per_iter_w - 1) / (2 * per_iter_w);
            grid_w = want < 1 ? 1 : (want > 16 ? 16 : (int)want);
        }
        if (grid_w > M2F_XCHG_MAX_CTAS) grid_w = M2F_XCHG_MAX_CTAS;
        const int64_t need_w = (per_rank_w + per_iter_w - 1) / per_iter_w;
        if (need_w < grid_w) grid_w = need_w < 1 ? 1 : (int)need_w;
        static const int prio_w = [] {
            int lo_p = 0, hi_p = 0;
            return cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p) == cudaSuccess ? hi_p : 0;
        }();
        cudaLaunchAttribute attr_w[1];
        attr_w[0].id = cudaLaunchAttributePriority;
        attr_w[0].val.priority = prio_w;
        cudaLaunchConfig_t cfg_w = {};
        cfg_w.gridDim = dim3((unsigned)grid_w);
        cfg_w.blockDim = dim3(512);
        cfg_w.stream = st;
        cfg_w.attrs = attr_w;
        cfg_w.numAttrs = 1;
        const int rank_w = g->rank, world_w = g->world;
        M2F_CUDA(cudaLaunchKernelEx(&cfg_w, grad_allreduce_bf16_kernel<512, 8>, mc_wire, pp, rank_w, world_w, lo, n8, scale));
        M2F_LAUNCH_CHECK();
        wire_unpack_kernel<<<(unsigned)cg, 256, 0, st>>>(my_wire, mine + lo, n8);
        M2F_LAUNCH_CHECK();
        return M2F_OK;
    }
    const int head = (int)((4 - (lo & 3)) & 3) < (hi - lo) ? (int)((4 - (lo & 3)) & 3) : (int)(hi - lo);
    const int64_t n4 = (hi - lo - head) / 4;
    const int tail = (int)(hi - lo - head - 4 * n4);
    // Measured (tools/exchange_sweep.py, profiles/r1_exchange_sweep_2gpu.txt): the multicast path is bound by the NVLink
    // traffic of the range (every GPU serves the whole range to the switch and receives the whole reduced range back:
    // ~400 GB/s of range bytes at 2 GPUs) and 8 CTAs of 512 threads x 8 vectors already saturate it; more CTAs only displace
    // more of the backward's GEMM CTAs (those fill the register file, so every resident exchange CTA costs one).  So: at
    // most 16 CTAs, each making at least two passes.  Every rank must use the same count (the barrier slots are per block):
    // it depends on the range, the world size and `ctas` only.
    const int64_t per_rank = (n4 + g->world - 1) / g->world;
    const int threads = g_xchg_threads, unroll = g_xchg_unroll;
    const int64_t per_iter = (int64_t)threads * unroll;  // vectors one CTA moves per loop iteration
    int grid;
    if (ctas > 0) {
        grid = ctas;
    } else {
        const int64_t want = (per_rank + 2 * per_iter - 1) / (2 * per_iter);
        grid = want < 1 ? 1 : (want > 16 ? 16 : (int)want);
    }
    if (grid > M2F_XCHG_MAX_CTAS) grid = M2F_XCHG_MAX_CTAS;
    const int64_t need = (per_rank + per_iter - 1) / per_iter;
    if (need < grid) grid = need < 1 ? 1 : (int)need;
    // The few exchange CTAs go in front of the backward's pending GEMM CTAs: the priority is set on the launch itself, so it
    // also holds for the kernel nodes of a captured graph, whatever stream they were captured from.
    static const int prio = [] {
        int lo_p = 0, hi_p = 0;
        return cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p) == cudaSuccess ? hi_p : 0;
    }();
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributePriority;
    attr[0].val.priority = prio;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    float* mc_arg = mc;
    const int rank = g->rank, world = g->world;
#define M2F_XCHG_LAUNCH(T, U)                                                                                                          \
    if (threads == T && unroll == U) {                                                                                                 \
        cfg.blockDim = dim3(T);                                                                                                        \
        if (mc) M2F_CUDA(cudaLaunchKernelEx(&cfg, grad_allreduce_kernel<true, T, U>, mc_arg, pp, rank, world, lo, head, n4, tail, scale, (int)flags));  \
        else    M2F_CUDA(cudaLaunchKernelEx(&cfg, grad_allreduce_kernel<false, T, U>, mc_arg, pp, rank, world, lo, head, n4, tail, scale, (int)flags)); \
    }
    M2F_XCHG_LAUNCH(512, 8)
    M2F_XCHG_LAUNCH(512, 4)
    M2F_XCHG_LAUNCH(256, 8)
    M2F_XCHG_LAUNCH(256, 16)
    M2F_XCHG_LAUNCH(1024, 4)
    M2F_XCHG_LAUNCH(128, 16)
#undef M2F_XCHG_LAUNCH
    M2F_LAUNCH_CHECK();
    return M2F_OK;
}

extern "C" int64_t m2f_exchange_trace(uint64_t* out_records) {
    if (!out_records) return 0;
    unsigned int n = 0;
    if (cudaMemcpyFromSymbol(&n, g_xchg_trace_n, sizeof(n)) != cudaSuccess) return 0;
    if (cudaMemcpyFromSymbol(out_records, g_xchg_trace, sizeof(unsigned long long) * M2F_XCHG_TRACE_RECORDS * 6) != cudaSuccess) return 0;
    return (int64_t)n;
}
