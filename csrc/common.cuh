// Shared device helpers: error plumbing, system-scope memory-model primitives, block reductions,
// the peer-memory block layout.  sm_100a only.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/psb200.h"

// ------------------------------------------------------------------ host error plumbing
void psb_set_error(const char* fmt, ...);
#define PSB_CHECK_CUDA(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            psb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return PSB_E_CUDA;                                                                 \
        }                                                                                      \
    } while (0)
#define PSB_REQUIRE(cond, ...)                                                                 \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            psb_set_error(__VA_ARGS__);                                                        \
            return PSB_E_INVALID;                                                              \
        }                                                                                      \
    } while (0)

// ------------------------------------------------------------------ peer-memory block layout
// One block per rank, identical layout everywhere (a "symmetric heap" of one object).
struct CommLayout {
    int nv;            // floats in the parameter vector: (F+1)*8
    int n4;            // float4s
    int slice4;        // float4s per shard
    int num_ranks;
    // byte offsets
    size_t off_status, off_arrivals, off_ticket, off_version, off_clocks;
    size_t off_rep, off_auth, off_acc, off_inbox, off_wclocks, off_pub, bytes;

    __host__ __device__ static CommLayout make(int num_features, int num_ranks) {
        CommLayout L;
        L.nv = (num_features + 1) * PSB_CLASS_PAD;
        L.n4 = L.nv / 4;
        L.slice4 = (L.n4 + num_ranks - 1) / num_ranks;
        L.num_ranks = num_ranks;
        size_t o = 0;
        L.off_status = o;   o += 256;                     // uint32[64]
        L.off_arrivals = o; o += 2 * 128;                 // one 128-byte line per parity
        L.off_ticket = o;   o += 2 * 128;
        L.off_version = o;  o += 128;                     // uint32[PSB_MAX_RANKS]
        L.off_clocks = o;   o += 128;
        L.off_rep = o;      o += (size_t)2 * L.nv * 4;
        L.off_auth = o;     o += (size_t)L.slice4 * 16;
        L.off_acc = o;      o += (size_t)2 * L.nv * 4;
        L.off_inbox = o;    o += (size_t)2 * num_ranks * L.slice4 * 16;
        o = (o + 127) & ~(size_t)127;
        L.off_wclocks = o;  o += (size_t)PSB_MAX_WORKERS * 4;  // uint32[PSB_MAX_WORKERS]: per-worker clocks
        L.off_pub = o;      o += (size_t)L.nv * 4;             // published snapshot (seqlock word: status()[8])
        L.bytes = (o + 255) & ~(size_t)255;
        return L;
    }
    __host__ __device__ int slice_begin(int s) const { return s * slice4 < n4 ? s * slice4 : n4; }
    __host__ __device__ int slice_end(int s) const { return (s + 1) * slice4 < n4 ? (s + 1) * slice4 : n4; }
};

struct CommBlock {
    char* base;
    CommLayout L;
    __host__ __device__ CommBlock(void* p, const CommLayout& l) : base((char*)p), L(l) {}
    __host__ __device__ unsigned int* status() const { return (unsigned int*)(base + L.off_status); }
    __host__ __device__ unsigned int* arrivals(int parity) const { return (unsigned int*)(base + L.off_arrivals + parity * 128); }
    __host__ __device__ unsigned int* ticket(int parity) const { return (unsigned int*)(base + L.off_ticket + parity * 128); }
    __host__ __device__ unsigned int* version() const { return (unsigned int*)(base + L.off_version); }
    __host__ __device__ unsigned int* clocks() const { return (unsigned int*)(base + L.off_clocks); }
    __host__ __device__ unsigned int* wclocks() const { return (unsigned int*)(base + L.off_wclocks); }
    __host__ __device__ float4* pub() const { return (float4*)(base + L.off_pub); }
    __host__ __device__ unsigned int* pub_seq() const { return status() + 8; }
    __host__ __device__ float4* rep(int parity) const { return (float4*)(base + L.off_rep) + (size_t)parity * L.n4; }
    __host__ __device__ float4* auth() const { return (float4*)(base + L.off_auth); }
    __host__ __device__ float4* acc(int parity) const { return (float4*)(base + L.off_acc) + (size_t)parity * L.n4; }
    __host__ __device__ float4* inbox(int parity, int src) const {
        return (float4*)(base + L.off_inbox) + ((size_t)parity * L.num_ranks + src) * L.slice4;
    }
};

#ifdef __CUDACC__
// ------------------------------------------------------------------ memory-model primitives
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_relaxed_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int atom_add_acqrel_sys(unsigned int* p, unsigned int v) {
    unsigned int old;
    asm volatile("atom.acq_rel.sys.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ unsigned int atom_add_acqrel_gpu(unsigned int* p, unsigned int v) {
    unsigned int old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// vectorised fire-and-forget reduction (sm_90+): one 16-byte RED per four parameters
__device__ __forceinline__ void red_add_v4(float4* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ void red_add_v4_sys(float4* p, float4 v) {
    asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}
// loads that must observe other SMs' / GPUs' writes: bypass the (non-coherent) L1
__device__ __forceinline__ float4 ld_cg_v4(const float4* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_v4(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void st_cg_v4(float4* p, float4 v) {
    asm volatile("st.global.cg.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// Spin until *p >= want (acquire, system scope).  Gives up after `limit` cycles and raises the
// block's error flag: a hung peer must not hang this GPU (watchdog).
__device__ __forceinline__ bool wait_ge_sys(const unsigned int* p, unsigned int want, unsigned long long limit,
                                            unsigned int* err_flag, unsigned int err_code) {
    unsigned long long t0 = clock64();
    while ((int)(ld_acquire_sys(p) - want) < 0) {
        if (limit && (unsigned long long)(clock64() - t0) > limit) {
            atomicExch(err_flag, err_code);
            return false;
        }
        __nanosleep(64);
    }
    return true;
}

// ------------------------------------------------------------------ debug timeline (block 0, -DPSB_TC_TRACE only)
#ifdef PSB_TC_TRACE
__shared__ long long psb_trace_t[48];
__shared__ int psb_trace_id[48];
__shared__ int psb_trace_n;
#define PSB_TRACE(id)                                                                   \
    do {                                                                                \
        if (threadIdx.x == 0 && blockIdx.x == 0 && psb_trace_n < 48) {                  \
            psb_trace_t[psb_trace_n] = clock64(), psb_trace_id[psb_trace_n] = (id), ++psb_trace_n; \
        }                                                                               \
    } while (0)
#else
#define PSB_TRACE(id) do { } while (0)
#endif

// ------------------------------------------------------------------ block reductions
// All threads get the same double.  `red` is shared scratch of >= 33 doubles.
__device__ __forceinline__ double block_sum(double v, double* red) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    if (w == 0) {
        double t = l < nw ? red[l] : 0.0;
        for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (l == 0) red[32] = t;
    }
    __syncthreads();
    return red[32];
}
// Ask L1 for this thread's strides of a parameter-sized vector.  The solver's vectors live in global memory and a
// pass over one is a chain of ~8 dependent load-use round trips per thread; the prefetches put those round trips
// in flight together without holding registers (batching the loads in registers spills: the kernels that use
// these helpers run at their register cap).
__device__ __forceinline__ void vec_prefetch(const float* p, int n4) {
    const float4* p4 = (const float4*)p;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) asm volatile("prefetch.global.L1 [%0];" ::"l"(p4 + i));
}

__device__ __forceinline__ double block_dot(const float* a, const float* b, int n4, double* red) {
    vec_prefetch(a, n4);
    if (b != a) vec_prefetch(b, n4);
    float s = 0.f;
    const float4* a4 = (const float4*)a;
    const float4* b4 = (const float4*)b;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 x = a4[i], y = b4[i];
        s += x.x * y.x + x.y * y.y + x.z * y.z + x.w * y.w;
    }
    return block_sum((double)s, red);
}
#endif
