// K1: device side of ingestion — scatter a staged batch of rows into the workers' dense fp16 ring
// slots.  The host (producer thread / DynamicBuffer) decides the slot of every row: the buffer policy
// of WorkerSamplingProcessor.java:50-113 collapses to "write at head, window = last count rows",
// so the device only needs (worker, slot) per row and the (head, count) pair per worker.
#include "common.cuh"

// staged row: word0 nnz, word1 label, word2 worker, word3 slot, then nnz x (int32 index, fp32 value)
__global__ void __launch_bounds__(256) psb_ring_scatter_kernel(const uint32_t* __restrict__ stage, int n,
                                                              int stride_words, __half* __restrict__ X,
                                                              int* __restrict__ labels, int cap, int F) {
    for (int r = blockIdx.x; r < n; r += gridDim.x) {
        const uint32_t* row = stage + (size_t)r * stride_words;
        const int nnz = (int)row[0], worker = (int)row[2], slot = (int)row[3];
        if (slot < 0) continue;                   // superseded by a later row of the same batch (block-uniform)
        __half* dst = X + ((size_t)worker * cap + slot) * F;
        // zero-fill then scatter: two phases, rows are short
        uint4* d16 = (uint4*)dst;
        for (int i = threadIdx.x; i < F / 8; i += blockDim.x) d16[i] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        for (int e = threadIdx.x; e < nnz; e += blockDim.x) {
            const int idx = (int)row[4 + 2 * e];
            const float v = __uint_as_float(row[5 + 2 * e]);
            if (idx >= 0 && idx < F) dst[idx] = __float2half_rn(v);
        }
        if (threadIdx.x == 0) labels[(size_t)worker * cap + slot] = (int)row[1];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) psb_ring_store_half_kernel(const __half* __restrict__ rows,
                                                                 const int* __restrict__ labels_in,
                                                                 const int* __restrict__ worker_slot, int n,
                                                                 __half* __restrict__ X, int* __restrict__ labels,
                                                                 int cap, int F) {
    for (int r = blockIdx.x; r < n; r += gridDim.x) {
        const int worker = worker_slot[2 * r], slot = worker_slot[2 * r + 1];
        if (slot < 0) continue;
        const uint4* src = (const uint4*)(rows + (size_t)r * F);
        uint4* dst = (uint4*)(X + ((size_t)worker * cap + slot) * F);
        for (int i = threadIdx.x; i < F / 8; i += blockDim.x) dst[i] = src[i];
        if (threadIdx.x == 0) labels[(size_t)worker * cap + slot] = labels_in[r];
    }
}

__global__ void __launch_bounds__(256) psb_ring_store_dense_kernel(const float* __restrict__ rows,
                                                                  const int* __restrict__ labels_in,
                                                                  const int* __restrict__ worker_slot, int n,
                                                                  __half* __restrict__ X, int* __restrict__ labels,
                                                                  int cap, int F) {
    for (int r = blockIdx.x; r < n; r += gridDim.x) {
        const int worker = worker_slot[2 * r], slot = worker_slot[2 * r + 1];
        if (slot < 0) continue;
        const float2* src = (const float2*)(rows + (size_t)r * F);
        __half2* dst = (__half2*)(X + ((size_t)worker * cap + slot) * F);
        for (int i = threadIdx.x; i < F / 2; i += blockDim.x) dst[i] = __float22half2_rn(src[i]);
        if (threadIdx.x == 0) labels[(size_t)worker * cap + slot] = labels_in[r];
    }
}

// Bitmap-packed rows (the sparse wire format of ingestion): per row F/32 presence words, an offset into one
// contiguous fp16 value stream, values in feature order.  One warp expands one row: lane l owns presence
// word l of every 32-word group, a warp scan of the popcounts gives it its position in the value stream, and
// it writes its 32 features (64 bytes) of the dense ring row.  Lossless with respect to the fp16 ring.
__global__ void __launch_bounds__(256) psb_ring_store_bitmap_kernel(const uint32_t* __restrict__ bitmap,
                                                                   const int* __restrict__ row_off,
                                                                   const __half* __restrict__ values,
                                                                   const int* __restrict__ labels_in,
                                                                   const int* __restrict__ worker_slot, int n,
                                                                   __half* __restrict__ X, int* __restrict__ labels,
                                                                   int cap, int F) {
    const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    const int words = F >> 5;
    for (int r = blockIdx.x * wpb + (threadIdx.x >> 5); r < n; r += gridDim.x * wpb) {
        const int worker = worker_slot[2 * r], slot = worker_slot[2 * r + 1];
        if (slot < 0) continue;                   // warp-uniform: one row per warp
        __half* dst = X + ((size_t)worker * cap + slot) * F;
        int base = row_off[r];
        for (int w0 = 0; w0 < words; w0 += 32) {
            const int w = w0 + lane;
            const uint32_t bits = w < words ? bitmap[(size_t)r * words + w] : 0u;
            const int cnt = __popc(bits);
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const __half* v = values + base + (incl - cnt);
            if (w < words) {
                __align__(16) __half out[32];
                int k = 0;
#pragma unroll
                for (int b = 0; b < 32; ++b) {
                    const bool on = (bits >> b) & 1u;
                    out[b] = on ? v[k] : __float2half_rn(0.f);
                    k += on;
                }
                uint4* d16 = (uint4*)(dst + (size_t)w * 32);
#pragma unroll
                for (int q = 0; q < 4; ++q) d16[q] = ((const uint4*)out)[q];
            }
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) labels[(size_t)worker * cap + slot] = labels_in[r];
    }
}

// Bitmap + per-row palette rows (psb_pack_palette_dense): one warp per row.  Lanes 0..15 hold the palette; lane l
// expands presence word l of every 32-word group: the warp scan of the popcounts gives it its first code index,
// every set bit fetches its nibble and the value comes from the palette lane by shuffle.  Raw rows (npal == 0)
// read their fp16 values directly.
__global__ void __launch_bounds__(256, 4) psb_ring_store_palette_kernel(const uint32_t* __restrict__ bitmap,
                                                                    const int* __restrict__ row_off,
                                                                    const unsigned char* __restrict__ payload,
                                                                    const int* __restrict__ labels_in,
                                                                    const int* __restrict__ worker_slot, int n,
                                                                    __half* __restrict__ X, int* __restrict__ labels,
                                                                    int cap, int F) {
    const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    const int words = F >> 5;
    for (int r = blockIdx.x * wpb + (threadIdx.x >> 5); r < n; r += gridDim.x * wpb) {
        const int worker = worker_slot[2 * r], slot = worker_slot[2 * r + 1];
        if (slot < 0) continue;
        __half* dst = X + ((size_t)worker * cap + slot) * F;
        const unsigned char* p = payload + row_off[r];
        const int npal = *(const unsigned short*)p;
        const unsigned short* vals = (const unsigned short*)(p + 2);           // palette, or the raw values
        const unsigned char* codes = p + 2 + 2 * npal;
        const unsigned int pal = lane < npal ? vals[lane] : 0u;
        int base = 0;
        for (int w0 = 0; w0 < words; w0 += 32) {
            const int w = w0 + lane;
            const uint32_t bits = w < words ? bitmap[(size_t)r * words + w] : 0u;
            const int cnt = __popc(bits);
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            int k = base + incl - cnt;
            __align__(16) unsigned short out[32];
            if (npal > 0) {                                                    // warp-uniform branch
                // This lane's <= 32 codes are nibbles [k, k + cnt) of the row's code stream.  Fetch the <= 20 bytes
                // that hold them with five aligned word loads (all issued at once) and line them up so that nibble
                // t of the lane sits at bit 4 (t & 7) of A[t >> 3]; a set bit b then finds its code with a popcount
                // instead of a dependent byte load per bit.
                const unsigned char* cb = codes + (k >> 1);
                const uintptr_t a0 = (uintptr_t)cb & ~(uintptr_t)3;
                const uint32_t* wp = (const uint32_t*)a0;
                const int nwords = cnt ? (int)(((uintptr_t)cb + (size_t)((cnt + (k & 1) + 1) >> 1) - a0 + 3) >> 2) : 0;
                uint32_t wd[6];
#pragma unroll
                for (int i = 0; i < 6; ++i) wd[i] = i < nwords ? __ldg(wp + i) : 0u;
                const uint32_t sh = (uint32_t)(((uintptr_t)cb & 3) * 8 + (k & 1) * 4);      // < 32
                uint32_t A[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) A[i] = __funnelshift_r(wd[i], wd[i + 1], sh);
#pragma unroll
                for (int b = 0; b < 32; ++b) {
                    const bool on = (bits >> b) & 1u;
                    const uint32_t t = (uint32_t)__popc(bits & ((1u << b) - 1u));          // codes before bit b
                    const uint32_t lo = (t & 8u) ? A[1] : A[0], hi = (t & 8u) ? A[3] : A[2];
                    const uint32_t word = (t & 16u) ? hi : lo;
                    const uint32_t code = on ? (word >> ((t & 7u) * 4u)) & 15u : 0u;
                    const unsigned int v = __shfl_sync(0xffffffffu, pal, (int)code);
                    out[b] = on ? (unsigned short)v : (unsigned short)0;
                }
            } else {
#pragma unroll
                for (int b = 0; b < 32; ++b) {
                    const bool on = (bits >> b) & 1u;
                    out[b] = on ? vals[k] : (unsigned short)0;
                    k += on;
                }
            }
            if (w < words) {
                uint4* d16 = (uint4*)(dst + (size_t)w * 32);
#pragma unroll
                for (int q = 0; q < 4; ++q) d16[q] = ((const uint4*)out)[q];
            }
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) labels[(size_t)worker * cap + slot] = labels_in[r];
    }
}

// Incremental window statistics: one block per (worker, 128-feature strip).  The rows that left / entered the window are
// dealt round-robin to the block's 8 warps; a lane owns 4 adjacent features (one 8-byte load per row, 256 contiguous
// bytes per warp) and keeps their partial sums in doubles; the warps' partials are combined through shared memory.
__global__ void __launch_bounds__(256) psb_window_stats_kernel(const __half* __restrict__ X,
                                                              const int4* __restrict__ change, int cap, int F,
                                                              double* __restrict__ sums, float* __restrict__ inv_std) {
    __shared__ double part[8][128][2];
    const int w = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int f0 = blockIdx.x * 128 + 4 * lane;                  // first of this lane's 4 features
    const bool live = f0 < F;                                    // (F % 4 == 0 is checked on the host)
    const int4 ch = change[w];
    const int old_head = ch.x, old_count = ch.y, added = ch.z, new_count = ch.w;
    const bool full = added < 0 || added > cap - old_count || new_count > old_count + added;
    // unwrapped coordinates: old window [oh - oc, oh), new window [nh - nc, nh) with nh = oh + added
    const int oh = old_head, nh = old_head + (added > 0 ? added : 0);
    int leave_lo = oh - old_count, leave_hi = min(oh, nh - new_count);       // rows that left
    int enter_lo = max(oh, nh - new_count), enter_hi = nh;                   // rows that entered
    if (full) leave_lo = leave_hi = 0, enter_lo = nh - new_count;
    double s1[4] = {0.0, 0.0, 0.0, 0.0}, s2[4] = {0.0, 0.0, 0.0, 0.0};
    const __half* base = X + (size_t)w * cap * F + f0;
    auto walk = [&](int lo, int hi, double sign) {
        for (int u0 = lo + warp; u0 < hi; u0 += 32) {           // 4 rows of this warp in flight
            uint2 raw[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int u = u0 + 8 * k;
                int slot = u % cap;
                if (slot < 0) slot += cap;
                raw[k] = (u < hi && live) ? *(const uint2*)(base + (size_t)slot * F) : make_uint2(0u, 0u);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 a = __half22float2(*(const __half2*)&raw[k].x), b = __half22float2(*(const __half2*)&raw[k].y);
                const double x[4] = {(double)a.x, (double)a.y, (double)b.x, (double)b.y};
#pragma unroll
                for (int e = 0; e < 4; ++e) s1[e] += sign * x[e], s2[e] += sign * x[e] * x[e];
            }
        }
    };
    walk(leave_lo, leave_hi, -1.0);
    walk(enter_lo, enter_hi, 1.0);
#pragma unroll
    for (int e = 0; e < 4; ++e) part[warp][4 * lane + e][0] = s1[e], part[warp][4 * lane + e][1] = s2[e];
    __syncthreads();
    if (threadIdx.x < 128) {
        const int f = blockIdx.x * 128 + threadIdx.x;
        if (f < F) {
            double2* sp = (double2*)sums + ((size_t)w * F + f);
            double2 acc = full ? make_double2(0.0, 0.0) : *sp;
#pragma unroll
            for (int g = 0; g < 8; ++g) acc.x += part[g][threadIdx.x][0], acc.y += part[g][threadIdx.x][1];
            *sp = acc;
            const double n = (double)new_count;
            const double var = new_count > 1 ? (acc.y - acc.x * acc.x / n) / (n - 1.0) : 0.0;
            inv_std[(size_t)w * F + f] = var > 1e-12 ? (float)(1.0 / sqrt(var)) : 0.f;
        }
    }
}

int psb_launch_window_stats(const void* d_X, const int* d_change, int num_workers, int cap, int F, double* d_sums,
                            float* d_inv_std, cudaStream_t stream) {
    if (num_workers <= 0) return 0;
    psb_window_stats_kernel<<<dim3((F + 127) / 128, num_workers), 256, 0, stream>>>(
        (const __half*)d_X, (const int4*)d_change, cap, F, d_sums, d_inv_std);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

__global__ void psb_apply_delta_kernel(float* __restrict__ w, const float* __restrict__ delta, float lr, int n4,
                                       int skip_index) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 a = ((float4*)w)[i];
    const float4 d = ((const float4*)delta)[i];
    float v[4] = {fmaf(lr, d.x, a.x), fmaf(lr, d.y, a.y), fmaf(lr, d.z, a.z), fmaf(lr, d.w, a.w)};
    if (skip_index >= 4 * i && skip_index < 4 * i + 4) v[skip_index - 4 * i] = (&a.x)[skip_index - 4 * i];
    ((float4*)w)[i] = make_float4(v[0], v[1], v[2], v[3]);
}

int psb_launch_ring_scatter(const uint32_t* d_stage, int n, int stride_words, void* d_X, int* d_labels, int cap, int F,
                            cudaStream_t stream) {
    if (n <= 0) return 0;
    psb_ring_scatter_kernel<<<n < 1184 ? n : 1184, 256, 0, stream>>>(d_stage, n, stride_words, (__half*)d_X, d_labels,
                                                                      cap, F);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
int psb_launch_ring_store_dense(const void* d_rows, int src_half, const int* d_labels_in, const int* d_ws, int n,
                                void* d_X, int* d_labels, int cap, int F, cudaStream_t stream) {
    if (n <= 0) return 0;
    const int grid = n < 2368 ? n : 2368;
    if (src_half)
        psb_ring_store_half_kernel<<<grid, 256, 0, stream>>>((const __half*)d_rows, d_labels_in, d_ws, n, (__half*)d_X,
                                                              d_labels, cap, F);
    else
        psb_ring_store_dense_kernel<<<grid, 256, 0, stream>>>((const float*)d_rows, d_labels_in, d_ws, n,
                                                               (__half*)d_X, d_labels, cap, F);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
int psb_launch_ring_store_bitmap(const uint32_t* d_bitmap, const int* d_row_off, const void* d_values,
                                 const int* d_labels_in, const int* d_ws, int n, void* d_X, int* d_labels, int cap, int F,
                                 cudaStream_t stream) {
    if (n <= 0) return 0;
    const int blocks = (n + 7) / 8;
    psb_ring_store_bitmap_kernel<<<blocks < 2368 ? blocks : 2368, 256, 0, stream>>>(
        d_bitmap, d_row_off, (const __half*)d_values, d_labels_in, d_ws, n, (__half*)d_X, d_labels, cap, F);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
int psb_launch_ring_store_palette(const uint32_t* d_bitmap, const int* d_row_off, const void* d_payload,
                                  const int* d_labels_in, const int* d_ws, int n, void* d_X, int* d_labels, int cap, int F,
                                  cudaStream_t stream) {
    if (n <= 0) return 0;
    const int blocks = (n + 7) / 8;
    psb_ring_store_palette_kernel<<<blocks < 2368 ? blocks : 2368, 256, 0, stream>>>(
        d_bitmap, d_row_off, (const unsigned char*)d_payload, d_labels_in, d_ws, n, (__half*)d_X, d_labels, cap, F);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
int psb_launch_apply_delta(float* d_w, const float* d_delta, float lr, int F, int skip_class, cudaStream_t stream) {
    const int n4 = (F + 1) * 2;
    const int skip = skip_class >= 0 ? F * 8 + skip_class : -1;
    psb_apply_delta_kernel<<<(n4 + 255) / 256, 256, 0, stream>>>(d_w, d_delta, lr, n4, skip);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
