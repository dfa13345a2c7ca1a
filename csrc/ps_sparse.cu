// Sharded sparse path (BASELINE config 5: 2^24 hashed features, server shards across the GPUs).
//
// The weight matrix W[F][8] is range-partitioned over the ranks (KeyRange, BaseMessage.java:24-27: the field
// the reference carries in every message but never uses, ServerApp.java:38); rank s owns rows
// [s*Fs, (s+1)*Fs) plus, on rank 0, the intercept row.  A worker step touches only the rows of its
// batch's non-zeros:
//   pull  : fused into the forward GEMV — every lane loads the 32-byte weight row of its non-zero straight
//           from the owning shard over NVLink (ld.relaxed.sys), no staging copy of the model;
//   push  : fused into the backward — the same lane fires two 16-byte red.global.add.v4.f32 of
//           -(eta / (B * numWorkers)) * v * r into the owner's row (the reduce AND the server apply,
//           ServerProcessor.java:148-151, in one instruction).
// Eventual / bounded-delay: one kernel does both.  Sequential: forward kernel, device-side phase barrier
// over peer memory, scatter kernel, phase barrier (all readers finish before any writer starts).
#include "common.cuh"

struct SparseComm {
    float* shard[PSB_MAX_RANKS];          // shard base pointers (mapped peers), (Fs + 1) rows x 8 floats
    unsigned int* phase[PSB_MAX_RANKS];   // phase[q][r]: rank r's phase counter as seen on rank q
    unsigned int* ticket;                 // local
    int rank, num_ranks, rows_per_shard;  // Fs
    unsigned long long spin_limit;
    unsigned int* status;                 // local error flag
};

__device__ __forceinline__ float* row_ptr(const SparseComm& c, int idx, int F) {
    if (idx >= F) return c.shard[0] + (size_t)c.rows_per_shard * 8;          // intercept row
    const int s = idx / c.rows_per_shard;
    return c.shard[s] + (size_t)(idx - s * c.rows_per_shard) * 8;
}

__device__ __forceinline__ bool phase_wait(const SparseComm& c, unsigned int want) {
    __shared__ int ok;
    if (threadIdx.x == 0) ok = 1;
    __syncthreads();
    if (threadIdx.x < c.num_ranks)
        if (!wait_ge_sys(c.phase[c.rank] + threadIdx.x, want, c.spin_limit, c.status, 3u)) ok = 0;
    __syncthreads();
    return ok != 0;
}
// last CTA of the grid publishes this rank's phase to every rank
__device__ __forceinline__ void phase_publish(const SparseComm& c, unsigned int value) {
    __shared__ unsigned int t;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) t = atom_add_acqrel_gpu(c.ticket, 1u);
    __syncthreads();
    if (t != gridDim.x - 1) return;
    if (threadIdx.x == 0) *c.ticket = 0u;
    fence_acq_rel_sys();
    if (threadIdx.x < c.num_ranks) st_release_sys(c.phase[threadIdx.x] + c.rank, value);
}

// mode bit 0: forward (logits -> softmax -> residuals R, loss); bit 1: scatter (push/apply from R or fused)
// wait_phase / publish_phase: 0 = none
__global__ void __launch_bounds__(256) psb_sparse_step_kernel(SparseComm c, int mode, unsigned int wait_phase,
                                                             unsigned int publish_phase, const long long* indptr,
                                                             const int* indices, const float* values,
                                                             const int* labels, int B, int F, int R, float scale,
                                                             float* Rbuf, float* loss_out) {
    if (wait_phase && !phase_wait(c, wait_phase)) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    float lsum = 0.f;
    for (int row = blockIdx.x * nwarp + warp; row < B; row += gridDim.x * nwarp) {
        const long long a = indptr[row], e = indptr[row + 1];
        float r[8];
        if (mode & 1) {
            float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (long long k = a + lane; k < e; k += 32) {
                const float v = values[k];
                const float4* w = (const float4*)row_ptr(c, indices[k], F);
                const float4 w0 = ld_relaxed_sys_v4(w), w1 = ld_relaxed_sys_v4(w + 1);       // the pull
                z[0] = fmaf(v, w0.x, z[0]), z[1] = fmaf(v, w0.y, z[1]), z[2] = fmaf(v, w0.z, z[2]);
                z[3] = fmaf(v, w0.w, z[3]), z[4] = fmaf(v, w1.x, z[4]), z[5] = fmaf(v, w1.y, z[5]);
                z[6] = fmaf(v, w1.z, z[6]), z[7] = fmaf(v, w1.w, z[7]);
            }
#pragma unroll
            for (int cc = 0; cc < 8; ++cc)
                for (int o = 16; o; o >>= 1) z[cc] += __shfl_xor_sync(0xffffffffu, z[cc], o);
            const float4* bw = (const float4*)row_ptr(c, F, F);
            const float4 b0 = ld_relaxed_sys_v4(bw), b1 = ld_relaxed_sys_v4(bw + 1);
            const float bias[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            float m = -INFINITY, p[8], sum = 0.f;
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                p[cc] = cc < R ? z[cc] + bias[cc] : -INFINITY;
                m = fmaxf(m, p[cc]);
            }
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                p[cc] = cc < R ? __expf(p[cc] - m) : 0.f;
                sum += p[cc];
            }
            const int y = labels[row];
            float ey = 1.f;
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                if (cc == y) ey = p[cc];
                r[cc] = p[cc] / sum - (cc == y ? 1.f : 0.f);
            }
            if (lane == 0) {
                lsum += __logf(sum) - __logf(fmaxf(ey, 1e-37f));
                if (Rbuf) {
                    ((float4*)Rbuf)[2 * row] = make_float4(r[0], r[1], r[2], r[3]);
                    ((float4*)Rbuf)[2 * row + 1] = make_float4(r[4], r[5], r[6], r[7]);
                }
            }
        } else {
            const float4 r0 = ((const float4*)Rbuf)[2 * row], r1 = ((const float4*)Rbuf)[2 * row + 1];
            r[0] = r0.x, r[1] = r0.y, r[2] = r0.z, r[3] = r0.w, r[4] = r1.x, r[5] = r1.y, r[6] = r1.z, r[7] = r1.w;
        }
        if (mode & 2) {                                                           // the push == the apply
            for (long long k = a + lane; k < e; k += 32) {
                const float sv = scale * values[k];
                float4* w = (float4*)row_ptr(c, indices[k], F);
                red_add_v4_sys(w, make_float4(sv * r[0], sv * r[1], sv * r[2], sv * r[3]));
                red_add_v4_sys(w + 1, make_float4(sv * r[4], sv * r[5], sv * r[6], sv * r[7]));
            }
            if (lane == 0) {
                float4* w = (float4*)row_ptr(c, F, F);
                red_add_v4_sys(w, make_float4(scale * r[0], scale * r[1], scale * r[2], scale * r[3]));
                red_add_v4_sys(w + 1, make_float4(scale * r[4], scale * r[5], scale * r[6], scale * r[7]));
            }
        }
    }
    if ((mode & 1) && lane == 0 && lsum != 0.f) atomicAdd(loss_out, lsum);
    if (publish_phase) phase_publish(c, publish_phase);
}

// argmax over the sharded weights -> confusion matrix (test-set evaluation of the sharded model)
__global__ void __launch_bounds__(256) psb_sparse_eval_kernel(SparseComm c, const long long* indptr, const int* indices,
                                                             const float* values, const int* labels, int B, int F,
                                                             int R, int* conf) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int row = blockIdx.x * nwarp + warp; row < B; row += gridDim.x * nwarp) {
        float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (long long k = indptr[row] + lane; k < indptr[row + 1]; k += 32) {
            const float v = values[k];
            const float4* w = (const float4*)row_ptr(c, indices[k], F);
            const float4 w0 = ld_relaxed_sys_v4(w), w1 = ld_relaxed_sys_v4(w + 1);
            z[0] = fmaf(v, w0.x, z[0]), z[1] = fmaf(v, w0.y, z[1]), z[2] = fmaf(v, w0.z, z[2]);
            z[3] = fmaf(v, w0.w, z[3]), z[4] = fmaf(v, w1.x, z[4]), z[5] = fmaf(v, w1.y, z[5]);
            z[6] = fmaf(v, w1.z, z[6]), z[7] = fmaf(v, w1.w, z[7]);
        }
#pragma unroll
        for (int cc = 0; cc < 8; ++cc)
            for (int o = 16; o; o >>= 1) z[cc] += __shfl_xor_sync(0xffffffffu, z[cc], o);
        if (lane == 0) {
            const float* bias = row_ptr(c, F, F);
            int best = 0;
            float bv = z[0] + bias[0];
            for (int cc = 1; cc < R; ++cc) {
                const float v = z[cc] + bias[cc];
                if (v > bv) bv = v, best = cc;
            }
            int y = labels[row];
            y = y < 0 ? 0 : (y > 7 ? 7 : y);
            atomicAdd(&conf[y * 8 + best], 1);
        }
    }
}

extern "C" {



static SparseComm to_dev(const PsbSparseComm* p) {
    SparseComm c;
    for (int i = 0; i < PSB_MAX_RANKS; ++i) c.shard[i] = (float*)p->shard[i], c.phase[i] = (unsigned int*)p->phase[i];
    c.ticket = (unsigned int*)p->ticket, c.status = (unsigned int*)p->status;
    c.rank = p->rank, c.num_ranks = p->num_ranks, c.rows_per_shard = p->rows_per_shard, c.spin_limit = p->spin_limit;
    return c;
}

int psb_sparse_step(const PsbSparseComm* comm, int mode, unsigned int wait_phase, unsigned int publish_phase,
                    const long long* d_indptr, const int* d_indices, const float* d_values, const int* d_labels,
                    int batch, int num_features, int num_class_rows, float scale, float* d_residuals, float* d_loss,
                    void* stream) {
    PSB_REQUIRE(comm && d_indptr && d_indices && d_values && d_labels && d_loss, "psb_sparse_step: null pointer");
    PSB_REQUIRE(comm->num_ranks >= 1 && comm->num_ranks <= PSB_MAX_RANKS, "psb_sparse_step: num_ranks out of range");
    PSB_REQUIRE((mode & 3) && (mode == 3 || d_residuals), "psb_sparse_step: split phases need a residual buffer");
    int blocks = (batch + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    psb_sparse_step_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(to_dev(comm), mode, wait_phase, publish_phase,
                                                                     d_indptr, d_indices, d_values, d_labels, batch,
                                                                     num_features, num_class_rows, scale, d_residuals,
                                                                     d_loss);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int psb_sparse_eval(const PsbSparseComm* comm, const long long* d_indptr, const int* d_indices, const float* d_values,
                    const int* d_labels, int n, int num_features, int num_class_rows, int* d_conf, void* stream) {
    PSB_REQUIRE(comm && d_indptr && d_indices && d_values && d_labels && d_conf, "psb_sparse_eval: null pointer");
    PSB_CHECK_CUDA(cudaMemsetAsync(d_conf, 0, 64 * sizeof(int), (cudaStream_t)stream));
    int blocks = (n + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    psb_sparse_eval_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(to_dev(comm), d_indptr, d_indices, d_values,
                                                                     d_labels, n, num_features, num_class_rows, d_conf);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
}

// =====================================================================================================
// Slab exchange: the same sharded step with every NVLink transfer made contiguous.
//
// Direct P2P above issues one 32-byte remote load and two 16-byte remote REDs per non-zero, which is bound by
// the link's request rate (measured 39 GB/s per direction at 8 GPUs), not its bandwidth.  Here a step is:
//   K1 bucket : every non-zero is assigned a position in a per-owner request list (warp-aggregated counters);
//               the list of local row ids goes to the owner (4-byte stores, consecutive positions);
//   K2 serve  : each owner gathers the requested rows from its shard (local random reads) and streams them, in
//               request order, into the requester's reply slab  (the PULL: contiguous remote stores);
//   K3 step   : forward GEMV from the local reply slabs, softmax, residuals; the per-non-zero updates
//               scale*v*r go, in the same request order, into the owner's update slab (the PUSH);
//   K5 apply  : each owner adds the update slabs into its shard with local REDs (duplicates are legal).
// Three device-side phase barriers per step order the stages across ranks; request lists are double
// buffered.  All pulls of a step complete before any apply starts, so the step is sequentially consistent.
struct SlabComm {
    SparseComm base;
    int* req[PSB_MAX_RANKS];             // req[s]: on rank s, [2][N][M] row ids requested from s
    unsigned int* req_cnt[PSB_MAX_RANKS];// on rank s, [2][N]
    float* rep[PSB_MAX_RANKS];           // rep[q]: on rank q, [N][M][8] rows received from each owner
    float* upd[PSB_MAX_RANKS];           // upd[s]: on rank s, [N][M][8] updates received from each requester
    unsigned int* cnt;                   // local [N] position counters
    int* pos;                            // local [nnz] position of every non-zero in its owner's list
    int cap;                             // M
    int parity;
};

// words 64.. of rank 0's req_cnt allocation: float biasbox[N][8], the ranks' summed intercept updates of a step
__device__ __forceinline__ float* slab_biasbox(const SlabComm& c, int q) { return (float*)(c.req_cnt[0] + 64) + q * 8; }

__device__ __forceinline__ void slab_tail_publish(const SlabComm& c, unsigned int value, bool send_counts,
                                                  bool send_bias = false) {
    __shared__ unsigned int t;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) t = atom_add_acqrel_gpu(c.base.ticket, 1u);
    __syncthreads();
    if (t != gridDim.x * gridDim.y - 1) return;
    if (threadIdx.x == 0) *c.base.ticket = 0u;
    if (send_counts && threadIdx.x < c.base.num_ranks) {
        const unsigned int n = atomicExch(c.cnt + threadIdx.x, 0u);           // read final count, reset for next step
        if (n > (unsigned)c.cap) atomicExch(c.base.status, 4u);
        st_relaxed_sys(c.req_cnt[threadIdx.x] + c.parity * c.base.num_ranks + c.base.rank, n);
    }
    if (send_bias && threadIdx.x < 8) {                // this rank's intercept update -> rank 0's bias box
        float* acc = (float*)(c.cnt + 8) + threadIdx.x;
        const float v = atomicExch(acc, 0.f);
        asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(slab_biasbox(c, c.base.rank) + threadIdx.x), "f"(v) : "memory");
    }
    fence_acq_rel_sys();
    __syncthreads();
    if (value && threadIdx.x < c.base.num_ranks) st_relaxed_sys(c.base.phase[threadIdx.x] + c.base.rank, value);
}

__global__ void __launch_bounds__(256) psb_slab_bucket_kernel(SlabComm c, const int* __restrict__ indices,
                                                             long long nnz, unsigned int publish) {
    const int lane = threadIdx.x & 31;
    const int N = c.base.num_ranks, Fs = c.base.rows_per_shard, me = c.base.rank;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k0 = (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31); k0 < nnz; k0 += stride) {
        const long long k = k0 + lane;
        const bool valid = k < nnz;
        const int idx = valid ? indices[k] : 0;
        const int s = idx / Fs;
        for (int o = 0; o < N; ++o) {
            const unsigned m = __ballot_sync(0xffffffffu, valid && s == o);
            if (!m) continue;
            const int leader = __ffs(m) - 1;
            unsigned int b = 0;
            if (lane == leader) b = atomicAdd(c.cnt + o, (unsigned)__popc(m));
            b = __shfl_sync(0xffffffffu, b, leader);
            if (valid && s == o) {
                const unsigned p = b + __popc(m & ((1u << lane) - 1u));
                c.pos[k] = (int)p;
                if (p < (unsigned)c.cap)
                    c.req[o][((size_t)c.parity * N + me) * c.cap + p] = idx - o * Fs;
            }
        }
    }
    slab_tail_publish(c, publish, true);
}

// blockIdx.y = requester q
__global__ void __launch_bounds__(256) psb_slab_serve_kernel(SlabComm c, unsigned int wait, unsigned int publish) {
    if (!phase_wait(c.base, wait)) return;
    const int N = c.base.num_ranks, me = c.base.rank, q = blockIdx.y;
    const unsigned n = min(ld_acquire_sys(c.req_cnt[me] + c.parity * N + q), (unsigned)c.cap);
    const int* req = c.req[me] + ((size_t)c.parity * N + q) * c.cap;
    const float4* W = (const float4*)c.base.shard[me];
    float4* out = (float4*)(c.rep[q] + (size_t)me * c.cap * 8);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int row = __ldcg(req + i);
        const float4 a = W[2 * (size_t)row], b = W[2 * (size_t)row + 1];
        st_cg_v4(out + 2 * (size_t)i, a);
        st_cg_v4(out + 2 * (size_t)i + 1, b);
    }
    slab_tail_publish(c, publish, false);
}

__global__ void __launch_bounds__(256) psb_slab_step_kernel(SlabComm c, unsigned int wait, unsigned int publish,
                                                           const long long* __restrict__ indptr,
                                                           const int* __restrict__ indices,
                                                           const float* __restrict__ values,
                                                           const int* __restrict__ labels, int B, int F, int R,
                                                           float scale, float* loss_out) {
    if (!phase_wait(c.base, wait)) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int Fs = c.base.rows_per_shard, me = c.base.rank;
    __shared__ float s_bias[8], s_bgrad[8];
    if (threadIdx.x < 8) {
        s_bias[threadIdx.x] = __ldcg(row_ptr(c.base, F, F) + threadIdx.x);
        s_bgrad[threadIdx.x] = 0.f;
    }
    __syncthreads();
    float lsum = 0.f, bg[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int row = blockIdx.x * nwarp + warp; row < B; row += gridDim.x * nwarp) {
        const long long a = indptr[row], e = indptr[row + 1];
        float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (long long k = a + lane; k < e; k += 32) {
            const float v = values[k];
            const int s = indices[k] / Fs;
            const float4* w = (const float4*)(c.rep[me] + ((size_t)s * c.cap + c.pos[k]) * 8);   // local reply slab
            const float4 w0 = ld_cg_v4(w), w1 = ld_cg_v4(w + 1);
            z[0] = fmaf(v, w0.x, z[0]), z[1] = fmaf(v, w0.y, z[1]), z[2] = fmaf(v, w0.z, z[2]);
            z[3] = fmaf(v, w0.w, z[3]), z[4] = fmaf(v, w1.x, z[4]), z[5] = fmaf(v, w1.y, z[5]);
            z[6] = fmaf(v, w1.z, z[6]), z[7] = fmaf(v, w1.w, z[7]);
        }
#pragma unroll
        for (int cc = 0; cc < 8; ++cc)
            for (int o = 16; o; o >>= 1) z[cc] += __shfl_xor_sync(0xffffffffu, z[cc], o);
        float m = -INFINITY, p[8], sum = 0.f, r[8];
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
            p[cc] = cc < R ? z[cc] + s_bias[cc] : -INFINITY;
            m = fmaxf(m, p[cc]);
        }
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
            p[cc] = cc < R ? __expf(p[cc] - m) : 0.f;
            sum += p[cc];
        }
        const int y = labels[row];
        float ey = 1.f;
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
            if (cc == y) ey = p[cc];
            r[cc] = p[cc] / sum - (cc == y ? 1.f : 0.f);
        }
        if (lane == 0) {
            lsum += __logf(sum) - __logf(fmaxf(ey, 1e-37f));
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) bg[cc] += r[cc];
        }
        for (long long k = a + lane; k < e; k += 32) {                   // the push: update slab of the owner
            const float sv = scale * values[k];
            const int s = indices[k] / Fs;
            float4* u = (float4*)(c.upd[s] + ((size_t)me * c.cap + c.pos[k]) * 8);
            st_cg_v4(u, make_float4(sv * r[0], sv * r[1], sv * r[2], sv * r[3]));
            st_cg_v4(u + 1, make_float4(sv * r[4], sv * r[5], sv * r[6], sv * r[7]));
        }
    }
    if (lane == 0) {
        if (lsum != 0.f) atomicAdd(loss_out, lsum);
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) atomicAdd(&s_bgrad[cc], bg[cc]);
    }
    __syncthreads();
    if (threadIdx.x < 8 && s_bgrad[threadIdx.x] != 0.f)                  // intercept update: reduced per rank, applied
        atomicAdd((float*)(c.cnt + 8) + threadIdx.x, scale * s_bgrad[threadIdx.x]);   // by rank 0 in the apply stage
    slab_tail_publish(c, publish, false, true);
}

// blockIdx.y = requester q
__global__ void __launch_bounds__(256) psb_slab_apply_kernel(SlabComm c, unsigned int wait) {
    if (!phase_wait(c.base, wait)) return;
    const int N = c.base.num_ranks, me = c.base.rank, q = blockIdx.y;
    const unsigned n = min(ld_acquire_sys(c.req_cnt[me] + c.parity * N + q), (unsigned)c.cap);
    const int* req = c.req[me] + ((size_t)c.parity * N + q) * c.cap;
    const float4* in = (const float4*)(c.upd[me] + (size_t)q * c.cap * 8);
    float4* W = (float4*)c.base.shard[me];
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int row = __ldcg(req + i);
        red_add_v4(W + 2 * (size_t)row, ld_cg_v4(in + 2 * (size_t)i));
        red_add_v4(W + 2 * (size_t)row + 1, ld_cg_v4(in + 2 * (size_t)i + 1));
    }
    if (me == 0 && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < 8) {      // intercept row: sum of the bias boxes
        float sum = 0.f;
        for (int r = 0; r < N; ++r) sum += __ldcg(slab_biasbox(c, r) + threadIdx.x);
        float* b = c.base.shard[0] + (size_t)c.base.rows_per_shard * 8 + threadIdx.x;
        *b = *b + sum;
    }
}

extern "C" {



// One sharded SGD step through the slab exchange.  `step_index` numbers the device-side phases.
int psb_sparse_step_slab(const PsbSlabComm* comm, int step_index, const long long* d_indptr, const int* d_indices,
                         const float* d_values, const int* d_labels, int batch, long long nnz, int num_features,
                         int num_class_rows, float scale, float* d_loss, int stage_mask, void* stream) {
    PSB_REQUIRE(comm && d_indptr && d_indices && d_values && d_labels && d_loss, "psb_sparse_step_slab: null pointer");
    const int N = comm->base.num_ranks;
    PSB_REQUIRE(N >= 1 && N <= PSB_MAX_RANKS, "psb_sparse_step_slab: num_ranks out of range");
    SlabComm c;
    c.base = to_dev(&comm->base);
    for (int i = 0; i < PSB_MAX_RANKS; ++i) {
        c.req[i] = (int*)comm->req[i], c.req_cnt[i] = (unsigned int*)comm->req_cnt[i];
        c.rep[i] = (float*)comm->rep[i], c.upd[i] = (float*)comm->upd[i];
    }
    c.cnt = (unsigned int*)comm->cnt, c.pos = (int*)comm->pos, c.cap = comm->cap, c.parity = step_index & 1;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned ph = 3u * (unsigned)step_index;
    const bool multi = N > 1;
    int blocks = (int)((nnz + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    if (stage_mask & 1) psb_slab_bucket_kernel<<<blocks, 256, 0, st>>>(c, d_indices, nnz, ph + 1);
    if (stage_mask & 2) psb_slab_serve_kernel<<<dim3(148, N), 256, 0, st>>>(c, multi ? ph + 1 : 0, ph + 2);
    int rb = (batch + 7) / 8;
    if (rb > 148 * 8) rb = 148 * 8;
    if (rb < 1) rb = 1;
    if (stage_mask & 4)
        psb_slab_step_kernel<<<rb, 256, 0, st>>>(c, multi ? ph + 2 : 0, ph + 3, d_indptr, d_indices, d_values, d_labels,
                                                 batch, num_features, num_class_rows, scale, d_loss);
    if (stage_mask & 8) psb_slab_apply_kernel<<<dim3(148, N), 256, 0, st>>>(c, multi ? ph + 3 : 0);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
}
