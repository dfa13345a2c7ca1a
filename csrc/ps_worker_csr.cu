// K2, sparse-row flavour: the worker's local solve (the same in-kernel L-BFGS / SGD of solver.cuh) on buffers that
// hold SPARSE rows, for models the dense tcgen05 kernel does not take (num_features > 1024, or not a multiple of
// 128) and for rows too sparse to be worth densifying.  The reference fits on sparse vectors of any width
// (LogisticRegressionTaskSpark.java:146-184: buffer -> Vectors.sparse -> LogisticRegression.fit).
//
// Input rings (ELL): per worker `cap` slots of at most K non-zeros: idx[w][cap][K] int32, val[w][cap][K] fp16,
// nnz[w][cap] int32, labels[w][cap]; the (head, count) window as for the dense rings.
//
// One CTA per worker.  An evaluation walks the window once, one warp per row:
//   forward : z = sum_k val_k * Weff[idx_k]   (Weff = x * 1/std, staged in shared memory when (F+1)*32 bytes fit,
//             read from L2 otherwise), warp reduction, softmax, residual r, loss;
//   backward: the same warp adds val_k * r into the gradient row of idx_k with two 16-byte red.global.add.v4.f32
//             (the worker's gradient vector lives in L2-resident scratch; shared-memory float atomics would be
//             CAS loops on this architecture);
// then one pass scales the raw sums by 1/(n std_f).  The per-feature statistics (Spark standardisation) are two
// more RED targets during a dedicated first walk.  pull / push / apply over peer memory are comm.cuh, exactly as
// in the dense kernels, so all consistency models and the per-worker clocks work unchanged.
#include "comm.cuh"
#include "solver.cuh"
#include "worker_common.cuh"

namespace {
struct CsrEval {
    const int* idx;            // this worker's ring [cap][K]
    const __half* val;
    const int* nnz;            // [cap]
    const int* labels;         // [cap]
    int start, count, cap, F, R, K;
    float* Ws;                 // effective weights [(F+1)*8]: shared memory, or global scratch when they do not fit
    bool ws_shared;
    const float* inv;          // global [F]: 1/std (or 1)
    float* sred;               // smem [16]: loss + per-class residual sums
    double* red;

    __device__ __forceinline__ int slot_of(int i) const {
        int s = start + i;
        return s >= cap ? s - cap : s;
    }

    // loss and gradient (scaled space) at x; all threads return the same loss
    __device__ double eval(const float* x, float* g) {
        const int tid = threadIdx.x, nthr = blockDim.x;
        const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
        const int n4 = (F + 1) * 2;
        // 1. effective weights w = x * inv (intercepts unscaled); gradient accumulator cleared
        for (int i = tid; i < n4; i += nthr) {
            const int f = i >> 1;
            const float4 v = ((const float4*)x)[i];
            const float sc = f < F ? inv[f] : 1.f;
            ((float4*)Ws)[i] = make_float4(v.x * sc, v.y * sc, v.z * sc, v.w * sc);
            st_cg_v4((float4*)g + i, make_float4(0.f, 0.f, 0.f, 0.f));
        }
        if (tid < 16) sred[tid] = 0.f;
        __threadfence();
        __syncthreads();
        // 2. one walk over the window: forward and backward of a row by the same warp
        float lsum = 0.f, rs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float4* W4 = (const float4*)Ws;
        for (int i = warp; i < count; i += nwarp) {
            const int slot = slot_of(i);
            const int* ri = idx + (size_t)slot * K;
            const __half* rv = val + (size_t)slot * K;
            const int n = min(nnz[slot], K);
            float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int k = lane; k < n; k += 32) {
                const float v = __half2float(rv[k]);
                const int f = ri[k];
                float4 a0, a1;
                if (ws_shared) a0 = W4[2 * f], a1 = W4[2 * f + 1];
                else a0 = ld_cg_v4(W4 + 2 * f), a1 = ld_cg_v4(W4 + 2 * f + 1);
                z[0] = fmaf(v, a0.x, z[0]), z[1] = fmaf(v, a0.y, z[1]), z[2] = fmaf(v, a0.z, z[2]);
                z[3] = fmaf(v, a0.w, z[3]), z[4] = fmaf(v, a1.x, z[4]), z[5] = fmaf(v, a1.y, z[5]);
                z[6] = fmaf(v, a1.z, z[6]), z[7] = fmaf(v, a1.w, z[7]);
            }
#pragma unroll
            for (int c = 0; c < 8; ++c)
                for (int o = 16; o; o >>= 1) z[c] += __shfl_xor_sync(0xffffffffu, z[c], o);
            float bias[8], r[8];
            if (ws_shared) {
#pragma unroll
                for (int c = 0; c < 8; ++c) bias[c] = Ws[F * 8 + c];
            } else {
                const float4 b0 = ld_cg_v4(W4 + 2 * F), b1 = ld_cg_v4(W4 + 2 * F + 1);
                bias[0] = b0.x, bias[1] = b0.y, bias[2] = b0.z, bias[3] = b0.w;
                bias[4] = b1.x, bias[5] = b1.y, bias[6] = b1.z, bias[7] = b1.w;
            }
            const float li = softmax_residual(z, bias, labels[slot], R, r);
            if (lane == 0) {
                lsum += li;
#pragma unroll
                for (int c = 0; c < 8; ++c) rs[c] += r[c];
            }
            for (int k = lane; k < n; k += 32) {                      // backward: g[idx] += val * r
                const float v = __half2float(rv[k]);
                float4* gr = (float4*)g + 2 * ri[k];
                red_add_v4(gr, make_float4(v * r[0], v * r[1], v * r[2], v * r[3]));
                red_add_v4(gr + 1, make_float4(v * r[4], v * r[5], v * r[6], v * r[7]));
            }
        }
        if (lane == 0) {
            atomicAdd(&sred[0], lsum);
#pragma unroll
            for (int c = 0; c < 8; ++c) atomicAdd(&sred[1 + c], rs[c]);
        }
        __threadfence();
        __syncthreads();
        // 3. raw sums -> scaled-space gradient
        const float invn = 1.f / (float)count;
        for (int i = tid; i < F * 2; i += nthr) {
            const float sc = inv[i >> 1] * invn;
            const float4 v = ld_cg_v4((const float4*)g + i);
            ((float4*)g)[i] = make_float4(v.x * sc, v.y * sc, v.z * sc, v.w * sc);
        }
        if (tid < 8) g[F * 8 + tid] = sred[1 + tid] * invn;
        const double loss = (double)sred[0] * (double)invn;
        __syncthreads();
        return loss;
    }
};

// inv[f] = 1/std_f over the window (unbiased, zeros included; 0 for constant columns) — column sums by REDs into
// acc[F][2] (a spare scratch vector), then converted in place into inv[F].
__device__ void csr_inv_std(const int* idx, const __half* val, const int* nnz, int start, int count, int cap, int F,
                            int K, int standardize, float* acc, float* inv) {
    const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    if (!standardize) {
        for (int f = tid; f < F; f += nthr) inv[f] = 1.f;
        return;
    }
    for (int i = tid; i < F / 2; i += nthr) st_cg_v4((float4*)acc + i, make_float4(0.f, 0.f, 0.f, 0.f));
    if ((F & 1) && tid == 0) acc[2 * (F - 1)] = 0.f, acc[2 * (F - 1) + 1] = 0.f;
    __threadfence();
    __syncthreads();
    for (int i = warp; i < count; i += nwarp) {
        int slot = start + i;
        if (slot >= cap) slot -= cap;
        const int n = min(nnz[slot], K);
        for (int k = lane; k < n; k += 32) {
            const float v = __half2float(val[(size_t)slot * K + k]);
            float* a = acc + 2 * (size_t)idx[(size_t)slot * K + k];
            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(a), "f"(v), "f"(v * v) : "memory");
        }
    }
    __threadfence();
    __syncthreads();
    const float n = (float)count;
    for (int f = tid; f < F; f += nthr) {
        float s1, s2;
        asm volatile("ld.global.cg.v2.f32 {%0, %1}, [%2];" : "=f"(s1), "=f"(s2) : "l"(acc + 2 * (size_t)f));
        const float var = count > 1 ? (s2 - s1 * s1 / n) / (n - 1.f) : 0.f;
        inv[f] = var > 1e-12f ? rsqrtf(var) : 0.f;
    }
}

// One clock of one worker; false when an in-kernel wait gave up (the CTA stops).
__device__ bool csr_clock(const PsbSolver& sv, const CommCtx& cx, int clock, int it, int K, const int* __restrict__ idx,
                          const __half* __restrict__ val, const int* __restrict__ nnz, const int* __restrict__ labels,
                          const int* __restrict__ win, const float* __restrict__ w_in, float* __restrict__ delta_out,
                          float* __restrict__ loss_out, int* __restrict__ info_out, float* __restrict__ scratch,
                          unsigned char* smem, bool ws_shared) {
    const int F = sv.num_features, R = sv.num_class_rows, cap = sv.cap;
    const int nv = (F + 1) * 8, n4 = nv / 4;
    const int w = blockIdx.x, gw = cx.c.worker_base + w;
    float* sred = (float*)smem;
    double* red = (double*)(sred + 16);
    float* Ws_s = (float*)(red + 34);
    PsbClockRecord* rec = cx.c.mode == PSB_COMM_WORKER && cx.c.hist
                              ? (PsbClockRecord*)cx.c.hist + (size_t)it * gridDim.x + w : nullptr;

    WorkerScratch ws(scratch + (size_t)w * PSB_SCRATCH_VECS * nv, nv);
    const int head = win[2 * w], count = win[2 * w + 1];
    int start = head - count;
    if (start < 0) start += cap;
    const int* iw = idx + (size_t)w * cap * K;
    const __half* vw = val + (size_t)w * cap * K;
    const int* nw = nnz + (size_t)w * cap;
    const int* lw = labels + (size_t)w * cap;

    const long long t_begin = clock64();
    float* w0 = ws.w0;
    if (cx.c.mode == PSB_COMM_NONE) {
        for (int i = threadIdx.x; i < n4; i += blockDim.x) ((float4*)w0)[i] = ((const float4*)w_in)[i];
        __syncthreads();
    } else if (!comm_pull(cx, clock, gw, (float4*)w0, rec)) {
        return false;
    }
    if (count <= 0) {
        if (threadIdx.x == 0) {
            loss_out[w] = nanf("");
            info_out[8 * w + 2] = -1, info_out[8 * w + 3] = 0;
            if (rec) rec->loss = nanf("");
        }
        if (cx.c.mode != PSB_COMM_NONE) comm_push(cx, clock, gw, nullptr);
        return true;
    }
    // the solver is told to keep PSB_HIST-1 pairs, which leaves the last pair's two vectors to this kernel: 1/std,
    // and the effective weights when they do not fit shared memory
    float* inv = ws.vec[7 + PSB_HIST - 1];
    float* Ws = ws_shared ? Ws_s : ws.vec[7 + 2 * PSB_HIST - 1];
    SolveVecs v;
    ws.bind(v);
    v.n4 = n4;
    csr_inv_std(iw, vw, nw, start, count, cap, F, K, sv.standardize, v.xt, inv);
    __syncthreads();
    const long long t_stats = clock64();
    to_scaled_space(v.x, w0, inv, F);
    __syncthreads();

    CsrEval ev = {iw, vw, nw, lw, start, count, cap, F, R, K, Ws, ws_shared, inv, sred, red};
    SolveResult res = local_solve(ev, v, sv.solver, sv.max_iter, sv.sgd_lr, red, false, 0.0, PSB_HIST - 1);
    const long long t_solve = clock64();

    float* delta = v.tmp;
    finish_delta(delta, v.x, w0, inv, F, R, sv.center && sv.solver == PSB_SOLVER_LBFGS);
    __syncthreads();
    if (threadIdx.x == 0) {
        loss_out[w] = (float)res.loss;
        if (rec) rec->loss = (float)res.loss;
        info_out[8 * w + 0] = res.evals, info_out[8 * w + 1] = res.iterations;
        info_out[8 * w + 2] = res.status, info_out[8 * w + 3] = count;
        info_out[8 * w + 4] = (int)((t_stats - t_begin) >> 6), info_out[8 * w + 5] = (int)((t_solve - t_stats) >> 6);
    }
    if (cx.c.mode == PSB_COMM_NONE) {
        float4* out = (float4*)(delta_out + (size_t)w * nv);
        for (int i = threadIdx.x; i < n4; i += blockDim.x) out[i] = ((const float4*)delta)[i];
    } else {
        comm_push(cx, clock, gw, (const float4*)delta);
    }
    if (threadIdx.x == 0) info_out[8 * w + 6] = (int)((clock64() - t_solve) >> 6), info_out[8 * w + 7] = (int)((clock64() - t_begin) >> 6);
    return true;
}
}  // namespace

__global__ void __launch_bounds__(512, 1)
psb_worker_csr_kernel(PsbSolver sv, CommCtx cx, int K, int ws_shared, const int* __restrict__ idx,
                      const __half* __restrict__ val, const int* __restrict__ nnz, const int* __restrict__ labels,
                      const int* __restrict__ win, const float* __restrict__ w_in, float* __restrict__ delta_out,
                      float* __restrict__ loss_out, int* __restrict__ info_out, float* __restrict__ scratch) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int nclk = cx.c.mode == PSB_COMM_WORKER && cx.c.num_clocks > 1 ? cx.c.num_clocks : 1;
    for (int it = 0; it < nclk; ++it) {
        if (!csr_clock(sv, cx, cx.c.clock + it, it, K, idx, val, nnz, labels, win, w_in, delta_out, loss_out, info_out,
                       scratch, smem, ws_shared != 0))
            return;
        __syncthreads();
    }
}

// staged rows (psb_ring_scatter's record format: word0 nnz, word1 label, word2 worker, word3 slot, then nnz x
// (int32 index, fp32 value)) -> ELL ring slots; rows with more than K non-zeros keep their first K
__global__ void __launch_bounds__(128) psb_ring_store_ell_kernel(const uint32_t* __restrict__ stage, int n,
                                                                int stride_words, int* __restrict__ idx,
                                                                __half* __restrict__ val, int* __restrict__ nnz,
                                                                int* __restrict__ labels, int cap, int K, int F) {
    for (int r = blockIdx.x; r < n; r += gridDim.x) {
        const uint32_t* row = stage + (size_t)r * stride_words;
        const int m = min((int)row[0], K), worker = (int)row[2], slot = (int)row[3];
        if (slot < 0) continue;
        const size_t base = ((size_t)worker * cap + slot) * K;
        for (int e = threadIdx.x; e < m; e += blockDim.x) {
            const int f = (int)row[4 + 2 * e];
            const bool ok = f >= 0 && f < F;
            idx[base + e] = ok ? f : 0;
            val[base + e] = __float2half_rn(ok ? __uint_as_float(row[5 + 2 * e]) : 0.f);
        }
        if (threadIdx.x == 0) nnz[(size_t)worker * cap + slot] = m, labels[(size_t)worker * cap + slot] = (int)row[1];
    }
}

extern "C" {

int psb_worker_step_csr(const PsbSolver* s, const PsbComm* comm, int num_workers, int max_nnz, const int* d_idx,
                        const void* d_val, const int* d_nnz, const int* d_labels, const int* d_win,
                        const float* d_w_in, float* d_delta, float* d_loss, int* d_info, float* d_scratch,
                        void* stream) {
    PSB_REQUIRE(s && comm, "psb_worker_step_csr: null solver/comm");
    PSB_REQUIRE(num_workers >= 1 && max_nnz >= 1, "psb_worker_step_csr: bad num_workers / max_nnz");
    PSB_REQUIRE(s->num_class_rows >= 2 && s->num_class_rows <= PSB_CLASS_PAD, "num_class_rows must be in [2,8]");
    PSB_REQUIRE(s->num_features >= 2 && s->cap >= 1, "bad num_features/cap");
    PSB_REQUIRE(d_idx && d_val && d_nnz && d_labels && d_win && d_loss && d_info && d_scratch,
                "psb_worker_step_csr: null buffer");
    PSB_REQUIRE(s->max_iter >= 0 && s->max_iter <= 64, "max_iter out of range");
    CommCtx cx;
    cx.c = *comm;
    if (comm->mode == PSB_COMM_NONE) {
        PSB_REQUIRE(d_w_in && d_delta, "comm mode NONE needs d_w_in and d_delta");
        cx.L = CommLayout::make(s->num_features, 1);
    } else {
        PSB_REQUIRE(comm->num_ranks >= 1 && comm->num_ranks <= PSB_MAX_RANKS, "num_ranks out of range");
        PSB_REQUIRE(comm->rank >= 0 && comm->rank < comm->num_ranks, "rank out of range");
        for (int q = 0; q < comm->num_ranks; ++q) PSB_REQUIRE(comm->peers[q], "peer block %d is null", q);
        cx.L = CommLayout::make(s->num_features, comm->num_ranks);
        if (comm->mode == PSB_COMM_WORKER) {
            PSB_REQUIRE(comm->total_workers >= 1 && comm->total_workers <= PSB_MAX_WORKERS, "total_workers out of range");
            PSB_REQUIRE(comm->worker_base >= 0 && comm->worker_base + num_workers <= comm->total_workers,
                        "workers outside the clock table");
        } else {
            PSB_REQUIRE(comm->workers_per_rank == num_workers, "workers_per_rank must equal num_workers");
            PSB_REQUIRE(comm->num_clocks <= 1, "several clocks per launch need comm mode PSB_COMM_WORKER");
        }
    }
    const size_t small = 16 * 4 + 34 * 8;
    const size_t wbytes = (size_t)(s->num_features + 1) * PSB_CLASS_PAD * 4;
    const int ws_shared = small + wbytes <= 200 * 1024;
    const size_t smem = small + (ws_shared ? wbytes : 0) + 16;
    PSB_CHECK_CUDA(cudaFuncSetAttribute(psb_worker_csr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (comm->mode == PSB_COMM_WORKER && comm->max_delay >= 0) {
        int dev = 0, sms = 0, per_sm = 0;
        PSB_CHECK_CUDA(cudaGetDevice(&dev));
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, psb_worker_csr_kernel, 512, smem);
        if (num_workers > sms * per_sm) {
            psb_set_error("bounded-delay launch of %d workers, but only %d sparse-row worker CTAs are resident at once",
                          num_workers, sms * per_sm);
            return PSB_E_LIMIT;
        }
    }
    psb_worker_csr_kernel<<<num_workers, 512, smem, (cudaStream_t)stream>>>(
        *s, cx, max_nnz, ws_shared, d_idx, (const __half*)d_val, d_nnz, d_labels, d_win, d_w_in, d_delta, d_loss, d_info,
        d_scratch);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int psb_ring_store_ell(const uint32_t* d_stage, int n, int stride_words, int* d_idx, void* d_val, int* d_nnz,
                       int* d_labels, int cap, int max_nnz, int num_features, void* stream) {
    PSB_REQUIRE(d_stage && d_idx && d_val && d_nnz && d_labels, "psb_ring_store_ell: null pointer");
    if (n <= 0) return 0;
    psb_ring_store_ell_kernel<<<n < 2368 ? n : 2368, 128, 0, (cudaStream_t)stream>>>(
        d_stage, n, stride_words, d_idx, (__half*)d_val, d_nnz, d_labels, cap, max_nnz, num_features);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
}
