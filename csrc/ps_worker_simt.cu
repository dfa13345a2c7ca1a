// K2 (SIMT flavour): one parameter-server iteration per worker CTA — pull, local solve on the dense
// fp16 ring buffer with fp32 CUDA-core math, push/apply.  This is the numerics reference of the
// tcgen05 kernel (ps_worker_tc.cu) and the fallback for shapes the tensor path does not take.
//
// Replaces LogisticRegressionTaskSpark.calculateGradients (LogisticRegressionTaskSpark.java:142-221):
// buffer -> (standardise) -> L-BFGS from the pulled weights -> (centre) -> delta = new - old.
#include "comm.cuh"
#include "solver.cuh"
#include "worker_common.cuh"

// ------------------------------------------------------------------ evaluator
struct SimtEval {
    const __half* X;        // this worker's ring [cap][F]
    const int* labels;      // [cap]
    int start, count, cap, F, R;
    float* Ws;              // smem [(F+1)*8]  effective (unscaled-space) weights
    float* Rs;              // smem [cap*8]    residuals, indexed by window position
    const float* inv;       // smem [F]        1/std (or 1)
    float* sred;            // smem [16]       loss + per-class residual sums
    double* red;

    __device__ __forceinline__ int slot_of(int i) const {
        int s = start + i;
        return s >= cap ? s - cap : s;
    }

    // loss and gradient (scaled space) at x; all threads return the same loss
    __device__ double eval(const float* x, float* g) {
        const int tid = threadIdx.x, nthr = blockDim.x;
        const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
        // 1. stage effective weights: w = x * inv (intercepts unscaled)
        for (int i = tid; i < (F + 1) * 2; i += nthr) {
            const int f = i >> 1;
            float4 v = ((const float4*)x)[i];
            const float sc = f < F ? inv[f] : 1.f;
            ((float4*)Ws)[i] = make_float4(v.x * sc, v.y * sc, v.z * sc, v.w * sc);
        }
        if (tid < 16) sred[tid] = 0.f;
        __syncthreads();
        // 2. forward: warp per row
        float lsum = 0.f, rs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float4* W4 = (const float4*)Ws;
        for (int i = warp; i < count; i += nwarp) {
            const int slot = slot_of(i);
            const __half2* xr = (const __half2*)(X + (size_t)slot * F);
            float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int f2 = lane; f2 < (F >> 1); f2 += 32) {
                const float2 xv = __half22float2(xr[f2]);
                const float4 a0 = W4[4 * f2], a1 = W4[4 * f2 + 1], b0 = W4[4 * f2 + 2], b1 = W4[4 * f2 + 3];
                z[0] = fmaf(xv.x, a0.x, z[0]), z[1] = fmaf(xv.x, a0.y, z[1]), z[2] = fmaf(xv.x, a0.z, z[2]);
                z[3] = fmaf(xv.x, a0.w, z[3]), z[4] = fmaf(xv.x, a1.x, z[4]), z[5] = fmaf(xv.x, a1.y, z[5]);
                z[6] = fmaf(xv.x, a1.z, z[6]), z[7] = fmaf(xv.x, a1.w, z[7]);
                z[0] = fmaf(xv.y, b0.x, z[0]), z[1] = fmaf(xv.y, b0.y, z[1]), z[2] = fmaf(xv.y, b0.z, z[2]);
                z[3] = fmaf(xv.y, b0.w, z[3]), z[4] = fmaf(xv.y, b1.x, z[4]), z[5] = fmaf(xv.y, b1.y, z[5]);
                z[6] = fmaf(xv.y, b1.z, z[6]), z[7] = fmaf(xv.y, b1.w, z[7]);
            }
#pragma unroll
            for (int c = 0; c < 8; ++c)
                for (int o = 16; o; o >>= 1) z[c] += __shfl_xor_sync(0xffffffffu, z[c], o);
            float r[8];
            const float li = softmax_residual(z, Ws + F * 8, labels[slot], R, r);
            if (lane == 0) {
                lsum += li;
#pragma unroll
                for (int c = 0; c < 8; ++c) rs[c] += r[c];
                ((float4*)Rs)[2 * i] = make_float4(r[0], r[1], r[2], r[3]);
                ((float4*)Rs)[2 * i + 1] = make_float4(r[4], r[5], r[6], r[7]);
            }
        }
        if (lane == 0) {
            atomicAdd(&sred[0], lsum);
#pragma unroll
            for (int c = 0; c < 8; ++c) atomicAdd(&sred[1 + c], rs[c]);
        }
        __syncthreads();
        // 3. backward: thread per feature pair, registers accumulate over rows (no atomics)
        const float invn = 1.f / (float)count;
        for (int fp = tid; fp < (F >> 1); fp += nthr) {
            float a[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) a[k] = 0.f;
            const __half2* xc = (const __half2*)X + fp;
            const size_t rowstride = (size_t)(F >> 1);
#pragma unroll 4
            for (int i = 0; i < count; ++i) {
                const float2 xv = __half22float2(xc[(size_t)slot_of(i) * rowstride]);
                const float4 r0 = ((const float4*)Rs)[2 * i], r1 = ((const float4*)Rs)[2 * i + 1];
                a[0] = fmaf(xv.x, r0.x, a[0]), a[1] = fmaf(xv.x, r0.y, a[1]), a[2] = fmaf(xv.x, r0.z, a[2]);
                a[3] = fmaf(xv.x, r0.w, a[3]), a[4] = fmaf(xv.x, r1.x, a[4]), a[5] = fmaf(xv.x, r1.y, a[5]);
                a[6] = fmaf(xv.x, r1.z, a[6]), a[7] = fmaf(xv.x, r1.w, a[7]);
                a[8] = fmaf(xv.y, r0.x, a[8]), a[9] = fmaf(xv.y, r0.y, a[9]), a[10] = fmaf(xv.y, r0.z, a[10]);
                a[11] = fmaf(xv.y, r0.w, a[11]), a[12] = fmaf(xv.y, r1.x, a[12]), a[13] = fmaf(xv.y, r1.y, a[13]);
                a[14] = fmaf(xv.y, r1.z, a[14]), a[15] = fmaf(xv.y, r1.w, a[15]);
            }
            const float s0 = inv[2 * fp] * invn, s1 = inv[2 * fp + 1] * invn;
            float4* g4 = (float4*)g + 4 * fp;
            g4[0] = make_float4(a[0] * s0, a[1] * s0, a[2] * s0, a[3] * s0);
            g4[1] = make_float4(a[4] * s0, a[5] * s0, a[6] * s0, a[7] * s0);
            g4[2] = make_float4(a[8] * s1, a[9] * s1, a[10] * s1, a[11] * s1);
            g4[3] = make_float4(a[12] * s1, a[13] * s1, a[14] * s1, a[15] * s1);
        }
        if (tid < 8) g[F * 8 + tid] = sred[1 + tid] * invn;
        const double loss = (double)sred[0] * (double)invn;
        __syncthreads();
        return loss;
    }
};

// ------------------------------------------------------------------ kernel
// One clock of one worker; false when an in-kernel wait gave up (the CTA stops).
__device__ bool simt_clock(const PsbSolver& sv, const CommCtx& cx, int clock, int it, const __half* __restrict__ X,
                           const int* __restrict__ labels, const int* __restrict__ win, const float* __restrict__ w_in,
                           float* __restrict__ delta_out, float* __restrict__ loss_out, int* __restrict__ info_out,
                           float* __restrict__ scratch, unsigned char* smem) {
    const int F = sv.num_features, R = sv.num_class_rows, cap = sv.cap;
    const int nv = (F + 1) * 8, n4 = nv / 4;
    const int w = blockIdx.x, gw = cx.c.worker_base + w;
    float* Ws = (float*)smem;
    float* Rs = Ws + nv;
    float* inv = Rs + (size_t)cap * 8;
    float* sred = inv + F;
    double* red = (double*)(sred + 16);
    PsbClockRecord* rec = cx.c.mode == PSB_COMM_WORKER && cx.c.hist
                              ? (PsbClockRecord*)cx.c.hist + (size_t)it * gridDim.x + w : nullptr;

    WorkerScratch ws(scratch + (size_t)w * PSB_SCRATCH_VECS * nv, nv);
    const int head = win[2 * w], count = win[2 * w + 1];
    int start = head - count;
    if (start < 0) start += cap;
    const __half* Xw = X + (size_t)w * cap * F;
    const int* lw = labels + (size_t)w * cap;

    // ---- pull
    const long long t_begin = clock64();
    float* w0 = ws.w0;
    if (cx.c.mode == PSB_COMM_NONE) {
        for (int i = threadIdx.x; i < n4; i += blockDim.x) ((float4*)w0)[i] = ((const float4*)w_in)[i];
        __syncthreads();
    } else if (!comm_pull(cx, clock, gw, (float4*)w0, rec)) {
        return false;
    }
    if (count <= 0) {                       // "There is no data for partition" (WorkerTrainingProcessor.java:131)
        if (threadIdx.x == 0) {
            loss_out[w] = nanf("");
            info_out[8 * w + 2] = -1, info_out[8 * w + 3] = 0;
            if (rec) rec->loss = nanf("");
        }
        if (cx.c.mode != PSB_COMM_NONE) comm_push(cx, clock, gw, nullptr);   // still takes part in the protocol
        return true;
    }

    // ---- standardisation statistics (unbiased sample std over the window)
    if (sv.standardize && sv.inv_std) {      // statistics maintained at ingestion time (psb_window_stats_update)
        for (int f = threadIdx.x; f < F; f += blockDim.x) inv[f] = sv.inv_std[(size_t)w * F + f];
    } else {
        feature_inv_std(Xw, start, count, cap, F, sv.standardize, inv);
    }
    __syncthreads();
    const long long t_stats = clock64();

    // ---- start point in scaled space: x = w0 * std
    SolveVecs v;
    ws.bind(v);
    v.n4 = n4;
    to_scaled_space(v.x, w0, inv, F);
    __syncthreads();

    SimtEval ev = {Xw, lw, start, count, cap, F, R, Ws, Rs, inv, sred, red};
    SolveResult res = local_solve(ev, v, sv.solver, sv.max_iter, sv.sgd_lr, red);
    const long long t_solve = clock64();

    // ---- back to model space, centre, delta = float(new) - old
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

__global__ void __launch_bounds__(512, 1)
psb_worker_simt_kernel(PsbSolver sv, CommCtx cx, const __half* __restrict__ X, const int* __restrict__ labels,
                       const int* __restrict__ win, const float* __restrict__ w_in, float* __restrict__ delta_out,
                       float* __restrict__ loss_out, int* __restrict__ info_out, float* __restrict__ scratch) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int nclk = (cx.c.mode == PSB_COMM_WORKER || cx.c.mode == PSB_COMM_SEQUENTIAL) && cx.c.num_clocks > 1 ? cx.c.num_clocks : 1;
    for (int it = 0; it < nclk; ++it) {
        if (!simt_clock(sv, cx, cx.c.clock + it, it, X, labels, win, w_in, delta_out, loss_out, info_out, scratch, smem))
            return;
        __syncthreads();
    }
}

static size_t simt_smem_bytes(const PsbSolver* s) {
    const int nv = (s->num_features + 1) * 8;
    return (size_t)nv * 4 + (size_t)s->cap * 32 + (size_t)s->num_features * 4 + 64 + 34 * 8;
}

// worker CTAs of this shape that can be resident on the current device at the same time
int psb_worker_simt_resident(const PsbSolver* s) {
    const size_t smem = simt_smem_bytes(s);
    if (smem > 226 * 1024) return 0;
    int dev = 0, sms = 0, per_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(psb_worker_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, psb_worker_simt_kernel, 512, smem);
    return sms * per_sm;
}

int psb_launch_worker_simt(const PsbSolver* s, const CommCtx& cx, int num_workers, const void* d_X,
                           const int* d_labels, const int* d_win, const float* d_w_in, float* d_delta,
                           float* d_loss, int* d_info, float* d_scratch, cudaStream_t stream) {
    const size_t smem = simt_smem_bytes(s);
    PSB_REQUIRE(s->num_features % 2 == 0, "num_features must be even (got %d)", s->num_features);
    if (smem > 226 * 1024) {
        psb_set_error("SIMT worker kernel needs %zu bytes of shared memory (> 227 KB): lower max_buffer_size (%d) "
                      "or num_features (%d)", smem, s->cap, s->num_features);
        return PSB_E_LIMIT;
    }
    PSB_CHECK_CUDA(cudaFuncSetAttribute(psb_worker_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    psb_worker_simt_kernel<<<num_workers, 512, smem, stream>>>(*s, cx, (const __half*)d_X, d_labels, d_win, d_w_in,
                                                                d_delta, d_loss, d_info, d_scratch);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
