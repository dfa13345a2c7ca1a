// K5: test-set evaluation — argmax(X W + b) -> confusion matrix -> weighted F1 and accuracy.
// Replaces LogisticRegressionTaskSpark.calculateEvaluationMetric + Metrics.from
// (LogisticRegressionTaskSpark.java:236-251, Metrics.java:15-24: two Spark jobs per call).
// SIMT flavour: warp per row over dense fp16 X with W staged in shared memory.  The tcgen05 flavour
// (ps_eval_tc.cu) produces the same confusion matrix; both share the finalise kernel below.
#include "common.cuh"

// blockIdx.y = model when several weight sets are scored in one launch; W2 (optional) is added element-wise.
__global__ void __launch_bounds__(256) psb_eval_simt_kernel(const __half* __restrict__ X, const int* __restrict__ labels,
                                                           int n, const float* __restrict__ W,
                                                           const float* __restrict__ W2, size_t model_stride, int F,
                                                           int R, int* __restrict__ conf) {
    extern __shared__ __align__(16) unsigned char smem[];
    float* Ws = (float*)smem;
    __shared__ int cm[64];
    W += (size_t)blockIdx.y * model_stride;
    if (W2) W2 += (size_t)blockIdx.y * model_stride;
    conf += (size_t)blockIdx.y * 64;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    for (int i = tid; i < (F + 1) * 2; i += blockDim.x) {
        float4 v = ((const float4*)W)[i];
        if (W2) {
            const float4 u = ((const float4*)W2)[i];
            v.x += u.x, v.y += u.y, v.z += u.z, v.w += u.w;
        }
        ((float4*)Ws)[i] = v;
    }
    if (tid < 64) cm[tid] = 0;
    __syncthreads();
    const float4* W4 = (const float4*)Ws;
    for (int row = blockIdx.x * nwarp + warp; row < n; row += gridDim.x * nwarp) {
        const __half2* xr = (const __half2*)(X + (size_t)row * F);
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
        if (lane == 0) {
            int best = 0;
            float bv = z[0] + Ws[F * 8];
#pragma unroll
            for (int c = 1; c < 8; ++c) {
                const float v = z[c] + Ws[F * 8 + c];
                if (c < R && v > bv) bv = v, best = c;       // first maximum wins, like argmax
            }
            int y = labels[row];
            y = y < 0 ? 0 : (y > 7 ? 7 : y);
            atomicAdd(&cm[y * 8 + best], 1);
        }
    }
    __syncthreads();
    if (tid < 64 && cm[tid]) atomicAdd(&conf[tid], cm[tid]);
}

// Support-weighted F1 over the true-label classes + accuracy, from the 8x8 confusion matrix (one block per model).
__global__ void psb_eval_finalize_kernel(const int* __restrict__ conf, float* __restrict__ out) {
    conf += (size_t)blockIdx.x * 64, out += (size_t)blockIdx.x * 2;
    const int c = threadIdx.x;                 // 32 threads; classes 0..7
    float tp = 0.f, sup = 0.f, pred = 0.f;
    if (c < 8) {
        tp = (float)conf[c * 8 + c];
        for (int k = 0; k < 8; ++k) sup += (float)conf[c * 8 + k], pred += (float)conf[k * 8 + c];
    }
    const float prec = pred > 0.f ? tp / pred : 0.f, rec = sup > 0.f ? tp / sup : 0.f;
    float f1w = (prec + rec > 0.f ? 2.f * prec * rec / (prec + rec) : 0.f) * sup;
    float total = sup, correct = tp;
    for (int o = 16; o; o >>= 1) {
        f1w += __shfl_xor_sync(0xffffffffu, f1w, o);
        total += __shfl_xor_sync(0xffffffffu, total, o);
        correct += __shfl_xor_sync(0xffffffffu, correct, o);
    }
    if (c == 0) {
        out[0] = total > 0.f ? f1w / total : 0.f;
        out[1] = total > 0.f ? correct / total : 0.f;
    }
}

int psb_launch_eval_simt(const void* d_X, const int* d_labels, int n, const float* d_w, const float* d_w2,
                         size_t model_stride, int num_models, int F, int R, int* d_conf, cudaStream_t stream) {
    const size_t smem = (size_t)(F + 1) * 32;
    if (smem > 200 * 1024) {
        psb_set_error("evaluation kernel: num_features %d does not fit shared memory", F);
        return PSB_E_LIMIT;
    }
    PSB_CHECK_CUDA(cudaFuncSetAttribute(psb_eval_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int blocks = (n + 7) / 8;
    const int cap_blocks = num_models > 1 ? 16 : 148 * 2;        // many models: the models fill the device
    if (blocks > cap_blocks) blocks = cap_blocks;
    if (blocks < 1) blocks = 1;
    psb_eval_simt_kernel<<<dim3(blocks, num_models), 256, smem, stream>>>((const __half*)d_X, d_labels, n, d_w, d_w2,
                                                                           model_stride, F, R, d_conf);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int psb_launch_eval_finalize(const int* d_conf, float* d_out, int num_models, cudaStream_t stream) {
    psb_eval_finalize_kernel<<<num_models, 32, 0, stream>>>(d_conf, d_out);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
