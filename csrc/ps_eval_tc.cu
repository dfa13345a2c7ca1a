// K5, tensor flavour: test-set evaluation as a TMA-fed tcgen05 GEMM with the argmax / confusion-matrix
// epilogue fused.  logits[128 rows x 16] = X_tile[128 x F] (fp16, streamed by TMA, SWIZZLE_128B) times
// Wsp[F x 16] where columns 0-7 hold the fp16 high parts of the 8 class columns and 8-15 the low parts
// (hi + lo carries ~22 bits of the fp32 weight); the accumulator lives in TMEM (128 lanes x 16 columns).
// One CTA per 128 rows and model; W is staged once per CTA and stays resident in shared memory.  blockIdx.y picks the
// model when several weight sets are scored in one launch (the worker log rows: every locally fitted model is
// scored on the test set, LogisticRegressionTaskSpark.java:186) — the test set is then served from L2.
// Replaces LogisticRegressionTaskSpark.calculateEvaluationMetric (LogisticRegressionTaskSpark.java:236-251).
#define PSB_HAVE_TC_EVAL 1
#include "tc.cuh"

#include <mutex>

// ------------------------------------------------------------------ host: tensor-map encoder
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled g_encode = nullptr;
static std::once_flag g_encode_once;

int psb_make_tmap_2d_f16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                         uint32_t box_cols) {
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            g_encode = (PFN_encodeTiled)fn;
    });
    if (!g_encode) {
        psb_set_error("cuTensorMapEncodeTiled is not available from this driver");
        return PSB_E_CUDA;
    }
    if ((uintptr_t)base % 16 != 0 || (cols * 2) % 16 != 0 || box_cols * 2 != 128) {
        psb_set_error("tensor map: base/stride must be 16-byte aligned and the box 128 bytes wide");
        return PSB_E_INVALID;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {cols * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)base, gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        psb_set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return PSB_E_CUDA;
    }
    return 0;
}

// ------------------------------------------------------------------ kernel
namespace {
constexpr int kStages = 4;
constexpr int kTileM = 128;
constexpr int kBlockK = 64;                       // fp16 elements per 128-byte swizzle row
constexpr int kABytes = kTileM * kBlockK * 2;     // 16 KB per stage
constexpr int kBChunkBytes = 16 * kBlockK * 2;    // 2 KB of Wsp per K block

// Stage Wsp (hi/lo split, K-major SWIZZLE_128B, 16 rows per K block) from fp32 W[(F+1)][8].
// W2 (optional) is added element-wise: a worker's local model is its start weights plus its delta.
__device__ __forceinline__ void stage_wsp(unsigned char* sB, const float* __restrict__ W, const float* __restrict__ W2,
                                          int F) {
    for (int f = threadIdx.x; f < F; f += blockDim.x) {
        float4 a = ((const float4*)W)[2 * f], b = ((const float4*)W)[2 * f + 1];
        if (W2) {
            const float4 a2 = ((const float4*)W2)[2 * f], b2 = ((const float4*)W2)[2 * f + 1];
            a.x += a2.x, a.y += a2.y, a.z += a2.z, a.w += a2.w, b.x += b2.x, b.y += b2.y, b.z += b2.z, b.w += b2.w;
        }
        const float w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        const int kb = f >> 6, k = f & 63, j = k >> 3, e = k & 7;
        unsigned char* chunk = sB + (size_t)kb * kBChunkBytes;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            __half hi, lo;
            tc::split_half(w[c], hi, lo);
            // row n = c (hi) lives in atom 0, row n = 8 + c (lo) in atom 1; row-in-atom = c
            *(__half*)(chunk + tc::sw128_off(c, j) + e * 2) = hi;
            *(__half*)(chunk + 1024 + tc::sw128_off(c, j) + e * 2) = lo;
        }
    }
}
}  // namespace

__global__ void __launch_bounds__(128, 1)
psb_eval_tc_kernel(const __grid_constant__ CUtensorMap tmX, const int* __restrict__ labels, int n,
                   const float* __restrict__ W, const float* __restrict__ W2, size_t model_stride, int F, int R,
                   int* __restrict__ conf) {
    W += (size_t)blockIdx.y * model_stride;
    if (W2) W2 += (size_t)blockIdx.y * model_stride;
    conf += (size_t)blockIdx.y * 64;
    // conf[63] (class 7 x class 7) can never be hit by labels 0..5: it doubles as the model's error word
    unsigned int* err = (unsigned int*)conf + 63;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int nkb = F / kBlockK;
    unsigned char* sA = smem;                                  // kStages x 16 KB
    unsigned char* sB = smem + kStages * kABytes;              // nkb x 2 KB
    uint64_t* bars = (uint64_t*)(sB + (size_t)nkb * kBChunkBytes);
    uint64_t* full = bars;                                     // [kStages]
    uint64_t* empty = bars + kStages;                          // [kStages]
    uint64_t* acc_full = bars + 2 * kStages;
    uint32_t* tmem_slot = (uint32_t*)(bars + 2 * kStages + 1);
    float* sbias = (float*)(tmem_slot + 2);
    int* cm = (int*)(sbias + 8);                               // [64]
    __shared__ int s_fail;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row0 = blockIdx.x * kTileM;

    if (tid == 0) {
        s_fail = 0;
        tc::tma_prefetch_desc(&tmX);
        for (int s = 0; s < kStages; ++s) tc::mbar_init(full + s, 1), tc::mbar_init(empty + s, 1);
        tc::mbar_init(acc_full, 1);
        tc::mbar_fence_init();
    }
    if (tid < 64) cm[tid] = 0;
    if (tid < 8) sbias[tid] = W[F * 8 + tid] + (W2 ? W2[F * 8 + tid] : 0.f);
    if (warp == 0) tc::tmem_alloc(tmem_slot, 32);
    stage_wsp(sB, W, W2, F);
    tc::fence_proxy_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {                                       // ---- TMA producer
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kStages;
                if (kb >= kStages && !tc::mbar_wait(empty + s, ((kb / kStages) - 1) & 1)) {
                    s_fail = 1;
                    break;
                }
                tc::mbar_arrive_expect_tx(full + s, kABytes);
                tc::tma_load_2d(sA + s * kABytes, &tmX, kb * kBlockK, row0, full + s);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {                                       // ---- MMA issuer
            constexpr uint32_t idesc = tc::idesc_f16(kTileM, 16, 0, 0);
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kStages;
                if (!tc::mbar_wait(full + s, (kb / kStages) & 1)) {
                    s_fail = 1;
                    break;
                }
                tc::tc_fence_after();
                const uint64_t ad = tc::desc_k_sw128(tc::smem_u32(sA + s * kABytes), 1024);
                const uint64_t bd = tc::desc_k_sw128(tc::smem_u32(sB + (size_t)kb * kBChunkBytes), 1024);
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k)         // +32 bytes along K per UMMA_K = 16
                    tc::umma_f16(tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | k) ? 1u : 0u);
                tc::umma_commit(empty + s);
            }
            tc::umma_commit(acc_full);
        }
        __syncwarp();
    }
    // ---- epilogue: all four warps, thread = row (TMEM lane)
    const bool ok = tc::mbar_wait(acc_full, 0);
    tc::tc_fence_after();
    if (ok && !s_fail) {
        float v[16];
        tc::tmem_ld_32x32b_x16(tmem + ((uint32_t)(warp * 32) << 16), v);
        const int row = row0 + tid;
        if (row < n) {
            int best = 0;
            float bv = v[0] + v[8] + sbias[0];
#pragma unroll
            for (int c = 1; c < 8; ++c) {
                const float z = v[c] + v[8 + c] + sbias[c];
                if (c < R && z > bv) bv = z, best = c;
            }
            int y = labels[row];
            y = y < 0 ? 0 : (y > 7 ? 7 : y);
            atomicAdd(&cm[y * 8 + best], 1);
        }
    } else if (tid == 0) {
        atomicExch(err, 100u);
    }
    tc::tc_fence_before();
    __syncthreads();
    if (tid < 64 && cm[tid]) atomicAdd(&conf[tid], cm[tid]);
    if (warp == 0) tc::tmem_dealloc(tmem, 32);
}

int psb_launch_eval_tc(const void* d_X, const int* d_labels, int n, const float* d_w, const float* d_w2,
                       size_t model_stride, int num_models, int F, int R, int* d_conf, cudaStream_t stream) {
    if (F % kBlockK != 0) {
        psb_set_error("tcgen05 evaluation kernel needs num_features %% 64 == 0 (got %d)", F);
        return PSB_E_INVALID;
    }
    const size_t smem = 1024 + (size_t)kStages * kABytes + (size_t)(F / kBlockK) * kBChunkBytes + 1024;
    if (smem > 220 * 1024) {
        psb_set_error("tcgen05 evaluation kernel: num_features %d does not fit shared memory", F);
        return PSB_E_LIMIT;
    }
    CUtensorMap tm;
    int rc = psb_make_tmap_2d_f16(&tm, d_X, (uint64_t)n, (uint64_t)F, kTileM, kBlockK);
    if (rc) return rc;
    PSB_CHECK_CUDA(cudaFuncSetAttribute(psb_eval_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    psb_eval_tc_kernel<<<dim3((n + kTileM - 1) / kTileM, num_models), 128, smem, stream>>>(
        tm, d_labels, n, d_w, d_w2, model_stride, F, R, d_conf);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
