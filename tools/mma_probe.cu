// Micro-benchmark: cycles per tcgen05.mma (kind::f16, cta_group::1) for the operand shapes / layouts the worker
// kernel uses and for the "swapped" formulation (the streamed rows as the wide B operand, the small weight /
// residual matrix as A).  Operands are resident in shared memory (no TMA), one elected thread issues `count`
// MMAs back to back, then commits; time = first issue -> commit arrival.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I. tools/mma_probe.cu -o tools/mma_probe -lcuda
#include <cstdio>
#include <cstdlib>
#include "../csrc/tc.cuh"
void psb_set_error(const char*, ...) {}

constexpr int CH = 16384;        // one [128 rows x 64 features] fp16 chunk
constexpr int NSLOT = 8;         // chunk slots cycled through (128 KB)

__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}

struct Cfg {
    int kind;      // 0 GEMM1 (A = rows K-major), 1 GEMM2 (A = rows MN-major), 2 swapped GEMM1 (B = rows K-major),
                   // 3 swapped GEMM2 (B = rows MN-major), 4 = kind 2 with A from TMEM, 5 = kind 3 with A from TMEM
    int M, N;
};

__global__ void __launch_bounds__(128, 1) probe(Cfg c, int groups, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* rows = smem;                         // NSLOT chunks
    unsigned char* small = smem + NSLOT * CH;           // 32 KB: the weight / residual operand
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    for (int i = threadIdx.x; i < (NSLOT * CH + 32768) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u;   // 1.0h
    if (threadIdx.x == 0) tc::mbar_init(&bar, 1), tc::mbar_fence_init();
    if (threadIdx.x < 32) tc::tmem_alloc(&tmem_slot, 512);
    tc::fence_proxy_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    const uint32_t rows0 = tc::smem_u32(rows), small0 = tc::smem_u32(small);
    long long t0 = 0, t1 = 0;
    int issued = 0;
    if (threadIdx.x < 32) {
        t0 = clock64();
        if (tc::elect_one()) {
            for (int g = 0; g < groups; ++g) {
                const uint32_t slot = rows0 + (uint32_t)(g % NSLOT) * CH;
                if (c.kind == 0) {               // 4 MMAs per chunk: A = chunk K-major SW128, B = 16 x 64 K-major SW128
                    const uint64_t ad = tc::desc_k_sw128(slot, 1024), bd = tc::desc_k_sw128(small0 + (g % 8) * 2048, 1024);
                    const uint32_t id = tc::idesc_f16(c.M, c.N, 0, 0);
                    for (int k = 0; k < 4; ++k) tc::umma_f16(tmem, ad + 2 * k, bd + 2 * k, id, 1u);
                    issued += 4;
                } else if (c.kind == 1) {        // 8 MMAs per chunk pair: A = two chunks MN-major SW128, B = Rsp MN-major none
                    const uint32_t pair = rows0 + (uint32_t)((2 * g) % NSLOT) * CH;
                    const uint64_t ad = tc::desc_mn_sw128(pair, CH, 1024), bd = tc::desc_mn_none(small0, 256, 128);
                    const uint32_t id = tc::idesc_f16(c.M, c.N, 1, 1);
                    for (int ks = 0; ks < 8; ++ks) tc::umma_f16(tmem + 16 * (g % 8), ad + (uint64_t)(ks * 128), bd + (uint64_t)(ks * 32), id, 1u);
                    issued += 8;
                } else if (c.kind == 2 || c.kind == 4) {   // swapped GEMM1: B = N rows x 64 k (K-major SW128), A = W^T [M x 64 k]
                    const uint32_t base = c.N > 128 ? rows0 + (uint32_t)((2 * g) % NSLOT) * CH : slot;
                    const uint64_t bd = tc::desc_k_sw128(base, 1024), ad = tc::desc_k_sw128(small0, 1024);
                    const uint32_t id = tc::idesc_f16(c.M, c.N, 0, 0);
                    for (int k = 0; k < 4; ++k) {
                        if (c.kind == 2) tc::umma_f16(tmem, ad + 2 * k, bd + 2 * k, id, 1u);
                        else umma_f16_ts(tmem, tmem + 256 + 8 * k, bd + 2 * k, id, 1u);
                    }
                    issued += 4;
                } else {                         // swapped GEMM2: B = 128 rows(K) x N features (MN-major SW128), A = R^T [M x 128 k]
                    const int nch = c.N / 64;
                    const uint32_t base = rows0 + (uint32_t)((nch * g) % NSLOT) * CH;
                    const uint64_t bd = tc::desc_mn_sw128(base, CH, 1024), ad = tc::desc_mn_none(small0, 1024, 128);
                    const uint32_t id = tc::idesc_f16(c.M, c.N, c.kind == 3 ? 1 : 0, 1);
                    for (int ks = 0; ks < 8; ++ks) {
                        if (c.kind == 3) tc::umma_f16(tmem, ad + (uint64_t)(ks * 128), bd + (uint64_t)(ks * 128), id, 1u);
                        else umma_f16_ts(tmem, tmem + 256 + 8 * ks, bd + (uint64_t)(ks * 128), id, 1u);
                    }
                    issued += 8;
                }
            }
            tc::umma_commit(&bar);
        }
        __syncwarp();
        const bool ok = tc::mbar_wait(&bar, 0);
        t1 = clock64();
        issued = __shfl_sync(0xffffffffu, __reduce_max_sync(0xffffffffu, issued), 0);
        if (threadIdx.x == 0) out[0] = ok ? t1 - t0 : -1, out[1] = issued;
    }
    tc::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
    const int ctas = argc > 1 ? atoi(argv[1]) : 1;
    long long* d;
    cudaMalloc(&d, 16);
    const size_t smem = 1024 + NSLOT * CH + 32768;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const Cfg cfgs[] = {{0, 128, 16}, {0, 128, 32}, {0, 128, 64},  {0, 64, 16},   {1, 128, 16},  {1, 128, 32},
                        {2, 128, 128}, {2, 128, 256}, {2, 64, 128}, {2, 64, 256},  {3, 64, 64},   {3, 64, 128},
                        {3, 64, 256},  {3, 128, 256}, {4, 128, 256}, {4, 64, 256}, {5, 64, 256},  {5, 128, 256}};
    const char* names[] = {"GEMM1  A=rows(K-major)  B=W", "GEMM2  A=rows(MN-major) B=R", "swapped GEMM1  A=W^T B=rows(K-major)",
                           "swapped GEMM2  A=R^T B=rows(MN-major)", "swapped GEMM1  A=W^T in TMEM", "swapped GEMM2  A=R^T in TMEM"};
    for (const Cfg& c : cfgs) {
        const int groups = 512;
        long long h[2] = {0, 0};
        for (int rep = 0; rep < 2; ++rep) {
            probe<<<ctas, 128, smem>>>(c, groups, d);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) {
                printf("%s M=%d N=%d: %s\n", names[c.kind], c.M, c.N, cudaGetErrorString(e));
                return 1;
            }
        }
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        // bytes of the streamed rows consumed per MMA: kinds 0/1: M x 16 x 2; swapped: N x 16 x 2
        const int row_bytes = (c.kind <= 1 ? c.M : c.N) * 32;
        if (h[0] < 0) printf("%-40s M=%3d N=%3d: TIMEOUT\n", names[c.kind], c.M, c.N);
        else printf("%-40s M=%3d N=%3d: %6.1f cycles/MMA, %5.1f row-bytes/cycle (%lld MMAs)\n", names[c.kind], c.M, c.N,
                    (double)h[0] / h[1], row_bytes * (double)h[1] / h[0], h[1]);
    }
    return 0;
}
