// Micro-benchmark: per-SM TMA streaming throughput, 2-D tensor-map boxes (128 rows x 128 B, row stride
// 2 KB) versus 1-D bulk copies of contiguous 16 KB tiles.  One producer thread, one consumer thread that
// releases each slot immediately.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I. tools/tma_probe.cu -o /tmp/tma_probe -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../csrc/tc.cuh"
void psb_set_error(const char*, ...) {}
int psb_make_tmap_2d_f16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols) {
    cuuint64_t gdim[2] = {cols, rows}; cuuint64_t gstride[1] = {cols * 2}; cuuint32_t box[2] = {box_cols, box_rows}; cuuint32_t estr[2] = {1, 1};
    return cuTensorMapEncodeTiled(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}
constexpr int CH = 16384;
template <int MODE>
__global__ void __launch_bounds__(128, 1) probe(const __grid_constant__ CUtensorMap tm, const unsigned char* base, int rows_per_cta, int F, int NS, long long* cycles, int passes = 1, int cta_rows_mod = 1 << 30) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full = (uint64_t*)(smem + (size_t)NS * CH); uint64_t* empty = full + NS;
    if (threadIdx.x == 0) { for (int s = 0; s < NS; ++s) tc::mbar_init(full + s, 1), tc::mbar_init(empty + s, 1); tc::mbar_fence_init(); }
    __syncthreads();
    const int nkb = F / 64, nslab = rows_per_cta / 128, per_pass = nkb * nslab, total = per_pass * passes;
    const int row0 = (blockIdx.x % cta_rows_mod) * rows_per_cta;
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        for (int i = 0; i < total; ++i) {
            const int slot = i % NS; const unsigned use = i / NS;
            if (use > 0) while (!tc::mbar_try_wait(empty + slot, (use - 1) & 1)) {}
            tc::mbar_arrive_expect_tx(full + slot, CH);
            const int sl = (i % per_pass) / nkb, c = i % nkb;
            if (MODE == 0) tc::tma_load_2d(smem + (size_t)slot * CH, &tm, c * 64, row0 + sl * 128, full + slot);
            else {
                const unsigned char* src = base + ((size_t)blockIdx.x * total + i) * CH;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tc::smem_u32(smem + (size_t)slot * CH)), "l"(src), "r"(CH), "r"(tc::smem_u32(full + slot)) : "memory");
            }
        }
    } else if (threadIdx.x == 32) {
        for (int i = 0; i < total; ++i) {
            const int slot = i % NS;
            while (!tc::mbar_try_wait(full + slot, (i / NS) & 1)) {}
            tc::mbar_arrive(empty + slot);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}
int main() {
    const int F = 1024, rows = 1024, ctas_max = 148;
    size_t bytes = (size_t)ctas_max * rows * F * 2;
    unsigned char* d; cudaMalloc(&d, bytes); cudaMemset(d, 1, bytes);
    long long* dc; cudaMalloc(&dc, ctas_max * 8);
    CUtensorMap tm; cuInit(0); int rc = psb_make_tmap_2d_f16(&tm, d, (uint64_t)ctas_max * rows, F, 128, 64); if (rc) { printf("tmap rc %d\n", rc); return 1; }
    for (int mode = 0; mode < 2; ++mode) for (int NS : {4, 10, 13}) for (int ctas : {1, 37, 148}) {
        size_t smem = 1024 + (size_t)NS * CH + 2 * NS * 8 + 64;
        if (mode == 0) cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); else cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) probe<0><<<ctas, 128, smem>>>(tm, d, rows, F, NS, dc); else probe<1><<<ctas, 128, smem>>>(tm, d, rows, F, NS, dc);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        cudaError_t err = cudaGetLastError();
        double gb = (double)ctas * rows * F * 2 / 1e9;
        printf("mode=%s NS=%d ctas=%d: %.1f us, %.1f GB/s total, %.1f GB/s per SM (%s)\n", mode ? "bulk1d" : "tensor2d", NS, ctas, ms * 1e3, gb / (ms * 1e-3), gb / (ms * 1e-3) / ctas, cudaGetErrorString(err));
    }
    // Long runs: `passes` sweeps over each CTA's own region.  region = 2 MB per CTA (296 MB, DRAM-bound) versus
    // 256 KB per CTA (38-76 MB, L2-resident after the first sweep): the second is the L2 -> SM streaming ceiling.
    cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + 4 * CH + 64 + 64);
    for (int ctas : {148, 296}) for (int rows_cta : {1024, 128}) {
        const int passes = rows_cta == 1024 ? 40 : 320;
        const int NS = 4; size_t smem = 1024 + (size_t)NS * CH + 2 * NS * 8 + 64;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        probe<0><<<ctas, 128, smem>>>(tm, d, rows_cta, F, NS, dc, 2, 148);
        cudaEventRecord(e0);
        probe<0><<<ctas, 128, smem>>>(tm, d, rows_cta, F, NS, dc, passes, 148);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double gb = (double)ctas * rows_cta * F * 2 * passes / 1e9;
        printf("long ctas=%d region=%d KB/CTA (footprint %.0f MB) passes=%d: %.1f us, %.2f TB/s total, %.1f GB/s per SM (%s)\n", ctas, rows_cta * 2, 148.0 * rows_cta * F * 2 / 1e6, passes, ms * 1e3, gb / ms, gb / (ms * 1e-3) / 148, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
