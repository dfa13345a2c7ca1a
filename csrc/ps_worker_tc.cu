// K2, tensor flavour: the worker's local solve with every pass over the input buffer running on
// tcgen05 tensor cores fed by TMA, accumulators in TMEM.  One CTA per worker, one CTA per SM.
//
// One loss/gradient evaluation streams the buffer ONCE, 64 rows (one "slab") at a time:
//   TMA     : slab = F/64 chunks of [64 rows x 64 features] fp16 (8 KB, SWIZZLE_128B) into a shared ring
//   GEMM1   : Z[64 x 16]   = slab (K-major A, M=64) x Wsp[F x 16]   (fp16 hi | lo halves of the weights)
//   epilogue: TMEM -> registers, + intercept, softmax over the R class rows, loss, residual r = p - y,
//             r -> fp16 hi | lo -> shared memory as a 16 x 64 MN-major B operand
//   GEMM2   : G[F x 16]   += slab^T (the SAME shared tiles read MN-major, M=128 features) x Rsp[64 x 16]
// so the gradient X^T R needs neither atomics nor a second read of X.  After the last slab the G
// accumulators (F/128 tiles x 16 TMEM columns) are drained, scaled by 1/(n std_f) and written to the
// solver's gradient vector.  A first pass with the same TMA ring computes the per-feature standard
// deviation (Spark standardisation) straight from shared memory.
//
// The solver driving these evaluations (L-BFGS + strong-Wolfe, or SGD) is solver.cuh; pull / push /
// apply over peer memory is comm.cuh.  Replaces LogisticRegressionTaskSpark.calculateGradients
// (LogisticRegressionTaskSpark.java:142-221) and the topics around it.
#define PSB_HAVE_TC_WORKER 1
#include "comm.cuh"
#include "solver.cuh"
#include "tc.cuh"
#include "worker_common.cuh"

namespace {
constexpr int kThreads = 256;
constexpr int kSlabRows = 64;
constexpr int kBlockK = 64;                          // features per chunk (128-byte swizzle row)
constexpr int kChunkBytes = kSlabRows * kBlockK * 2;  // 8 KB
constexpr int kBChunkBytes = 16 * kBlockK * 2;        // 2 KB of Wsp per chunk
constexpr int kRBytes = 2048;                         // Rsp: 8 k-groups x (2 n-groups x 128 B)
constexpr int kTmemCols = 256;
constexpr int kColZ = 0;
constexpr int kColG = 32;

struct TcShared {
    unsigned char* ring;      // NS x 8 KB, 1024-aligned
    unsigned char* sB;        // nkb x 2 KB
    unsigned char* sR;        // 2 KB
    float* inv;               // [F]
    float* sred;              // [16]
    double* red;              // [34]
    uint64_t* full;           // [NS]
    uint64_t* empty;          // [NS]
    uint64_t* z_full;
    uint64_t* r_full;
    uint64_t* g_full;
    uint32_t* tmem_slot;
    volatile int* fail;
    volatile unsigned int* chunk_base;   // chunks streamed so far (ring position), shared by all roles
};

__device__ __forceinline__ bool wait_or_fail(uint64_t* bar, uint32_t parity, volatile int* fail) {
    const unsigned long long t0 = clock64();
    while (!tc::mbar_try_wait(bar, parity)) {
        if (*fail) return false;
        if ((unsigned long long)(clock64() - t0) > 4000000000ull) {
            *fail = 1;
            return false;
        }
    }
    return true;
}

struct TcEval {
    const CUtensorMap* tmX;
    const int* labels;        // this worker's [cap]
    int row_base;             // w * cap (row coordinate of slot 0 in the tensor map)
    int start, count, cap, F, R, NS;
    uint32_t tmem;
    TcShared sh;

    __device__ __forceinline__ bool row_valid(int slot) const {
        int d = slot - start;
        if (d < 0) d += cap;
        return d < count;
    }
    __device__ __forceinline__ bool slab_active(int b) const {
        // window = [start, start+count) mod cap; slab = [64b, 64b+64)
        const int lo = b * kSlabRows, hi = lo + kSlabRows;
        const int e = start + count;
        if (e <= cap) return lo < e && hi > start;
        return hi > start || lo < e - cap;
    }

    // ---- TMA producer (one thread): all chunks of all active slabs
    __device__ void produce() const {
        const int nkb = F / kBlockK;
        unsigned int pc = *sh.chunk_base;
        for (int b = 0; b < cap / kSlabRows; ++b) {
            if (!slab_active(b)) continue;
            for (int c = 0; c < nkb; ++c, ++pc) {
                const int slot = pc % NS;
                const unsigned use = pc / NS;
                if (use > 0 && !wait_or_fail(sh.empty + slot, (use - 1) & 1, sh.fail)) return;
                tc::mbar_arrive_expect_tx(sh.full + slot, kChunkBytes);
                tc::tma_load_2d(sh.ring + (size_t)slot * kChunkBytes, tmX, c * kBlockK, row_base + b * kSlabRows,
                                sh.full + slot);
            }
        }
    }

    // ---- statistics pass consumer: warp `wi` of 4 takes chunks c with c % 4 == wi; lane = feature pair
    __device__ void stats_consume(int wi, int lane) const {
        const int nkb = F / kBlockK;
        const unsigned int base = *sh.chunk_base;
        float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
        float q0[4] = {0.f, 0.f, 0.f, 0.f}, q1[4] = {0.f, 0.f, 0.f, 0.f};
        unsigned int q = 0;
        for (int b = 0; b < cap / kSlabRows; ++b) {
            if (!slab_active(b)) continue;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c = wi + 4 * k;
                if (c >= nkb) break;
                const unsigned int pc = base + q * nkb + c;
                const int slot = pc % NS;
                if (!wait_or_fail(sh.full + slot, (pc / NS) & 1, sh.fail)) return;
                const unsigned char* tile = sh.ring + (size_t)slot * kChunkBytes;
#pragma unroll 8
                for (int r = 0; r < kSlabRows; ++r) {
                    if (!row_valid(b * kSlabRows + r)) continue;
                    const __half2 h = *(const __half2*)(tile + tc::sw128_off(r, lane >> 2) + (lane & 3) * 4);
                    const float2 x = __half22float2(h);
                    s0[k] += x.x, q0[k] = fmaf(x.x, x.x, q0[k]);
                    s1[k] += x.y, q1[k] = fmaf(x.y, x.y, q1[k]);
                }
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(sh.empty + slot);
            }
            ++q;
        }
        const float n = (float)count;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int c = wi + 4 * k;
            if (c >= nkb) break;
            const int f = c * kBlockK + 2 * lane;
            const float v0 = count > 1 ? (q0[k] - s0[k] * s0[k] / n) / (n - 1.f) : 0.f;
            const float v1 = count > 1 ? (q1[k] - s1[k] * s1[k] / n) / (n - 1.f) : 0.f;
            sh.inv[f] = v0 > 1e-12f ? rsqrtf(v0) : 0.f;
            sh.inv[f + 1] = v1 > 1e-12f ? rsqrtf(v1) : 0.f;
        }
    }

    // ---- MMA issuer (one thread)
    __device__ void mma_loop() {
        const int nkb = F / kBlockK;
        constexpr uint32_t idesc1 = tc::idesc_f16(64, 16, 0, 0);     // GEMM1: A K-major, B K-major
        constexpr uint32_t idesc2 = tc::idesc_f16(128, 16, 1, 1);    // GEMM2: A MN-major, B MN-major
        const unsigned int base = *sh.chunk_base;
        unsigned int q = 0;
        const uint32_t sB0 = tc::smem_u32(sh.sB), sR0 = tc::smem_u32(sh.sR), ring0 = tc::smem_u32(sh.ring);
        for (int b = 0; b < cap / kSlabRows; ++b) {
            if (!slab_active(b)) continue;
            const unsigned int pc0 = base + q * nkb;
            // GEMM1 over the slab's chunks as they land
            for (int c = 0; c < nkb; ++c) {
                const unsigned int pc = pc0 + c;
                const int slot = pc % NS;
                if (!wait_or_fail(sh.full + slot, (pc / NS) & 1, sh.fail)) return;
                tc::tc_fence_after();
                const uint64_t ad = tc::desc_k_sw128(ring0 + slot * kChunkBytes, 1024);
                const uint64_t bd = tc::desc_k_sw128(sB0 + c * kBChunkBytes, 1024);
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k)
                    tc::umma_f16(tmem + kColZ, ad + 2 * k, bd + 2 * k, idesc1, (c | k) ? 1u : 0u);
            }
            tc::umma_commit(sh.z_full);
            // residuals of this slab are written by the epilogue warps
            if (!wait_or_fail(sh.r_full, rphase & 1, sh.fail)) return;
            ++rphase;
            tc::tc_fence_after();
            // GEMM2: G[128 features x 16] += slab^T[128 x 16 rows] x Rsp[16 rows x 16], 4 K-steps of 16 rows
            for (int j = 0; j < nkb / 2; ++j) {
                const int slot = (pc0 + 2 * j) % NS;            // chunks 2j and 2j+1 are adjacent in the ring
#pragma unroll
                for (int ks = 0; ks < kSlabRows / 16; ++ks) {
                    const uint64_t ad = tc::desc_mn_sw128(ring0 + slot * kChunkBytes + ks * 2048, kChunkBytes, 1024);
                    const uint64_t bd = tc::desc_mn_none(sR0 + ks * 512, 256, 128);
                    tc::umma_f16(tmem + kColG + 16 * j, ad, bd, idesc2, (q | ks) ? 1u : 0u);
                }
            }
            for (int c = 0; c < nkb; ++c) tc::umma_commit(sh.empty + (pc0 + c) % NS);
            ++q;
        }
        tc::umma_commit(sh.g_full);
    }

    // ---- epilogue warps (4 warps; wi = TMEM sub-partition = warp % 4)
    __device__ void epilogue(int wi, int lane, float* g, float invn) {
        const int nkb = F / kBlockK;
        float lsum = 0.f, rs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float* bias = sh.sred + 16;            // intercepts staged behind the reduction words
        for (int b = 0; b < cap / kSlabRows; ++b) {
            if (!slab_active(b)) continue;
            if (!wait_or_fail(sh.z_full, zphase & 1, sh.fail)) return;
            ++zphase;
            tc::tc_fence_after();
            float v[16];
            tc::tmem_ld_32x32b_x16(tmem + kColZ + ((uint32_t)(wi * 32) << 16), v);
            if (lane < 16) {                         // M=64: rows 16*wi .. 16*wi+15 sit in lanes 0..15
                const int r = wi * 16 + lane;
                const int slot = b * kSlabRows + r;
                float rr[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (row_valid(slot)) {
                    float z[8];
#pragma unroll
                    for (int c = 0; c < 8; ++c) z[c] = v[c] + v[8 + c];
                    lsum += softmax_residual(z, bias, labels[slot], R, rr);
#pragma unroll
                    for (int c = 0; c < 8; ++c) rs[c] += rr[c];
                }
                // Rsp as MN-major no-swizzle B: core matrix (k-group kg, n-group ng) at kg*256 + ng*128,
                // row k%8 at +16 bytes each; n-group 0 = fp16 high parts, n-group 1 = low parts
                __align__(16) __half hi[8], lo[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) tc::split_half(rr[c], hi[c], lo[c]);
                unsigned char* dst = sh.sR + (r >> 3) * 256 + (r & 7) * 16;
                *(uint4*)dst = *(const uint4*)hi;
                *(uint4*)(dst + 128) = *(const uint4*)lo;
            }
            tc::fence_proxy_async_smem();
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(sh.r_full);
        }
        if (lane < 16) {
            atomicAdd(&sh.sred[0], lsum);
#pragma unroll
            for (int c = 0; c < 8; ++c) atomicAdd(&sh.sred[1 + c], rs[c]);
        }
        // drain the gradient accumulators: thread = feature row of each 128-feature tile
        if (!wait_or_fail(sh.g_full, gphase & 1, sh.fail)) return;
        ++gphase;
        tc::tc_fence_after();
        for (int j = 0; j < nkb / 2; ++j) {
            float v[16];
            tc::tmem_ld_32x32b_x16(tmem + kColG + 16 * j + ((uint32_t)(wi * 32) << 16), v);
            const int f = j * 128 + wi * 32 + lane;
            const float sc = sh.inv[f] * invn;
            float4* g4 = (float4*)g + 2 * f;
            g4[0] = make_float4((v[0] + v[8]) * sc, (v[1] + v[9]) * sc, (v[2] + v[10]) * sc, (v[3] + v[11]) * sc);
            g4[1] = make_float4((v[4] + v[12]) * sc, (v[5] + v[13]) * sc, (v[6] + v[14]) * sc, (v[7] + v[15]) * sc);
        }
        tc::tc_fence_before();
    }

    unsigned int zphase, gphase, rphase;   // per-thread running barrier phases (persist across evaluations)

    __device__ int num_active_slabs() const {
        int n = 0;
        for (int b = 0; b < cap / kSlabRows; ++b) n += slab_active(b) ? 1 : 0;
        return n;
    }

    // ---- statistics pass (all threads enter)
    __device__ void stats_pass(int standardize) {
        const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
        if (!standardize) {
            for (int f = tid; f < F; f += blockDim.x) sh.inv[f] = 1.f;
            __syncthreads();
            return;
        }
        if (warp == 0) {
            if (lane == 0) produce();
            __syncwarp();
        } else if (warp >= 2 && warp < 6) {
            stats_consume(warp - 2, lane);
        }
        __syncthreads();
        if (tid == 0) *sh.chunk_base += (unsigned)num_active_slabs() * (F / kBlockK);
        __syncthreads();
    }

    // ---- loss + gradient at x (scaled space); all threads enter, all get the same loss
    __device__ double eval(const float* x, float* g) {
        const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
        // Wsp = (x * inv) split into fp16 hi | lo, K-major SWIZZLE_128B, 16 rows per 64-feature chunk
        for (int f = tid; f < F; f += blockDim.x) {
            const float4 a = ((const float4*)x)[2 * f], b = ((const float4*)x)[2 * f + 1];
            const float sc = sh.inv[f];
            const float w[8] = {a.x * sc, a.y * sc, a.z * sc, a.w * sc, b.x * sc, b.y * sc, b.z * sc, b.w * sc};
            const int kb = f >> 6, k = f & 63, j = k >> 3, e = k & 7;
            unsigned char* chunk = sh.sB + (size_t)kb * kBChunkBytes;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                __half hi, lo;
                tc::split_half(w[c], hi, lo);
                *(__half*)(chunk + tc::sw128_off(c, j) + e * 2) = hi;
                *(__half*)(chunk + 1024 + tc::sw128_off(c, j) + e * 2) = lo;
            }
        }
        if (tid < 16) sh.sred[tid] = 0.f;
        if (tid < 8) sh.sred[16 + tid] = x[F * 8 + tid];          // intercepts (unscaled)
        tc::fence_proxy_async_smem();
        __syncthreads();
        const float invn = 1.f / (float)count;
        if (warp == 0) {
            if (lane == 0) produce();
            __syncwarp();
        } else if (warp == 1) {
            if (lane == 0) mma_loop();
            __syncwarp();
        } else if (warp < 6) {
            epilogue(warp & 3, lane, g, invn);
        }
        __syncthreads();
        if (tid == 0) *sh.chunk_base += (unsigned)num_active_slabs() * (F / kBlockK);
        if (tid < 8) g[F * 8 + tid] = sh.sred[1 + tid] * invn;
        const double loss = (double)sh.sred[0] * (double)invn;
        __syncthreads();
        return loss;
    }
};
}  // namespace

__global__ void __launch_bounds__(kThreads, 1)
psb_worker_tc_kernel(const __grid_constant__ CUtensorMap tmX, PsbSolver sv, CommCtx cx, int NS,
                     const int* __restrict__ labels, const int* __restrict__ win, const float* __restrict__ w_in,
                     float* __restrict__ delta_out, float* __restrict__ loss_out, int* __restrict__ info_out,
                     float* __restrict__ scratch) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int F = sv.num_features, R = sv.num_class_rows, cap = sv.cap;
    const int nv = (F + 1) * 8, n4 = nv / 4, nkb = F / kBlockK;
    const int w = blockIdx.x, tid = threadIdx.x, warp = tid >> 5;
    __shared__ int s_fail;
    __shared__ unsigned int s_chunk_base;

    TcShared sh;
    sh.ring = smem;
    sh.sB = sh.ring + (size_t)NS * kChunkBytes;
    sh.sR = sh.sB + (size_t)nkb * kBChunkBytes;
    sh.inv = (float*)(sh.sR + kRBytes);
    sh.sred = sh.inv + F;
    sh.red = (double*)(sh.sred + 32);
    sh.full = (uint64_t*)(sh.red + 34);
    sh.empty = sh.full + NS;
    sh.z_full = sh.empty + NS;
    sh.r_full = sh.z_full + 1;
    sh.g_full = sh.r_full + 1;
    sh.tmem_slot = (uint32_t*)(sh.g_full + 1);
    sh.fail = &s_fail;
    sh.chunk_base = &s_chunk_base;

    if (tid == 0) {
        s_fail = 0, s_chunk_base = 0;
        tc::tma_prefetch_desc(&tmX);
        for (int s = 0; s < NS; ++s) tc::mbar_init(sh.full + s, 1), tc::mbar_init(sh.empty + s, 1);
        tc::mbar_init(sh.z_full, 1);
        tc::mbar_init(sh.r_full, 4);
        tc::mbar_init(sh.g_full, 1);
        tc::mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(sh.tmem_slot, kTmemCols);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *sh.tmem_slot;

    WorkerScratch ws(scratch + (size_t)w * PSB_SCRATCH_VECS * nv, nv);
    const int head = win[2 * w], count = win[2 * w + 1];
    int start = head - count;
    if (start < 0) start += cap;
    const int* lw = labels + (size_t)w * cap;

    // ---- pull
    float* w0 = ws.w0;
    bool alive = true;
    if (cx.c.mode == PSB_COMM_NONE) {
        for (int i = tid; i < n4; i += blockDim.x) ((float4*)w0)[i] = ((const float4*)w_in)[i];
        __syncthreads();
    } else {
        alive = comm_pull(cx, (float4*)w0);
    }
    if (alive && count <= 0) {
        if (tid == 0) {
            loss_out[w] = nanf("");
            info_out[4 * w + 2] = -1, info_out[4 * w + 3] = 0;
        }
        if (cx.c.mode != PSB_COMM_NONE) {
            for (int i = tid; i < n4; i += blockDim.x) ((float4*)ws.vec[5])[i] = make_float4(0, 0, 0, 0);
            __syncthreads();
            comm_push(cx, (const float4*)ws.vec[5]);
        }
        alive = false;
    }
    if (alive) {
        TcEval ev;
        ev.tmX = &tmX, ev.labels = lw, ev.row_base = w * cap;
        ev.start = start, ev.count = count, ev.cap = cap, ev.F = F, ev.R = R, ev.NS = NS;
        ev.tmem = tmem, ev.sh = sh, ev.zphase = 0, ev.gphase = 0, ev.rphase = 0;

        ev.stats_pass(sv.standardize);

        SolveVecs v;
        ws.bind(v);
        v.n4 = n4;
        to_scaled_space(v.x, w0, sh.inv, F);
        __syncthreads();

        SolveResult res = local_solve(ev, v, sv.solver, sv.max_iter, sv.sgd_lr, sh.red);

        float* delta = v.tmp;
        finish_delta(delta, v.x, w0, sh.inv, F, R, sv.center && sv.solver == PSB_SOLVER_LBFGS);
        __syncthreads();
        if (tid == 0) {
            loss_out[w] = s_fail ? nanf("") : (float)res.loss;
            info_out[4 * w + 0] = res.evals, info_out[4 * w + 1] = res.iterations;
            info_out[4 * w + 2] = s_fail ? -2 : res.status, info_out[4 * w + 3] = count;
        }
        if (cx.c.mode == PSB_COMM_NONE) {
            float4* out = (float4*)(delta_out + (size_t)w * nv);
            for (int i = tid; i < n4; i += blockDim.x) out[i] = ((const float4*)delta)[i];
        } else {
            comm_push(cx, (const float4*)delta);
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, kTmemCols);
}

int psb_launch_worker_tc(const PsbSolver* s, const CommCtx& cx, int num_workers, const void* d_X, const int* d_labels,
                         const int* d_win, const float* d_w_in, float* d_delta, float* d_loss, int* d_info,
                         float* d_scratch, cudaStream_t stream) {
    const int F = s->num_features, cap = s->cap;
    if (F % 128 != 0 || F > 1024 || cap % kSlabRows != 0) {
        psb_set_error("tcgen05 worker kernel needs num_features %% 128 == 0, <= 1024 and max_buffer_size %% 64 == 0 "
                      "(got %d, %d): use the SIMT kernel for this shape", F, cap);
        return PSB_E_INVALID;
    }
    const int nkb = F / kBlockK;
    int NS = nkb + 4;                                           // one resident slab + 4 chunks of prefetch
    size_t fixed = (size_t)nkb * kBChunkBytes + kRBytes + (size_t)F * 4 + 32 * 4 + 34 * 8 + 64;
    while ((size_t)(NS + 2) * kChunkBytes + fixed + (size_t)(NS + 2) * 16 + 2048 <= 225 * 1024 && NS + 2 <= 2 * nkb)
        NS += 2;
    const size_t smem = 1024 + (size_t)NS * kChunkBytes + fixed + (size_t)NS * 16 + 64;
    CUtensorMap tm;
    int rc = psb_make_tmap_2d_f16(&tm, d_X, (uint64_t)num_workers * cap, (uint64_t)F, kSlabRows, kBlockK);
    if (rc) return rc;
    PSB_CHECK_CUDA(cudaFuncSetAttribute(psb_worker_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    psb_worker_tc_kernel<<<num_workers, kThreads, smem, stream>>>(tm, *s, cx, NS, d_labels, d_win, d_w_in, d_delta,
                                                                  d_loss, d_info, d_scratch);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
