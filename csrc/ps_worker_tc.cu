// K2, tensor flavour: the worker's local solve with every pass over the input buffer running on
// tcgen05 tensor cores fed by TMA, accumulators in TMEM.  One CTA per worker; two CTAs share an SM when there
// are more workers than SMs (4-slot rings), otherwise one CTA per SM with a 10-slot ring.
//
// One loss/gradient evaluation is one continuous TMA stream through a shared-memory ring of
// [128 rows x 64 features] fp16 chunks (16 KB, SWIZZLE_128B); a "slab" is 128 rows = F/64 chunks:
//   A(s)    : slab s, consumed K-major by GEMM1   Z_s[128 x 16] = slab x Wsp[F x 16]
//             (Wsp = fp16 hi | lo halves of the 8 weight columns, resident in shared memory)
//   epilogue: TMEM -> registers, + intercept, softmax over the R class rows, loss, residual r = p - y,
//             r -> fp16 hi | lo -> shared memory as a 16 x 128 MN-major B operand
//   B(s)    : slab s AGAIN (an L2 hit: A(s) is fetched with an evict-last policy, B(s) right behind it with
//             evict-first), the same bytes consumed MN-major by GEMM2
//             G[F x 16] += slab^T x Rsp_s   (M = 128 features per tile, K = 128 rows)
// The stream order is A(0) B(0) A(1) B(1) ...: while the epilogue of slab s runs, the ring keeps filling
// with B(s) and A(s+1); every chunk is released as soon as its MMAs retire, so the ring never drains
// and DRAM sees each row once per evaluation.  X^T R needs no atomics.  After the last slab the G accumulators
// (F/128 tiles x 16 TMEM columns) are drained, scaled by 1/(n std_f) and written to the solver's
// gradient vector.  The per-feature standard deviation (Spark standardisation) is computed by two
// otherwise idle warps straight from the shared-memory chunks of the FIRST evaluation (half of the chunks on
// their A use, half on their B use).  Consecutive passes visit the slabs in opposite directions (L2 reuse).
// Build-time instrumentation (off by default): -DPSB_TC_PROFILE=1|2|3 (cycles per role into info[4..7]),
// -DPSB_TC_TRACE (block-0 timeline, printf), -DPSB_TC_EXPERIMENT_NO_TMA / _NO_MMA / _NO_STATS_LDS (timing-only
// builds with wrong numerics); drivers in tools/ (mma_role_profile.py, trace_step.py, cta_span.py, phase_times.py).
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
constexpr int kSlabRows = 128;
constexpr int kBlockK = 64;                           // features per chunk (128-byte swizzle row)
constexpr int kChunkBytes = kSlabRows * kBlockK * 2;   // 16 KB
constexpr int kBChunkBytes = 16 * kBlockK * 2;         // 2 KB of Wsp per chunk
constexpr int kRBytes = 4096;                          // Rsp: 16 k-groups x (2 n-groups x 128 B)
constexpr int kTmemCols = 256;                         // two co-resident CTAs fit the SM's 512 columns
constexpr int kColZ = 0;                               // Z_0 at 0, Z_1 at 16
constexpr int kColG = 32;                              // G tile j at 32 + 16*j
// (Rotating consecutive MMAs over independent accumulators was measured: no gain — an SS-mode
//  128x16x16 MMA costs ~80 cycles of operand fetch regardless of accumulator dependencies.)

// Dynamic shared memory of the kernel.  The carve-up is kept as 32-bit offsets from this symbol and turned into
// pointers by the accessors below: pointers derived from a __shared__ symbol stay in the shared window for the
// compiler (LDS / STS / ATOMS with 32-bit addresses), where pointers parked in a struct degrade to generic
// 64-bit ones — and seventeen of those cost thirty-four registers of a kernel that runs at the 128-register cap.
extern __shared__ __align__(1024) unsigned char psb_tc_smem[];

struct TcShared {
    uint32_t o_ring, o_sB, o_sR, o_inv, o_sred, o_red, o_bar;   // offsets from psb_tc_smem
    int kNS;
    __device__ __forceinline__ unsigned char* ring() const { return psb_tc_smem + o_ring; }      // kNS x 16 KB, 1024-aligned
    __device__ __forceinline__ unsigned char* sB() const { return psb_tc_smem + o_sB; }          // nkb x 2 KB
    __device__ __forceinline__ unsigned char* sR() const { return psb_tc_smem + o_sR; }          // 4 KB Rsp + 4 KB statistics sums
    __device__ __forceinline__ float* inv() const { return (float*)(psb_tc_smem + o_inv); }      // [F]
    __device__ __forceinline__ float* sred() const { return (float*)(psb_tc_smem + o_sred); }    // [32]
    __device__ __forceinline__ double* red() const { return (double*)(psb_tc_smem + o_red); }    // [34]
    __device__ __forceinline__ uint64_t* bars() const { return (uint64_t*)(psb_tc_smem + o_bar); }
    __device__ __forceinline__ uint64_t* full() const { return bars(); }                          // [kNS]
    __device__ __forceinline__ uint64_t* empty() const { return bars() + kNS; }                   // [kNS]
    __device__ __forceinline__ uint64_t* z_full() const { return bars() + 2 * kNS; }              // [2]
    __device__ __forceinline__ uint64_t* r_full() const { return bars() + 2 * kNS + 2; }          // [2]
    __device__ __forceinline__ uint64_t* g_full() const { return bars() + 2 * kNS + 4; }
    __device__ __forceinline__ uint64_t* sdone() const { return bars() + 2 * kNS + 5; }           // [kNS] statistics warps done with the slot
    __device__ __forceinline__ uint64_t* inv_ready() const { return bars() + 3 * kNS + 5; }       // 1/std written (fused first evaluation)
    __device__ __forceinline__ uint32_t* tmem_slot() const { return (uint32_t*)(bars() + 3 * kNS + 6); }
};
// control words shared by all roles
__shared__ int psb_tc_fail;                    // a bounded wait gave up
__shared__ unsigned int psb_tc_chunk_base;     // ring position (chunks streamed so far)
__shared__ int psb_tc_need_redo;               // a constant non-zero column was found: redo the first evaluation

// `site` identifies the wait in the failure report (info[1] = -site when a worker's step gave up)
__device__ __forceinline__ bool wait_or_fail(uint64_t* bar, uint32_t parity, volatile int* fail, int site = 1) {
    if (tc::mbar_try_wait(bar, parity)) return true;            // common case: no clock read, no fail poll
    const unsigned long long t0 = clock64();
    while (!tc::mbar_try_wait(bar, parity)) {
        if (*fail) return false;
        if ((unsigned long long)(clock64() - t0) > 4000000000ull) {
            *fail = site;
            return false;
        }
    }
    return true;
}

// What one pass over the buffer does.  a: A stream + logits GEMM; b: B stream + gradient GEMM;
// fused: standardisation statistics ride on the A stream (first evaluation only);
// zstore: the epilogue also stores the raw logits (no intercept) of every valid row there;
// zsrc: residuals come from cached logits instead of TMEM (gradient-only pass).
struct PassCfg {
    bool a, b, fused;
    float* zstore;
    const float* zsrc;
};

struct TcEval {
    const CUtensorMap* tmX;
    const int* labels;        // this worker's [cap]
    int row_base;             // w * cap (row coordinate of slot 0 in the tensor map)
    int start, count, cap, F, R, kNS;      // kNS: ring slots (even: chunk pairs stay adjacent)
    uint32_t tmem;
    TcShared sh;
    unsigned int zuse[2], ruse[2], guse;   // per-thread running barrier use counts (persist across evaluations)
    int pass_no;                           // passes run so far (every thread counts them): parity = slab direction
    float* Zx;                             // cached logits of the current point / of the direction, [count][8] each,
    float* Zd;                             // indexed by window position (null: no caching, every trial is a full evaluation)
#ifdef PSB_TC_PROFILE
    long long prof[4];                     // cycles: [0] wait full(A) [1] wait r_full [2] wait full(B) [3] role total
#endif

    __device__ __forceinline__ int window_pos(int slot) const {      // position of a ring slot in the window (>= count: outside)
        int d = slot - start;
        if (d < 0) d += cap;
        return d;
    }
    __device__ __forceinline__ bool row_valid(int slot) const { return window_pos(slot) < count; }
    __device__ __forceinline__ bool slab_active(int b) const {
        const int lo = b * kSlabRows, hi = lo + kSlabRows;
        const int e = start + count;
        if (e <= cap) return lo < e && hi > start;
        return hi > start || lo < e - cap;
    }
    // true when the window cannot start or end strictly inside slab b (so first/last row valid => all valid)
    __device__ __forceinline__ bool slab_rows_contiguous(int b) const {
        const int lo = b * kSlabRows, hi = lo + kSlabRows;
        int e = start + count;
        if (e > cap) e -= cap;
        const bool start_inside = start > lo && start < hi;
        const bool end_inside = e > lo && e < hi;
        return count == cap || !(start_inside || end_inside);
    }
    __device__ int num_active_slabs() const {
        int n = 0;
        for (int b = 0; b < cap / kSlabRows; ++b) n += slab_active(b) ? 1 : 0;
        return n;
    }
    // slab index of the q-th slab this pass visits.  Passes alternate direction: the slabs a pass ends on are the
    // ones the next pass starts with, still in L2 (126 MB / 296 co-resident workers ~ 1.6 slabs each).
    __device__ int active_slab(int q) const {
        if (pass_no & 1) q = num_active_slabs() - 1 - q;
        for (int b = 0; b < cap / kSlabRows; ++b)
            if (slab_active(b) && q-- == 0) return b;
        return 0;
    }

    // ---- TMA producer (one thread).  with_b: stream every slab twice, A(s) then B(s).
    // fused: the statistics warps also follow every chunk (and read it once, on its A or its B use), so a slot is
    // refilled only after their sdone arrival for its previous occupant as well.
    __device__ void produce(const PassCfg& cfg) {
        const bool with_a = cfg.a, with_b = cfg.b, fused = cfg.fused;
        const int nkb = F / kBlockK, S = num_active_slabs();
        // ring position as running (slot, use) counters: this thread is alone on its warp, so the latency of a
        // runtime integer division per chunk (~100 cycles each) would pace the whole stream
        const unsigned int pc0 = psb_tc_chunk_base;
        int slot = (int)(pc0 % (unsigned)kNS);
        unsigned int use = pc0 / (unsigned)kNS;
        int idx = 0;                          // chunks issued in this pass
        unsigned int sd_par = 0u;             // per-slot parity of the statistics warps' sdone barrier
        auto load = [&](int b, int c, uint64_t policy) -> bool {
#if PSB_TC_PROFILE == 2
            const long long te = clock64();
#endif
            if (use > 0 && !wait_or_fail(sh.empty() + slot, (use - 1) & 1, &psb_tc_fail, 2)) return false;
#if PSB_TC_PROFILE == 2
            const long long ts = clock64();
            prof[0] += ts - te;
#endif
            if (fused && idx >= kNS) {
                // The statistics warp of this slot has SEEN the previous occupant (it arrives on sdone after every use
                // of its slots, whether it read the chunk or only followed the barrier): a hard handshake.  Waiting
                // only after the uses it reads left the others to a timing margin — a statistics warp held up for
                // more than one trip around the ring lost a phase of the slot's `full` barrier for good.
                if (!wait_or_fail(sh.sdone() + slot, (sd_par >> slot) & 1u, &psb_tc_fail, 3)) return false;
                sd_par ^= 1u << slot;
            }
#if PSB_TC_PROFILE == 2
            prof[1] += clock64() - ts;
#endif
#ifdef PSB_TC_EXPERIMENT_NO_TMA                // timing experiment only: no data movement, barriers still cycle
            tc::mbar_arrive(sh.full() + slot);
#else
            tc::mbar_arrive_expect_tx(sh.full() + slot, kChunkBytes);
            tc::tma_load_2d_hint(sh.ring() + (size_t)slot * kChunkBytes, tmX, c * kBlockK, row_base + b * kSlabRows,
                                 sh.full() + slot, policy);
#endif
            ++idx;
            if (++slot == kNS) slot = 0, ++use;
            return true;
        };
        // (L2 prefetch of the coming A chunks with cp.async.bulk.prefetch was measured twice: the whole next slab
        //  issued from this thread, 0.585 -> 0.796 ms per clock at 296 workers; eight chunks ahead of the MMA warp
        //  from an idle thread, 0.492 -> 0.492 — the A stream is not waiting on DRAM latency.)
        for (int q = 0; q < S; ++q) {
            const int b = active_slab(q);
            if (with_a)
                for (int c = 0; c < nkb; ++c)
                    if (!load(b, c, with_b ? tc::kL2EvictLast : tc::kL2EvictFirst)) return;
            if (with_b)
                for (int c = 0; c < nkb; ++c)
                    if (!load(b, c, q + 2 >= S ? tc::kL2EvictLast : tc::kL2EvictFirst)) return;   // tail slabs: next pass starts here
        }
    }

    // ---- statistics fused into the first evaluation: 2 warps (wi = 0, 1) read the chunks with c % 2 == wi
    //      while the MMA warp consumes them — half of them during the A stream, the other half during the B
    //      stream.  Each half-warp takes every other row, a lane reads 8 bytes (4 features) per row; per-chunk
    //      register sums are folded into shared per-feature accumulators (sum in the spare half of the Rsp
    //      region, sum of squares in inv[], converted in place at the end).  The warp waits on EVERY use of its
    //      slots, reading or not: every use of a ring slot has the same chunk parity, so a warp that follows
    //      all of them is never more than one mbarrier phase ahead.
    __device__ void stats_fused(int wi, int lane) const {
        const int nkb = F / kBlockK, S = num_active_slabs();
        // this warp's chunks are base + wi, base + wi + 2, ... through every A(q) B(q): one stride-2 walk of the ring
        const unsigned int pc0 = psb_tc_chunk_base + (unsigned)wi;
        int slot = (int)(pc0 % (unsigned)kNS);
        uint32_t phase = (pc0 / (unsigned)kNS) & 1u;
        float* ssum = (float*)(sh.sR() + kRBytes);
        float* ssq = sh.inv();
        const int h = lane >> 4, j8 = lane & 15;
#ifdef PSB_TC_TRACE
        long long st_wait = 0, st_read = 0;
        int st_chunks = 0;
#endif
        for (int q = 0; q < S; ++q) {
            const int b = active_slab(q);
            const bool full_slab = row_valid(b * kSlabRows) && row_valid(b * kSlabRows + kSlabRows - 1) &&
                                   slab_rows_contiguous(b);
            // half of this warp's chunks are read on their A use (c % 4 < 2), the other half on their B use — the
            // bytes are the same, and the reads no longer pace the A stream
            for (int pass = 0; pass < 2; ++pass) {
                for (int c = wi; c < nkb; c += 2) {
#ifdef PSB_TC_TRACE
                    const long long tw0 = clock64();
#endif
                    if (!wait_or_fail(sh.full() + slot, phase, &psb_tc_fail, 4)) return;
#ifdef PSB_TC_TRACE
                    const long long tw1 = clock64();
                    st_wait += tw1 - tw0;
#endif
                    if (((c >> 1) & 1) == pass) {
                        const uint32_t tile = tc::smem_u32(sh.ring()) + (uint32_t)slot * kChunkBytes + (j8 & 1) * 8;
                        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#ifdef PSB_TC_EXPERIMENT_NO_STATS_LDS          // timing experiment only (wrong statistics)
                        if (slot < 0)
#endif
                        // branch-free unrolled body (the row mask of a partial slab is a separate loop), so that the
                        // eight loads of a group are in flight together and their arithmetic interleaves
                        auto row = [&](const uint2 raw, const float m) {
                            float2 x = __half22float2(*(const __half2*)&raw.x), y = __half22float2(*(const __half2*)&raw.y);
                            x.x *= m, x.y *= m, y.x *= m, y.y *= m;
                            a0 += x.x, a1 += x.y, a2 += y.x, a3 += y.y;
                            q0 = fmaf(x.x, x.x, q0), q1 = fmaf(x.y, x.y, q1), q2 = fmaf(y.x, y.x, q2), q3 = fmaf(y.y, y.y, q3);
                        };
                        if (full_slab) {
#pragma unroll 1
                            for (int i0 = 0; i0 < kSlabRows / 2; i0 += 8) {
                                uint2 raw[8];
#pragma unroll
                                for (int u = 0; u < 8; ++u) raw[u] = tc::lds_v2(tile + tc::sw128_off(2 * (i0 + u) + h, j8 >> 1));
#pragma unroll
                                for (int u = 0; u < 8; ++u) row(raw[u], 1.f);
                            }
                        } else {
#pragma unroll 4
                            for (int i = 0; i < kSlabRows / 2; ++i) {
                                const int r = 2 * i + h;
                                row(tc::lds_v2(tile + tc::sw128_off(r, j8 >> 1)), row_valid(b * kSlabRows + r) ? 1.f : 0.f);
                            }
                        }
                        a0 += __shfl_xor_sync(0xffffffffu, a0, 16), a1 += __shfl_xor_sync(0xffffffffu, a1, 16);
                        a2 += __shfl_xor_sync(0xffffffffu, a2, 16), a3 += __shfl_xor_sync(0xffffffffu, a3, 16);
                        q0 += __shfl_xor_sync(0xffffffffu, q0, 16), q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
                        q2 += __shfl_xor_sync(0xffffffffu, q2, 16), q3 += __shfl_xor_sync(0xffffffffu, q3, 16);
                        if (h == 0) {
                            const uint32_t ps = tc::smem_u32(ssum + c * kBlockK + 4 * j8);
                            const uint32_t pq = tc::smem_u32(ssq + c * kBlockK + 4 * j8);
                            const float4 u = tc::lds_f4(ps), v = tc::lds_f4(pq);
                            tc::sts_f4(ps, make_float4(u.x + a0, u.y + a1, u.z + a2, u.w + a3));
                            tc::sts_f4(pq, make_float4(v.x + q0, v.y + q1, v.z + q2, v.w + q3));
                        }
#ifdef PSB_TC_TRACE
                        st_read += clock64() - tw1, ++st_chunks;
#endif
                    }
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(sh.sdone() + slot);   // seen (and, on its read use, done with) the slot
                    slot += 2;
                    if (slot >= kNS) slot -= kNS, phase ^= 1u;
                }
            }
        }
        __syncwarp();
#ifdef PSB_TC_TRACE
        if (blockIdx.x == 0 && lane == 0)
            printf("stats warp %d: %d chunks read, %lld cycles reading (%lld per chunk), %lld cycles waiting\n", wi, st_chunks,
                   st_read, st_read / (st_chunks ? st_chunks : 1), st_wait);
#endif
        const float n = (float)count;
        bool redo = false;
        for (int c = wi; c < nkb; c += 2) {
            const int f = c * kBlockK + 2 * lane;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const float s1 = ssum[f + e], s2 = ssq[f + e];
                const float var = count > 1 ? (s2 - s1 * s1 / n) / (n - 1.f) : 0.f;
                const float iv = var > 1e-12f ? rsqrtf(var) : 0.f;
                ssq[f + e] = iv;                               // inv[] in place
                // a constant NON-zero column gets weight 0 under standardisation, but this pass used w0 for it
                redo = redo || (iv == 0.f && s2 != 0.f);
            }
        }
        if (redo) psb_tc_need_redo = 1;
        __threadfence_block();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(sh.inv_ready());
    }

    // ---- MMA issuer: the whole warp walks the ring in stream order (warp-uniform state), one elected
    //      lane issues the tcgen05 instructions
    __device__ void mma_loop(const PassCfg& cfg) {
        const int nkb = F / kBlockK, S = __shfl_sync(0xffffffffu, num_active_slabs(), 0);
        constexpr uint32_t idesc1 = tc::idesc_f16(128, 16, 0, 0);    // GEMM1: A K-major, B K-major
        constexpr uint32_t idesc2 = tc::idesc_f16(128, 16, 1, 1);    // GEMM2: A MN-major, B MN-major
        // ring position as running (slot, phase) counters — no per-chunk integer divisions on the issue path
        const unsigned int pc0 = __shfl_sync(0xffffffffu, psb_tc_chunk_base, 0);
        int slot = (int)(pc0 % (unsigned)kNS);
        uint32_t phase = (pc0 / (unsigned)kNS) & 1u;
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
        const uint32_t sB0 = tc::smem_u32(sh.sB()), sR0 = tc::smem_u32(sh.sR()), ring0 = tc::smem_u32(sh.ring());

        auto gemm1 = [&](int q) -> bool {                       // Z_{q&1} = A(q) x Wsp
            for (int c = 0; c < nkb; ++c) {
#ifdef PSB_TC_PROFILE
                const long long tw = clock64();
#endif
                if (!wait_or_fail(sh.full() + slot, phase, &psb_tc_fail, 5)) return false;
#ifdef PSB_TC_PROFILE
                prof[0] += clock64() - tw;
#endif
                tc::tc_fence_after();
                const uint64_t ad = tc::desc_k_sw128(ring0 + slot * kChunkBytes, 1024);
                const uint64_t bd = tc::desc_k_sw128(sB0 + c * kBChunkBytes, 1024);
                if (tc::elect_one()) {
#ifndef PSB_TC_EXPERIMENT_NO_MMA               // timing experiment only: commits without MMAs
#pragma unroll
                    for (int k = 0; k < kBlockK / 16; ++k)
                        tc::umma_f16(tmem_u + kColZ + 16 * (q & 1), ad + 2 * k, bd + 2 * k, idesc1, (c | k) ? 1u : 0u);
#endif
                    tc::umma_commit(sh.empty() + slot);
                    if (c == nkb - 1) tc::umma_commit(sh.z_full() + (q & 1));
                }
                __syncwarp();
                if (++slot == kNS) slot = 0, phase ^= 1u;
            }
            return true;
        };
        auto gemm2 = [&](int q) -> bool {                       // G += B(q)^T x Rsp_{q&1}
#ifdef PSB_TC_PROFILE
            const long long tr = clock64();
#endif
            if (!wait_or_fail(sh.r_full(), ruse[0] & 1, &psb_tc_fail, 6)) return false;
            ++ruse[0];
#ifdef PSB_TC_PROFILE
            prof[1] += clock64() - tr;
#endif
            tc::tc_fence_after();
            for (int j = 0; j < nkb / 2; ++j) {                 // chunks 2j, 2j+1 sit in adjacent slots (kNS even)
#ifdef PSB_TC_PROFILE
                const long long tb = clock64();
#endif
                if (!wait_or_fail(sh.full() + slot, phase, &psb_tc_fail, 7)) return false;
                if (!wait_or_fail(sh.full() + slot + 1, phase, &psb_tc_fail, 8)) return false;
#ifdef PSB_TC_PROFILE
                prof[2] += clock64() - tb;
#endif
                tc::tc_fence_after();
                const uint64_t ad0 = tc::desc_mn_sw128(ring0 + slot * kChunkBytes, kChunkBytes, 1024);
                const uint64_t bd0 = tc::desc_mn_none(sR0, 256, 128);   // one Rsp buffer: epilogue(q+1) starts after GEMM2(q)
                if (tc::elect_one()) {
#ifndef PSB_TC_EXPERIMENT_NO_MMA
#pragma unroll
                    for (int ks = 0; ks < kSlabRows / 16; ++ks)      // +2048 B of A, +512 B of B per 16 rows
                        tc::umma_f16(tmem_u + kColG + 16 * j, ad0 + (uint64_t)(ks * 128), bd0 + (uint64_t)(ks * 32),
                                     idesc2, (q | ks) ? 1u : 0u);
#endif
                    tc::umma_commit(sh.empty() + slot);
                    tc::umma_commit(sh.empty() + slot + 1);
                }
                __syncwarp();
                slot += 2;
                if (slot == kNS) slot = 0, phase ^= 1u;
            }
            return true;
        };
        for (int q = 0; q < S; ++q) {
            if (cfg.a && !gemm1(q)) return;
            if (cfg.b && !gemm2(q)) return;
            if (cfg.b && !cfg.a) {           // gradient-only pass: tell the residual warps that Rsp(q) has been consumed
                if (tc::elect_one()) tc::umma_commit(sh.z_full() + (q & 1));
                __syncwarp();
            }
        }
        if (cfg.b && tc::elect_one()) tc::umma_commit(sh.g_full());
        __syncwarp();
    }

    // ---- epilogue warps (4 warps; wi = TMEM sub-partition = warp % 4); thread = row of the slab
    __device__ void epilogue(int wi, int lane, float* g, float invn, const PassCfg& cfg) {
        const bool fused = cfg.fused;
        const int nkb = F / kBlockK, S = num_active_slabs();
        float lsum = 0.f, rs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float bias[8];                               // intercepts staged behind the reduction words
#pragma unroll
        for (int c = 0; c < 8; ++c) bias[c] = sh.sred()[16 + c];
        const int r = wi * 32 + lane;
        // the label of the NEXT slab's row is requested before waiting for this slab's logits, so that its latency
        // is not part of the z_full -> r_full path the MMA warp sits on
        int lab_next = 0;
        if (S > 0) {
            const int s0 = active_slab(0) * kSlabRows + r;
            lab_next = row_valid(s0) ? labels[s0] : 0;
        }
        for (int q = 0; q < S; ++q) {
            const int b = active_slab(q), zb = q & 1;
            const int slot = b * kSlabRows + r;
            const bool valid = row_valid(slot);
            const int lab = lab_next;
            if (q + 1 < S) {
                const int s1 = active_slab(q + 1) * kSlabRows + r;
                lab_next = row_valid(s1) ? labels[s1] : 0;
            }
            float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (cfg.a) {
                if (!wait_or_fail(sh.z_full() + zb, zuse[zb] & 1, &psb_tc_fail, 9)) return;
                ++zuse[zb];
                tc::tc_fence_after();
                float v[16];
                tc::tmem_ld_32x32b_x16(tmem + kColZ + 16 * zb + ((uint32_t)(wi * 32) << 16), v);
#pragma unroll
                for (int c = 0; c < 8; ++c) z[c] = v[c] + v[8 + c];
                if (cfg.zstore && valid) {
                    float4* zo = (float4*)cfg.zstore + 2 * window_pos(slot);
                    zo[0] = make_float4(z[0], z[1], z[2], z[3]);
                    zo[1] = make_float4(z[4], z[5], z[6], z[7]);
                }
            } else {
                // gradient-only pass: nothing paces these warps but the single Rsp buffer — wait until GEMM2(q-1)
                // has consumed it (the MMA warp commits z_full[(q-1)&1] behind that GEMM)
                if (q >= 1) {
                    if (!wait_or_fail(sh.z_full() + ((q - 1) & 1), zuse[(q - 1) & 1] & 1, &psb_tc_fail, 10)) return;
                    ++zuse[(q - 1) & 1];
                }
            }
            if (!cfg.a && valid) {
                const float4* zi = (const float4*)cfg.zsrc + 2 * window_pos(slot);
                const float4 z0 = zi[0], z1 = zi[1];
                z[0] = z0.x, z[1] = z0.y, z[2] = z0.z, z[3] = z0.w, z[4] = z1.x, z[5] = z1.y, z[6] = z1.z, z[7] = z1.w;
            }
            float rr[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (valid) {
                lsum += softmax_residual(z, bias, lab, R, rr);
#pragma unroll
                for (int c = 0; c < 8; ++c) rs[c] += rr[c];
            }
            // Rsp as MN-major no-swizzle B: core matrix (k-group kg, n-group ng) at kg*256 + ng*128,
            // row k%8 at +16 bytes each; n-group 0 = fp16 high parts, n-group 1 = low parts
            __align__(16) __half hi[8], lo[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) tc::split_half(rr[c], hi[c], lo[c]);
            const uint32_t dst = tc::smem_u32(sh.sR()) + (r >> 3) * 256 + (r & 7) * 16;
            tc::sts_v4(dst, *(const uint4*)hi);
            tc::sts_v4(dst + 128, *(const uint4*)lo);
            tc::fence_proxy_async_smem();
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(sh.r_full());
        }
        if (!cfg.a && S >= 1) {                  // consume the last "Rsp consumed" commit of a gradient-only pass
            if (!wait_or_fail(sh.z_full() + ((S - 1) & 1), zuse[(S - 1) & 1] & 1, &psb_tc_fail, 11)) return;
            ++zuse[(S - 1) & 1];
        }
        for (int o = 16; o; o >>= 1) {
            lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
#pragma unroll
            for (int c = 0; c < 8; ++c) rs[c] += __shfl_xor_sync(0xffffffffu, rs[c], o);
        }
        if (lane == 0) {
            atomicAdd(&sh.sred()[0], lsum);
#pragma unroll
            for (int c = 0; c < 8; ++c) atomicAdd(&sh.sred()[1 + c], rs[c]);
        }
        // drain the gradient accumulators: thread = feature row of each 128-feature tile
        if (fused && !wait_or_fail(sh.inv_ready(), 0, &psb_tc_fail, 12)) return;
        if (!wait_or_fail(sh.g_full(), guse & 1, &psb_tc_fail, 13)) return;
        ++guse;
        tc::tc_fence_after();
        for (int j = 0; j < nkb / 2; ++j) {
            float v[16];
            tc::tmem_ld_32x32b_x16(tmem + kColG + 16 * j + ((uint32_t)(wi * 32) << 16), v);
            const int f = j * 128 + wi * 32 + lane;
            const float sc = sh.inv()[f] * invn;
            float4* g4 = (float4*)g + 2 * f;
            g4[0] = make_float4((v[0] + v[8]) * sc, (v[1] + v[9]) * sc, (v[2] + v[10]) * sc, (v[3] + v[11]) * sc);
            g4[1] = make_float4((v[4] + v[12]) * sc, (v[5] + v[13]) * sc, (v[6] + v[14]) * sc, (v[7] + v[15]) * sc);
        }
        tc::tc_fence_before();
    }

    // ---- no standardisation: 1/std = 1 (all threads enter)
    __device__ void unit_scale() {
        for (int f = threadIdx.x; f < F; f += blockDim.x) sh.inv()[f] = 1.f;
        __syncthreads();
    }

    // ---- one configured pass over the buffer; all threads enter, all get the same loss.
    // x: the point whose weights feed the logits GEMM (when cfg.a) and whose intercepts enter the softmax.
    // In the fused pass x holds the UNSCALED start weights w0 (1/std is being computed during the pass).
    __device__ double run_pass(const float* x, float* g, const PassCfg& cfg) {
        const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#if PSB_TC_PROFILE == 3                        // thread 0: [0] weight staging [1] role section [2] pass tail
        const long long tq0 = clock64();
#endif
        PSB_TRACE(20);
        if (cfg.a) {
            // Wsp = (x * inv) split into fp16 hi | lo, K-major SWIZZLE_128B, 16 rows per 64-feature chunk
            vec_prefetch(x, F * 2);
            for (int f = tid; f < F; f += blockDim.x) {
                const float4 a = ((const float4*)x)[2 * f], b = ((const float4*)x)[2 * f + 1];
                const float sc = cfg.fused ? 1.f : sh.inv()[f];
                const float w[8] = {a.x * sc, a.y * sc, a.z * sc, a.w * sc, b.x * sc, b.y * sc, b.z * sc, b.w * sc};
                const int kb = f >> 6, k = f & 63, j = k >> 3, e = k & 7;
                const uint32_t chunk = tc::smem_u32(sh.sB()) + (uint32_t)kb * kBChunkBytes + e * 2;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    __half hi, lo;
                    tc::split_half(w[c], hi, lo);
                    tc::sts_u16(chunk + tc::sw128_off(c, j), __half_as_ushort(hi));
                    tc::sts_u16(chunk + 1024 + tc::sw128_off(c, j), __half_as_ushort(lo));
                }
            }
        }
        if (tid < 16) sh.sred()[tid] = 0.f;
        if (tid < 8) sh.sred()[16 + tid] = x[F * 8 + tid];          // intercepts (unscaled)
        if (cfg.fused)
            for (int f = tid; f < F; f += blockDim.x) sh.inv()[f] = 0.f, ((float*)(sh.sR() + kRBytes))[f] = 0.f;
        tc::fence_proxy_async_smem();
        __syncthreads();
        const float invn = 1.f / (float)count;
        PSB_TRACE(21);
#if PSB_TC_PROFILE == 3
        const long long tq1 = clock64();
        prof[0] += tq1 - tq0;
#endif
        if (warp == 0) {
#if PSB_TC_PROFILE == 2
            const long long tp0 = clock64();
#endif
            if (lane == 0) produce(cfg);
#if PSB_TC_PROFILE == 2
            prof[3] += clock64() - tp0;
#endif
            __syncwarp();
        } else if (warp == 1) {
#if PSB_TC_PROFILE == 1
            const long long tm0 = clock64();
#endif
            mma_loop(cfg);
#if PSB_TC_PROFILE == 1
            prof[3] += clock64() - tm0;
#endif
            __syncwarp();
        } else if (warp < 6) {
            epilogue(warp & 3, lane, g, invn, cfg);
        } else if (cfg.fused) {
            stats_fused(warp - 6, lane);
        }
        __syncthreads();
        PSB_TRACE(22);
#if PSB_TC_PROFILE == 3
        const long long tq2 = clock64();
        prof[1] += tq2 - tq1;
#endif
        if (tid == 0)
            psb_tc_chunk_base += (unsigned)((cfg.a ? 1 : 0) + (cfg.b ? 1 : 0)) * (unsigned)num_active_slabs() * (F / kBlockK);
        if (cfg.b && tid < 8) g[F * 8 + tid] = sh.sred()[1 + tid] * invn;
        const double loss = (double)sh.sred()[0] * (double)invn;
        ++pass_no;
        __syncthreads();
#if PSB_TC_PROFILE == 3
        prof[2] += clock64() - tq2;
        prof[3] += 1;
#endif
        return loss;
    }

    // classic interface: loss + gradient at x (scaled space), one A+B pass
    __device__ double eval(const float* x, float* g) {
        const PassCfg cfg = {true, true, false, Zx, nullptr};   // keeps the logit cache current when there is one
        return run_pass(x, g, cfg);
    }
    __device__ double eval_store(const float* x, float* g, float* zbuf) {     // A+B pass, logits parked in zbuf
        const PassCfg cfg = {true, true, false, zbuf, nullptr};
        return run_pass(x, g, cfg);
    }
    __device__ double eval_first_fused(const float* w0, float* g) {
        const PassCfg cfg = {true, true, true, Zx, nullptr};
        return run_pass(w0, g, cfg);
    }

    // ---- cached-logit interface used by local_solve_cached (solver.cuh)
    // the Zd buffer holds the logits of x + t1 d: turn it into z(d) = (Zt - Zx) / t1 (logits are linear)
    __device__ void derive_direction_logits(double t1) {
        const float it = (float)(1.0 / t1);
        for (int i = threadIdx.x; i < count; i += blockDim.x) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float4 a = ((const float4*)Zx)[2 * i + h];
                float4 b = ((float4*)Zd)[2 * i + h];
                b.x = (b.x - a.x) * it, b.y = (b.y - a.y) * it, b.z = (b.z - a.z) * it, b.w = (b.w - a.w) * it;
                ((float4*)Zd)[2 * i + h] = b;
            }
        }
        __syncthreads();
    }
    __device__ void gradient_at(const float* x, float* g) {        // residuals from Zx, G = X^T R: B stream only
        const PassCfg cfg = {false, true, false, nullptr, Zx};
        run_pass(x, g, cfg);
    }
    // phi(t) and phi'(t) of the line search from the cached logits: z = Zx + t Zd (+ intercepts of x + t d)
    __device__ void trial(double t, const float* x, const float* d, double& f, double& dd) {
        const float ft = (float)t;
        float bias[8], bd[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) bd[c] = d[F * 8 + c], bias[c] = fmaf(ft, bd[c], x[F * 8 + c]);
        float lsum = 0.f, dsum = 0.f;
        for (int i = threadIdx.x; i < count; i += blockDim.x) {
            int slot = start + i;
            if (slot >= cap) slot -= cap;
            const float4 a0 = ((const float4*)Zx)[2 * i], a1 = ((const float4*)Zx)[2 * i + 1];
            const float4 d0 = ((const float4*)Zd)[2 * i], d1 = ((const float4*)Zd)[2 * i + 1];
            const float zd[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
            const float z[8] = {fmaf(ft, zd[0], a0.x), fmaf(ft, zd[1], a0.y), fmaf(ft, zd[2], a0.z), fmaf(ft, zd[3], a0.w),
                                fmaf(ft, zd[4], a1.x), fmaf(ft, zd[5], a1.y), fmaf(ft, zd[6], a1.z), fmaf(ft, zd[7], a1.w)};
            float rr[8];
            lsum += softmax_residual(z, bias, labels[slot], R, rr);
#pragma unroll
            for (int c = 0; c < 8; ++c) dsum = fmaf(rr[c], zd[c] + bd[c], dsum);
        }
        const double n = (double)count;
        f = block_sum((double)lsum, sh.red()) / n;
        dd = block_sum((double)dsum, sh.red()) / n;
#ifdef PSB_TC_DEBUG
        if (blockIdx.x == 0 && threadIdx.x == 0) printf("trial t=%g f=%g dd=%g\n", t, f, dd);
#endif
    }
    __device__ void advance(double alpha) {                        // Zx <- Zx + alpha Zd
        const float fa = (float)alpha;
        for (int i = threadIdx.x; i < count; i += blockDim.x) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float4 a = ((float4*)Zx)[2 * i + h];
                const float4 b = ((const float4*)Zd)[2 * i + h];
                a.x = fmaf(fa, b.x, a.x), a.y = fmaf(fa, b.y, a.y), a.z = fmaf(fa, b.z, a.z), a.w = fmaf(fa, b.w, a.w);
                ((float4*)Zx)[2 * i + h] = a;
            }
        }
    }
};
}  // namespace

__global__ void __launch_bounds__(kThreads, 2)
psb_worker_tc_kernel(const __grid_constant__ CUtensorMap tmX, PsbSolver sv, CommCtx cx, int kNS,
                     const int* __restrict__ labels, const int* __restrict__ win, const float* __restrict__ w_in,
                     float* __restrict__ delta_out, float* __restrict__ loss_out, int* __restrict__ info_out,
                     float* __restrict__ scratch) {
    const int F = sv.num_features, R = sv.num_class_rows, cap = sv.cap;
    const int nv = (F + 1) * 8, n4 = nv / 4, nkb = F / kBlockK;
    const int w = blockIdx.x, tid = threadIdx.x, warp = tid >> 5;

    TcShared sh;
    sh.kNS = kNS;
    sh.o_ring = (1024u - (tc::smem_u32(psb_tc_smem) & 1023u)) & 1023u;      // the ring wants 1024-byte alignment
    sh.o_sB = sh.o_ring + (uint32_t)kNS * kChunkBytes;
    sh.o_sR = sh.o_sB + (uint32_t)nkb * kBChunkBytes;
    sh.o_inv = sh.o_sR + 2 * kRBytes;
    sh.o_sred = sh.o_inv + (uint32_t)F * 4;
    sh.o_red = sh.o_sred + 32 * 4;
    sh.o_bar = sh.o_red + 34 * 8;

#ifdef PSB_TC_TRACE
    unsigned long long gt_start = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_start));
    if (tid == 0) psb_trace_n = 0;
    __syncthreads();
    PSB_TRACE(0);
#endif
    // Every clock starts from fresh barriers and ring counters (the roles' use counts restart with the clock), so a
    // launch that runs several clocks per worker (PSB_COMM_WORKER) re-initialises them between two clocks.
    auto init_barriers = [&](bool again) {
        if (again)
            for (int i = 0; i < 3 * kNS + 6; ++i) tc::mbar_inval(sh.bars() + i);
        for (int s = 0; s < kNS; ++s) tc::mbar_init(sh.full() + s, 1), tc::mbar_init(sh.empty() + s, 1);
        for (int i = 0; i < 2; ++i) tc::mbar_init(sh.z_full() + i, 1), tc::mbar_init(sh.r_full() + i, 4);
        tc::mbar_init(sh.g_full(), 1);
        for (int s = 0; s < kNS; ++s) tc::mbar_init(sh.sdone() + s, 1);
        tc::mbar_init(sh.inv_ready(), 2);
        tc::mbar_fence_init();
    };
    if (tid == 0) {
        psb_tc_fail = 0, psb_tc_chunk_base = 0, psb_tc_need_redo = 0;
        tc::tma_prefetch_desc(&tmX);
        init_barriers(false);
    }
    if (warp == 0) tc::tmem_alloc(sh.tmem_slot(), kTmemCols);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *sh.tmem_slot();

    WorkerScratch ws(scratch + (size_t)w * PSB_SCRATCH_VECS * nv, nv);
    const int head = win[2 * w], count = win[2 * w + 1];
    int start = head - count;
    if (start < 0) start += cap;
    const int* lw = labels + (size_t)w * cap;
    const int gw = cx.c.worker_base + w;
    const int nclk = (cx.c.mode == PSB_COMM_WORKER || cx.c.mode == PSB_COMM_SEQUENTIAL) && cx.c.num_clocks > 1 ? cx.c.num_clocks : 1;

  for (int it = 0; it < nclk; ++it) {
    const int clock = cx.c.clock + it;
    PsbClockRecord* rec = cx.c.mode == PSB_COMM_WORKER && cx.c.hist
                              ? (PsbClockRecord*)cx.c.hist + (size_t)it * gridDim.x + w : nullptr;
    if (it > 0) {
        tc::tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            psb_tc_chunk_base = 0, psb_tc_need_redo = 0;
            init_barriers(true);
        }
        __syncthreads();
        tc::tc_fence_after();
    }
    // ---- pull
    PSB_TRACE(1);
    const long long t_begin = clock64();
    float* w0 = ws.w0;
    bool alive = true;
    if (cx.c.mode == PSB_COMM_NONE) {
        for (int i = tid; i < n4; i += blockDim.x) ((float4*)w0)[i] = ((const float4*)w_in)[i];
        __syncthreads();
    } else {
        alive = comm_pull(cx, clock, gw, (float4*)w0, rec);
        if (!alive) break;                                        // an in-kernel wait gave up: the CTA stops
    }
    if (alive && count <= 0) {
        if (tid == 0) {
            loss_out[w] = nanf("");
            info_out[8 * w + 2] = -1, info_out[8 * w + 3] = 0;
            if (rec) rec->loss = nanf("");
        }
        if (cx.c.mode != PSB_COMM_NONE) comm_push(cx, clock, gw, nullptr);   // still takes part in the protocol
        alive = false;
    }
    if (alive) {
        TcEval ev;
        ev.tmX = &tmX, ev.labels = lw, ev.row_base = w * cap;
        ev.start = start, ev.count = count, ev.cap = cap, ev.F = F, ev.R = R, ev.kNS = kNS;
        ev.tmem = tmem, ev.sh = sh;
        ev.zuse[0] = ev.zuse[1] = ev.ruse[0] = ev.ruse[1] = ev.guse = 0;
        ev.pass_no = 0;
        // logit cache in two spare solver vectors (the L-BFGS pairs of index PSB_HIST-1 are never reached when
        // max_iter <= PSB_HIST-1): [count][8] floats each (indexed by window position) must fit one parameter vector
        const bool cached = sv.solver == PSB_SOLVER_LBFGS && count * 8 <= nv && sv.max_iter <= PSB_HIST - 1;
        ev.Zx = cached ? ws.vec[7 + PSB_HIST - 1] : nullptr;
        ev.Zd = cached ? ws.vec[7 + 2 * PSB_HIST - 1] : nullptr;
#ifdef PSB_TC_PROFILE
        ev.prof[0] = ev.prof[1] = ev.prof[2] = ev.prof[3] = 0;
#endif

        SolveVecs v;
        ws.bind(v);
        v.n4 = n4;
        // First evaluation with the standardisation statistics fused in (one pass over the buffer
        // instead of two): at the start point the effective weights are w0 itself.
        bool have_f0 = false;
        double f0 = 0.0;
        PSB_TRACE(2);
        if (sv.standardize && sv.inv_std) {
            // statistics maintained at ingestion time (psb_window_stats_update): no statistics pass, the first
            // evaluation is an ordinary one at the scaled start point
            for (int f = tid; f < F; f += blockDim.x) sh.inv()[f] = sv.inv_std[(size_t)w * F + f];
            __syncthreads();
        } else if (sv.standardize) {
            f0 = ev.eval_first_fused(w0, v.g);
            have_f0 = !psb_tc_need_redo;
        } else {
            ev.unit_scale();
        }
        const long long t_stats = clock64();
        PSB_TRACE(3);
        to_scaled_space(v.x, w0, sh.inv(), F);
        __syncthreads();
        PSB_TRACE(4);

        SolveResult res;
        if (cached) {
            if (!have_f0) f0 = ev.eval(v.x, v.g);                    // (also fills the logit cache)
            res = local_solve_cached(ev, v, sv.max_iter, sh.red(), f0);
        } else {
            res = local_solve(ev, v, sv.solver, sv.max_iter, sv.sgd_lr, sh.red(), have_f0, f0);
        }
        const long long t_solve = clock64();
        PSB_TRACE(5);

        float* delta = v.tmp;
        finish_delta(delta, v.x, w0, sh.inv(), F, R, sv.center && sv.solver == PSB_SOLVER_LBFGS);
        __syncthreads();
        PSB_TRACE(6);
        if (tid == 0) {
            loss_out[w] = psb_tc_fail ? nanf("") : (float)res.loss;
            if (rec) rec->loss = loss_out[w];
            info_out[8 * w + 0] = res.evals, info_out[8 * w + 1] = psb_tc_fail ? -psb_tc_fail : res.iterations;
            info_out[8 * w + 2] = psb_tc_fail ? -2 : res.status, info_out[8 * w + 3] = count;
            info_out[8 * w + 4] = (int)((t_stats - t_begin) >> 6), info_out[8 * w + 5] = (int)((t_solve - t_stats) >> 6);
        }
        if (cx.c.mode == PSB_COMM_NONE) {
            float4* out = (float4*)(delta_out + (size_t)w * nv);
            for (int i = tid; i < n4; i += blockDim.x) out[i] = ((const float4*)delta)[i];
        } else {
            // a step whose barrier wait gave up has no result: it still takes part in the protocol, with no delta
            comm_push(cx, clock, gw, psb_tc_fail ? nullptr : (const float4*)delta);
        }
        if (tid == 0) info_out[8 * w + 6] = (int)((clock64() - t_solve) >> 6), info_out[8 * w + 7] = (int)((clock64() - t_begin) >> 6);
        PSB_TRACE(7);
#ifdef PSB_TC_TRACE
        if (tid == 0 && blockIdx.x == 0)
            for (int i = 1; i < psb_trace_n; ++i)
                printf("trace %2d -> %2d : %7lld cycles\n", psb_trace_id[i - 1], psb_trace_id[i], psb_trace_t[i] - psb_trace_t[i - 1]);
#endif
#ifdef PSB_TC_PROFILE
        __syncthreads();
        if (tid == (PSB_TC_PROFILE >= 2 ? 0 : 32)) for (int k = 0; k < 4; ++k) info_out[8 * w + 4 + k] = (int)(ev.prof[k] >> 6);
#endif
    }
    if (psb_tc_fail) break;                                       // (uniform: written before the last barrier of the clock)
  }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, kTmemCols);
#ifdef PSB_TC_TRACE                            // wall-clock (ns, low 31 bits) of this CTA's first and last instruction
    if (tid == 0) {
        unsigned long long gt_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_end));
        info_out[8 * w + 4] = (int)(gt_start & 0x7fffffffull), info_out[8 * w + 5] = (int)(gt_end & 0x7fffffffull);
    }
#endif
}

// Ring depth and dynamic shared memory for a launch of `num_workers` CTAs on a device with `num_sms` SMs.
// More workers than SMs: 4-slot rings so that two CTAs share an SM and one worker's vector / epilogue phases
// overlap the other's streaming; otherwise one CTA per SM with a deep ring.
static void tc_launch_shape(int F, int num_workers, int num_sms, int* kNS_out, size_t* smem_out) {
    const int nkb = F / kBlockK;
    auto smem_for = [&](int ns) {
        return 1024 + (size_t)ns * kChunkBytes + (size_t)nkb * kBChunkBytes + 2 * kRBytes + (size_t)F * 4 + 32 * 4 +
               34 * 8 + (size_t)(3 * ns + 6) * 8 + 64;
    };
    int kNS = 10;
    while (kNS > 2 && smem_for(kNS) > 225 * 1024) kNS -= 2;
    if (num_workers > num_sms) {
        kNS = 4;
        while (kNS > 2 && 2 * (smem_for(kNS) + 1024) > 227 * 1024) kNS -= 2;
    }
    *kNS_out = kNS, *smem_out = smem_for(kNS);
}

// worker CTAs that can be resident at the same time when `num_workers` of them are launched together
int psb_worker_tc_resident(const PsbSolver* s, int num_workers) {
    int dev = 0, sms = 0, per_sm = 0, kNS = 0;
    size_t smem = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    tc_launch_shape(s->num_features, num_workers, sms, &kNS, &smem);
    // CTAs per SM from the resources themselves: registers, shared memory (1 KB reserved per CTA), threads, and
    // tensor memory (kTmemCols of the SM's 512 columns per CTA — not known to the occupancy calculator, which also
    // answers 1 for this kernel on the 580 driver although two CTAs per SM do run together)
    cudaFuncAttributes fa;
    int regs_sm = 65536, smem_sm = 233472, thr_sm = 2048;
    if (cudaFuncGetAttributes(&fa, psb_worker_tc_kernel) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&regs_sm, cudaDevAttrMaxRegistersPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&thr_sm, cudaDevAttrMaxThreadsPerMultiProcessor, dev);
    const int by_regs = regs_sm / (((fa.numRegs + 7) / 8 * 8) * kThreads);
    const int by_smem = (int)(smem_sm / (smem + fa.sharedSizeBytes + 1024));
    per_sm = by_regs < by_smem ? by_regs : by_smem;
    if (per_sm > thr_sm / kThreads) per_sm = thr_sm / kThreads;
    if (per_sm > 512 / kTmemCols) per_sm = 512 / kTmemCols;
    return sms * per_sm;
}

int psb_launch_worker_tc(const PsbSolver* s, const CommCtx& cx, int num_workers, const void* d_X, const int* d_labels,
                         const int* d_win, const float* d_w_in, float* d_delta, float* d_loss, int* d_info,
                         float* d_scratch, cudaStream_t stream) {
    const int F = s->num_features, cap = s->cap;
    if (F % 128 != 0 || F > 1024 || cap % kSlabRows != 0) {
        psb_set_error("tcgen05 worker kernel needs num_features %% 128 == 0, <= 1024 and max_buffer_size %% 128 == 0 "
                      "(got %d, %d): use the SIMT kernel for this shape", F, cap);
        return PSB_E_INVALID;
    }
    int num_sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    int kNS = 0;
    size_t smem = 0;
    tc_launch_shape(F, num_workers, num_sms, &kNS, &smem);
    CUtensorMap tm;
    int rc = psb_make_tmap_2d_f16(&tm, d_X, (uint64_t)num_workers * cap, (uint64_t)F, kSlabRows, kBlockK);
    if (rc) return rc;
    PSB_CHECK_CUDA(cudaFuncSetAttribute(psb_worker_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    psb_worker_tc_kernel<<<num_workers, kThreads, smem, stream>>>(tm, *s, cx, kNS, d_labels, d_win, d_w_in, d_delta, d_loss,
                                                                  d_info, d_scratch);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}
