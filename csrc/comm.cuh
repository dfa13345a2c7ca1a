// Fused pull / push / apply over NVLink peer memory (device side).
//
// Replaces: WEIGHTS_TOPIC broadcast (ServerProcessor.java:172-182), GRADIENTS_TOPIC push
// (WorkerTrainingProcessor.java:95-97), the server's w += lr*g loop (ServerProcessor.java:148-151,
// 225-228) and the MessageTracker gating (MessageTracker.java:69-87).  There is no server process:
// the parameter vector is cut into num_ranks shards; rank s owns shard s.
//
// Sequential (BSP) protocol for clock v, parity p = v & 1:
//   pull   : wait until version[s] >= v for every shard s (local acquire spins), read replica[p].
//   push   : workers of a GPU reduce their deltas into acc[p] with 16-byte REDs (L2); the last one to
//            finish (ticket) stores the GPU's slice s of the sum into owner s's inbox[p][rank] over
//            P2P, fences, and bumps owner s's arrival counter.
//   apply  : whoever observes the last arrival of shard s (ticket == num_ranks-1) sums the inbox
//            slices in rank order (deterministic), computes w = replica[p] + lr * sum and writes it
//            into replica[p^1] of EVERY rank, then releases version[s] = v+1 on every rank.
// A weights "message" is therefore a replica half plus a version word; a gradient "message" is an
// inbox slice plus an arrival count.  Quiescent snapshot: replica[p] is never written during clock v.
//
// Async protocol (eventual, or bounded delay k):
//   pull   : wait until min_q clock[q] >= v - k (skipped when k < 0), read every shard from its owner.
//   push   : the GPU's last worker REDs lr * (sum of deltas) straight into the owners' shards
//            (apply == the reduction), fences, then publishes clock[rank] = v+1 on every rank.
//
// Per-worker protocol (PSB_COMM_WORKER; eventual, or bounded delay k) — the reference's granularity: the
// MessageTracker keeps a clock per PARTITION and the server applies every gradient message on arrival
// (MessageTracker.java:42-88, ServerProcessor.java:145-151).  Every worker CTA is its own actor:
//   gate   : wait until min over ALL workers' clock words >= v - k (skipped when k < 0); the table is
//            replicated in every rank's block, so the spin is on local memory.
//   pull   : read every shard from its owner (weights may move while being read: allowed in these models).
//   push   : RED lr * delta straight into the owners' shards, fence, publish wclock[worker] = v+1 on every rank.
// A launch may run several clocks per worker (PsbComm.num_clocks): fast workers run ahead of slow ones inside
// the launch, up to k clocks.  The workers of a launch must be co-resident when k >= 0 (the launcher checks).
#pragma once
#include "common.cuh"

#define PSB_ERR_PULL_TIMEOUT 1u
#define PSB_ERR_GATE_TIMEOUT 2u

struct CommCtx {
    PsbComm c;
    CommLayout L;
};

// ---- per-worker gate: all threads of the CTA.  Returns false when the wait gave up.  rec (thread 0, optional)
//      receives what the worker saw when it was let through.
__device__ __forceinline__ bool worker_gate(const CommCtx& cx, int clock, int gw, PsbClockRecord* rec) {
    const PsbComm& c = cx.c;
    CommBlock me(c.peers[c.rank], cx.L);
    if (c.delay_cycles && (c.delay_worker == gw || c.delay_worker == -2)) {   // injected straggler
        // -2: every worker sleeps a different pseudo-random fraction of delay_cycles before every clock (jitter tests)
        const unsigned long long want = c.delay_worker == -2
            ? c.delay_cycles * (unsigned long long)((((unsigned)gw * 2654435761u) ^ ((unsigned)clock * 40503u)) >> 8 & 63u) / 64ull
            : c.delay_cycles;
        const unsigned long long t0 = clock64();
        while ((unsigned long long)(clock64() - t0) < want) __nanosleep(1000);
    }
    const unsigned int* tab = me.wclocks();
    const bool gated = c.max_delay >= 0 && clock > c.max_delay;
    const unsigned want = gated ? (unsigned)(clock - c.max_delay) : 0u;
    __shared__ unsigned int s_min, s_max;
    const unsigned long long t0 = clock64();
    for (;;) {
        unsigned mn = 0xffffffffu, mx = 0u;
        for (int i = threadIdx.x; i < c.total_workers; i += blockDim.x) {
            const unsigned v = ld_relaxed_sys(tab + i);
            mn = min(mn, v), mx = max(mx, v);
        }
        const int ok = !gated || mn == 0xffffffffu || (int)(mn - want) >= 0;
        if (__syncthreads_and(ok)) {
            if (rec) {                                            // block-wide min / max of what was just read
                if (threadIdx.x == 0) s_min = 0xffffffffu, s_max = 0u;
                __syncthreads();
                mn = __reduce_min_sync(0xffffffffu, mn), mx = __reduce_max_sync(0xffffffffu, mx);
                if ((threadIdx.x & 31) == 0) atomicMin(&s_min, mn), atomicMax(&s_max, mx);
                __syncthreads();
                if (threadIdx.x == 0) rec->clock = clock, rec->lag = clock - (int)s_min, rec->lead = (int)s_max - clock;
            }
            break;
        }
        const int gave_up = c.spin_limit && (unsigned long long)(clock64() - t0) > c.spin_limit;
        if (__syncthreads_or(gave_up)) {
            if (threadIdx.x == 0) atomicExch(me.status(), PSB_ERR_GATE_TIMEOUT);
            return false;
        }
        __nanosleep(256);
    }
    fence_acq_rel_sys();
    return true;
}

// ---- pull: all threads of the CTA; loads the start weights of clock `clock` into dst (global or shared).
__device__ __forceinline__ bool comm_pull(const CommCtx& cx, int clock, int gw, float4* dst, PsbClockRecord* rec) {
    const PsbComm& c = cx.c;
    CommBlock me(c.peers[c.rank], cx.L);
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (c.mode == PSB_COMM_SEQUENTIAL) {
        if (threadIdx.x < c.num_ranks) {
            if (!wait_ge_sys(me.version() + threadIdx.x, (unsigned)clock, c.spin_limit, me.status(),
                             PSB_ERR_PULL_TIMEOUT))
                s_ok = 0;
        }
        __syncthreads();
        if (!s_ok) return false;
        const float4* src = me.rep(clock & 1);
        for (int i = threadIdx.x; i < cx.L.n4; i += blockDim.x) dst[i] = ld_cg_v4(src + i);
    } else {
        if (c.mode == PSB_COMM_WORKER) {
            if (!worker_gate(cx, clock, gw, rec)) return false;
        } else if (c.max_delay >= 0 && clock > c.max_delay) {
            if (threadIdx.x < c.num_ranks) {
                if (!wait_ge_sys(me.clocks() + threadIdx.x, (unsigned)(clock - c.max_delay), c.spin_limit,
                                 me.status(), PSB_ERR_GATE_TIMEOUT))
                    s_ok = 0;
            }
            __syncthreads();
            if (!s_ok) return false;
        }
        for (int s = 0; s < c.num_ranks; ++s) {
            CommBlock owner(c.peers[s], cx.L);
            const int b = cx.L.slice_begin(s), e = cx.L.slice_end(s);
            const float4* src = owner.auth();
            for (int i = b + threadIdx.x; i < e; i += blockDim.x) dst[i] = ld_relaxed_sys_v4(src + (i - b));
        }
    }
    __syncthreads();
    return true;
}

// ---- push + apply: all threads of the CTA; delta (n4 float4s, global or shared) is this worker's result of
//      clock `clock` (null: the worker had no rows; it still takes part in the protocol).
__device__ __forceinline__ void comm_push(const CommCtx& cx, int clock, int gw, const float4* delta) {
    const PsbComm& c = cx.c;
    const CommLayout& L = cx.L;
    CommBlock me(c.peers[c.rank], L);
    const int p = clock & 1;
    __shared__ unsigned int s_ticket;
    __shared__ unsigned int s_apply_mask;

    if (c.mode == PSB_COMM_WORKER) {
        // this worker's gradient message, applied on arrival: w += lr * delta in the owners' shards
        if (delta) {
            for (int s = 0; s < c.num_ranks; ++s) {
                float4* w = CommBlock(c.peers[s], L).auth();
                const int b = L.slice_begin(s), e = L.slice_end(s);
                for (int i = b + threadIdx.x; i < e; i += blockDim.x) {
                    const float4 g = delta[i];
                    red_add_v4_sys(w + (i - b), make_float4(c.lr * g.x, c.lr * g.y, c.lr * g.z, c.lr * g.w));
                }
            }
        }
        fence_acq_rel_sys();
        __syncthreads();
        if (threadIdx.x < c.num_ranks)
            st_release_sys(CommBlock(c.peers[threadIdx.x], L).wclocks() + gw, (unsigned)clock + 1u);
        return;
    }

    // 1. intra-GPU reduction in L2 (system scope when the block lives on another GPU)
    float4* acc = me.acc(p);
    if (delta) {
        if (c.home_remote)
            for (int i = threadIdx.x; i < L.n4; i += blockDim.x) red_add_v4_sys(acc + i, delta[i]);
        else
            for (int i = threadIdx.x; i < L.n4; i += blockDim.x) red_add_v4(acc + i, delta[i]);
    }
    if (c.home_remote) fence_acq_rel_sys();
    else __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        s_ticket = c.home_remote ? atom_add_acqrel_sys(me.ticket(p), 1u) : atom_add_acqrel_gpu(me.ticket(p), 1u);
        s_apply_mask = 0u;
    }
    __syncthreads();
    if (s_ticket != (unsigned)c.workers_per_rank - 1u) return;   // not the last worker of this GPU
    if (c.home_remote) fence_acq_rel_sys();
    else __threadfence();

    if (c.mode == PSB_COMM_SEQUENTIAL) {
        // 2. mailbox stores: my GPU's sum, cut into shards, into each owner's inbox[p][rank]
        for (int i = threadIdx.x; i < L.n4; i += blockDim.x) {
            const int s = i / L.slice4;
            float4* box = CommBlock(c.peers[s], L).inbox(p, c.rank);
            st_cg_v4(box + (i - s * L.slice4), ld_cg_v4(acc + i));
            st_cg_v4(acc + i, make_float4(0.f, 0.f, 0.f, 0.f));          // ready for clock v+2
        }
        if (threadIdx.x == 0) *me.ticket(p) = 0u;
        fence_acq_rel_sys();
        __syncthreads();
        // 3. arrival tickets, one thread per shard
        if (threadIdx.x < c.num_ranks) {
            CommBlock owner(c.peers[threadIdx.x], L);
            const unsigned t = atom_add_acqrel_sys(owner.arrivals(p), 1u);
            if (t == (unsigned)c.num_ranks - 1u) atomicOr(&s_apply_mask, 1u << threadIdx.x);
        }
        __syncthreads();
        unsigned mask = s_apply_mask;
        if (!mask) return;
        fence_acq_rel_sys();
        // 4. apply the shards whose last arrival we saw, broadcast into everyone's next replica.
        //    One flat loop over all parameters so that every thread has its N remote inbox loads of several
        //    elements in flight at once (one NVLink round trip for the whole apply, not one per shard).
        const float4* w_old = me.rep(p);
        const bool publish = c.publish && c.num_ranks == 1;      // seqlock: odd while the snapshot is being written
        if (publish) {
            if (threadIdx.x == 0) st_relaxed_sys(me.pub_seq(), 2u * (unsigned)clock + 1u);
            fence_acq_rel_sys();
            __syncthreads();
        }
        for (int i = threadIdx.x; i < L.n4; i += blockDim.x) {
            const int s = i / L.slice4;
            if (!(mask & (1u << s))) continue;
            CommBlock owner(c.peers[s], L);
            const int off = i - s * L.slice4;
            float4 g[PSB_MAX_RANKS];
#pragma unroll
            for (int q = 0; q < PSB_MAX_RANKS; ++q)
                g[q] = q < c.num_ranks ? ld_relaxed_sys_v4(owner.inbox(p, q) + off) : make_float4(0.f, 0.f, 0.f, 0.f);
            float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int q = 0; q < PSB_MAX_RANKS; ++q) sum.x += g[q].x, sum.y += g[q].y, sum.z += g[q].z, sum.w += g[q].w;
            const float4 w = ld_cg_v4(w_old + i);
            const float4 nw = make_float4(fmaf(c.lr, sum.x, w.x), fmaf(c.lr, sum.y, w.y), fmaf(c.lr, sum.z, w.z),
                                          fmaf(c.lr, sum.w, w.w));
#pragma unroll
            for (int q = 0; q < PSB_MAX_RANKS; ++q)
                if (q < c.num_ranks) st_cg_v4(CommBlock(c.peers[q], L).rep(p ^ 1) + i, nw);
            if (publish) st_cg_v4(me.pub() + i, nw);
        }
        fence_acq_rel_sys();
        __syncthreads();
        if (publish && threadIdx.x == 0) st_release_sys(me.pub_seq(), 2u * (unsigned)clock + 2u);
        if (threadIdx.x < c.num_ranks && (mask & (1u << threadIdx.x))) {
            const int s = threadIdx.x;
            st_relaxed_sys(CommBlock(c.peers[s], L).arrivals(p), 0u);
            fence_acq_rel_sys();                 // one release fence, then N independent flag stores in flight
            for (int q = 0; q < c.num_ranks; ++q)
                st_relaxed_sys(CommBlock(c.peers[q], L).version() + s, (unsigned)clock + 1u);
        }
    } else {
        // async: apply == reduction, straight into the owners' authoritative shards
        for (int s = 0; s < c.num_ranks; ++s) {
            CommBlock owner(c.peers[s], L);
            const int b = L.slice_begin(s), e = L.slice_end(s);
            float4* w = owner.auth();
            for (int i = b + threadIdx.x; i < e; i += blockDim.x) {
                float4 g = ld_cg_v4(acc + i);
                st_cg_v4(acc + i, make_float4(0.f, 0.f, 0.f, 0.f));
                red_add_v4_sys(w + (i - b), make_float4(c.lr * g.x, c.lr * g.y, c.lr * g.z, c.lr * g.w));
            }
        }
        if (threadIdx.x == 0) *me.ticket(p) = 0u;
        fence_acq_rel_sys();
        __syncthreads();
        if (threadIdx.x < c.num_ranks)
            st_release_sys(CommBlock(c.peers[threadIdx.x], L).clocks() + c.rank, (unsigned)clock + 1u);
    }
}
