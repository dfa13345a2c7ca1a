// extern "C" surface of libpsb200.so: argument checks, error text, launches.  See include/psb200.h.
#include <stdarg.h>
#include <string.h>

#include "comm.cuh"

static thread_local char g_err[512] = "";
void psb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int psb_launch_worker_simt(const PsbSolver*, const CommCtx&, int, const void*, const int*, const int*, const float*,
                           float*, float*, int*, float*, cudaStream_t);
int psb_launch_worker_tc(const PsbSolver*, const CommCtx&, int, const void*, const int*, const int*, const float*,
                         float*, float*, int*, float*, cudaStream_t);
int psb_worker_simt_resident(const PsbSolver*);
int psb_worker_tc_resident(const PsbSolver*, int);
int psb_launch_eval_simt(const void*, const int*, int, const float*, const float*, size_t, int, int, int, int*,
                         cudaStream_t);
int psb_launch_eval_tc(const void*, const int*, int, const float*, const float*, size_t, int, int, int, int*,
                       cudaStream_t);
int psb_launch_eval_finalize(const int*, float*, int, cudaStream_t);
int psb_launch_ring_scatter(const uint32_t*, int, int, void*, int*, int, int, cudaStream_t);
int psb_launch_ring_store_dense(const void*, int, const int*, const int*, int, void*, int*, int, int, cudaStream_t);
int psb_launch_ring_store_bitmap(const uint32_t*, const int*, const void*, const int*, const int*, int, void*, int*, int, int,
                                 cudaStream_t);
int psb_launch_ring_store_palette(const uint32_t*, const int*, const void*, const int*, const int*, int, void*, int*, int, int,
                                  cudaStream_t);
int psb_launch_apply_delta(float*, const float*, float, int, int, cudaStream_t);
int psb_launch_window_stats(const void*, const int*, int, int, int, double*, float*, cudaStream_t);

extern "C" {

const char* psb_last_error(void) { return g_err; }
int psb_version(void) { return 100; }

size_t psb_comm_block_bytes(int num_features, int num_ranks) { return CommLayout::make(num_features, num_ranks).bytes; }

__global__ void psb_comm_init_kernel(CommBlock blk, const float4* __restrict__ w0, int rank) {
    const CommLayout& L = blk.L;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < L.n4) {
        const float4 v = w0[i];
        blk.rep(0)[i] = v;
        blk.rep(1)[i] = v;
        blk.pub()[i] = v;
        const int b = L.slice_begin(rank), e = L.slice_end(rank);
        if (i >= b && i < e) blk.auth()[i - b] = v;
    }
}

int psb_comm_init_block(void* d_block, const float* d_w0, int num_features, int num_ranks, int rank, void* stream) {
    PSB_REQUIRE(d_block && d_w0, "psb_comm_init_block: null pointer");
    PSB_REQUIRE(num_ranks >= 1 && num_ranks <= PSB_MAX_RANKS, "num_ranks must be in [1,%d]", PSB_MAX_RANKS);
    PSB_REQUIRE(rank >= 0 && rank < num_ranks, "rank out of range");
    const CommLayout L = CommLayout::make(num_features, num_ranks);
    cudaStream_t st = (cudaStream_t)stream;
    PSB_CHECK_CUDA(cudaMemsetAsync(d_block, 0, L.bytes, st));
    psb_comm_init_kernel<<<(L.n4 + 255) / 256, 256, 0, st>>>(CommBlock(d_block, L), (const float4*)d_w0, rank);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

float* psb_comm_replica(void* d_block, int num_features, int num_ranks, int parity) {
    return (float*)CommBlock(d_block, CommLayout::make(num_features, num_ranks)).rep(parity & 1);
}
float* psb_comm_shard_weights(void* d_block, int num_features, int num_ranks) {
    return (float*)CommBlock(d_block, CommLayout::make(num_features, num_ranks)).auth();
}
unsigned int* psb_comm_status(void* d_block, int num_features, int num_ranks) {
    return CommBlock(d_block, CommLayout::make(num_features, num_ranks)).status();
}

unsigned int* psb_comm_version_words(void* d_block, int num_features, int num_ranks) {
    return CommBlock(d_block, CommLayout::make(num_features, num_ranks)).version();
}
unsigned int* psb_comm_rank_clocks(void* d_block, int num_features, int num_ranks) {
    return CommBlock(d_block, CommLayout::make(num_features, num_ranks)).clocks();
}
unsigned int* psb_comm_worker_clocks(void* d_block, int num_features, int num_ranks) {
    return CommBlock(d_block, CommLayout::make(num_features, num_ranks)).wclocks();
}

float* psb_comm_published(void* d_block, int num_features, int num_ranks) {
    return (float*)CommBlock(d_block, CommLayout::make(num_features, num_ranks)).pub();
}
unsigned int* psb_comm_published_seq(void* d_block, int num_features, int num_ranks) {
    return CommBlock(d_block, CommLayout::make(num_features, num_ranks)).pub_seq();
}

__global__ void psb_comm_set_clocks_kernel(CommBlock blk, unsigned int clock) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < PSB_MAX_RANKS) blk.version()[i] = clock, blk.clocks()[i] = clock;
    if (i == 0) *blk.pub_seq() = 2u * clock;
    if (i < PSB_MAX_WORKERS) blk.wclocks()[i] = clock;
}
int psb_comm_set_clocks(void* d_block, int num_features, int num_ranks, unsigned int clock, void* stream) {
    PSB_REQUIRE(d_block, "psb_comm_set_clocks: null block");
    psb_comm_set_clocks_kernel<<<(PSB_MAX_WORKERS + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        CommBlock(d_block, CommLayout::make(num_features, num_ranks)), clock);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// One warp waits (bounded) until every shard's version word of this block has reached `clock`: work enqueued behind it
// on the stream (evaluation, checkpoint copies) then reads the complete weights of that clock from the replica.
__global__ void psb_comm_wait_version_kernel(CommBlock blk, int num_ranks, unsigned int clock,
                                             unsigned long long spin_limit) {
    if ((int)threadIdx.x < num_ranks)
        wait_ge_sys(blk.version() + threadIdx.x, clock, spin_limit, blk.status(), PSB_ERR_PULL_TIMEOUT);
}
int psb_comm_wait_version(void* d_block, int num_features, int num_ranks, unsigned int clock,
                          unsigned long long spin_limit, void* stream) {
    PSB_REQUIRE(d_block && num_ranks >= 1 && num_ranks <= PSB_MAX_RANKS, "psb_comm_wait_version: bad arguments");
    psb_comm_wait_version_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
        CommBlock(d_block, CommLayout::make(num_features, num_ranks)), num_ranks, clock, spin_limit);
    PSB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// Reads the block's status word (synchronises the stream): 0, or PSB_E_TIMEOUT with the reason in psb_last_error().
int psb_comm_check(void* d_block, int num_features, int num_ranks, void* stream) {
    PSB_REQUIRE(d_block, "psb_comm_check: null block");
    unsigned int st = 0;
    PSB_CHECK_CUDA(cudaMemcpyAsync(&st, CommBlock(d_block, CommLayout::make(num_features, num_ranks)).status(), 4,
                                   cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PSB_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (st == 0) return 0;
    psb_set_error("in-kernel wait gave up (status %u: %s)", st,
                  st == PSB_ERR_PULL_TIMEOUT ? "the weights of this clock never arrived — a peer stopped pushing"
                  : st == PSB_ERR_GATE_TIMEOUT ? "the bounded-delay clock gate never opened — a peer fell behind and stopped"
                                               : "unknown");
    return PSB_E_TIMEOUT;
}

size_t psb_worker_scratch_bytes(int num_features) {
    return (size_t)PSB_SCRATCH_VECS * (num_features + 1) * PSB_CLASS_PAD * sizeof(float);
}

int psb_worker_step(const PsbSolver* s, const PsbComm* comm, int num_workers, int impl, const void* d_X,
                    const int* d_labels, const int* d_win, const float* d_w_in, float* d_delta, float* d_loss,
                    int* d_info, float* d_scratch, void* stream) {
    PSB_REQUIRE(s && comm, "psb_worker_step: null solver/comm");
    PSB_REQUIRE(num_workers >= 1, "num_workers must be >= 1");
    PSB_REQUIRE(s->num_class_rows >= 2 && s->num_class_rows <= PSB_CLASS_PAD, "num_class_rows must be in [2,8]");
    PSB_REQUIRE(s->num_features >= 2 && s->cap >= 1, "bad num_features/cap");
    PSB_REQUIRE(d_X && d_labels && d_win && d_loss && d_info && d_scratch, "psb_worker_step: null buffer");
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
            // a launch is any group of this rank's workers: [worker_base, worker_base + num_workers)
            PSB_REQUIRE(comm->total_workers >= 1 && comm->total_workers <= PSB_MAX_WORKERS,
                        "total_workers must be in [1,%d] (got %d)", PSB_MAX_WORKERS, comm->total_workers);
            PSB_REQUIRE(comm->worker_base >= 0 && comm->worker_base + num_workers <= comm->total_workers,
                        "workers [%d,%d) are outside the clock table of %d workers", comm->worker_base,
                        comm->worker_base + num_workers, comm->total_workers);
            PSB_REQUIRE(comm->num_clocks >= 0 && comm->num_clocks <= 1 << 20, "num_clocks out of range");
        } else {
            PSB_REQUIRE(comm->workers_per_rank == num_workers, "workers_per_rank must equal num_workers");
            PSB_REQUIRE(comm->num_clocks <= 1 || comm->mode == PSB_COMM_SEQUENTIAL,
                        "several clocks per launch need comm mode PSB_COMM_WORKER or PSB_COMM_SEQUENTIAL");
            PSB_REQUIRE(!comm->publish || (comm->mode == PSB_COMM_SEQUENTIAL && comm->num_ranks == 1),
                        "the published snapshot is kept in sequential mode on one rank only");
        }
    }
    if ((comm->mode == PSB_COMM_WORKER && comm->max_delay >= 0) ||
        (comm->mode == PSB_COMM_SEQUENTIAL && comm->num_clocks > 1)) {
        // gated workers spin on each other's clock words: every CTA of the launch has to be on the GPU at once
        const int fit = psb_worker_max_resident(s, impl, num_workers);
        if (num_workers > fit) {
            psb_set_error("launch of %d workers that wait on each other (bounded delay, or several sequential clocks "
                          "per launch), but only %d worker CTAs of this shape are resident at once on this device: "
                          "launch fewer workers per GPU", num_workers, fit);
            return PSB_E_LIMIT;
        }
    }
    if (impl == 0)
        return psb_launch_worker_simt(s, cx, num_workers, d_X, d_labels, d_win, d_w_in, d_delta, d_loss, d_info,
                                      d_scratch, (cudaStream_t)stream);
    return psb_launch_worker_tc(s, cx, num_workers, d_X, d_labels, d_win, d_w_in, d_delta, d_loss, d_info, d_scratch,
                                (cudaStream_t)stream);
}

int psb_worker_max_resident(const PsbSolver* s, int impl, int num_workers) {
    PSB_REQUIRE(s && num_workers >= 1, "psb_worker_max_resident: bad arguments");
    return impl == 0 ? psb_worker_simt_resident(s) : psb_worker_tc_resident(s, num_workers);
}

int psb_apply_delta(float* d_w, const float* d_delta, float lr, int num_features, int skip_class, void* stream) {
    PSB_REQUIRE(d_w && d_delta, "psb_apply_delta: null pointer");
    return psb_launch_apply_delta(d_w, d_delta, lr, num_features, skip_class, (cudaStream_t)stream);
}

int psb_eval(const void* d_X, const int* d_labels, int n, const float* d_w, int num_features, int num_class_rows,
             int* d_conf, float* d_out, int impl, void* stream) {
    PSB_REQUIRE(d_X && d_labels && d_w && d_conf && d_out, "psb_eval: null pointer");
    PSB_REQUIRE(n >= 0 && num_features >= 2 && num_features % 2 == 0, "psb_eval: bad shape");
    PSB_REQUIRE(num_class_rows >= 2 && num_class_rows <= PSB_CLASS_PAD, "num_class_rows must be in [2,8]");
    return psb_eval_models(d_X, d_labels, n, d_w, nullptr, 0, 1, num_features, num_class_rows, d_conf, d_out, impl,
                           stream);
}

int psb_eval_models(const void* d_X, const int* d_labels, int n, const float* d_w, const float* d_w_add,
                    size_t model_stride_floats, int num_models, int num_features, int num_class_rows, int* d_conf,
                    float* d_out, int impl, void* stream) {
    PSB_REQUIRE(d_X && d_labels && d_w && d_conf && d_out, "psb_eval_models: null pointer");
    PSB_REQUIRE(n >= 0 && num_features >= 2 && num_features % 2 == 0, "psb_eval_models: bad shape");
    PSB_REQUIRE(num_class_rows >= 2 && num_class_rows <= PSB_CLASS_PAD, "num_class_rows must be in [2,8]");
    PSB_REQUIRE(num_models >= 1 && num_models <= 65535, "num_models must be in [1,65535]");
    PSB_REQUIRE(model_stride_floats % 4 == 0, "model stride must be a multiple of 4 floats");
    cudaStream_t st = (cudaStream_t)stream;
    PSB_CHECK_CUDA(cudaMemsetAsync(d_conf, 0, (size_t)num_models * 64 * sizeof(int), st));
    int rc = 0;
    if (n > 0)
        rc = impl == 0 ? psb_launch_eval_simt(d_X, d_labels, n, d_w, d_w_add, model_stride_floats, num_models,
                                              num_features, num_class_rows, d_conf, st)
                       : psb_launch_eval_tc(d_X, d_labels, n, d_w, d_w_add, model_stride_floats, num_models,
                                            num_features, num_class_rows, d_conf, st);
    if (rc) return rc;
    return psb_launch_eval_finalize(d_conf, d_out, num_models, st);
}

int psb_eval_finalize(const int* d_conf, float* d_out, void* stream) {
    PSB_REQUIRE(d_conf && d_out, "psb_eval_finalize: null pointer");
    return psb_launch_eval_finalize(d_conf, d_out, 1, (cudaStream_t)stream);
}

int psb_ring_scatter(const uint32_t* d_stage, int n, int stride_words, void* d_X, int* d_labels, int cap,
                     int num_features, void* stream) {
    PSB_REQUIRE(d_stage && d_X && d_labels, "psb_ring_scatter: null pointer");
    PSB_REQUIRE(num_features % 8 == 0, "num_features must be a multiple of 8");
    return psb_launch_ring_scatter(d_stage, n, stride_words, d_X, d_labels, cap, num_features, (cudaStream_t)stream);
}
int psb_ring_store_dense(const void* d_rows, int src_half, const int* d_labels_in, const int* d_worker_slot, int n,
                         void* d_X, int* d_labels, int cap, int num_features, void* stream) {
    PSB_REQUIRE(d_rows && d_labels_in && d_worker_slot && d_X && d_labels, "psb_ring_store_dense: null pointer");
    PSB_REQUIRE(num_features % 8 == 0, "num_features must be a multiple of 8");
    return psb_launch_ring_store_dense(d_rows, src_half, d_labels_in, d_worker_slot, n, d_X, d_labels, cap,
                                       num_features, (cudaStream_t)stream);
}

int psb_ring_store_bitmap(const uint32_t* d_bitmap, const int* d_row_off, const void* d_values, const int* d_labels_in,
                          const int* d_worker_slot, int n, void* d_X, int* d_labels, int cap, int num_features,
                          void* stream) {
    PSB_REQUIRE(d_bitmap && d_row_off && d_values && d_labels_in && d_worker_slot && d_X && d_labels,
                "psb_ring_store_bitmap: null pointer");
    PSB_REQUIRE(num_features % 32 == 0, "psb_ring_store_bitmap: num_features must be a multiple of 32");
    return psb_launch_ring_store_bitmap(d_bitmap, d_row_off, d_values, d_labels_in, d_worker_slot, n, d_X, d_labels, cap,
                                        num_features, (cudaStream_t)stream);
}

int psb_ring_store_palette(const uint32_t* d_bitmap, const int* d_row_off, const void* d_payload, const int* d_labels_in,
                           const int* d_worker_slot, int n, void* d_X, int* d_labels, int cap, int num_features,
                           void* stream) {
    PSB_REQUIRE(d_bitmap && d_row_off && d_payload && d_labels_in && d_worker_slot && d_X && d_labels,
                "psb_ring_store_palette: null pointer");
    PSB_REQUIRE(num_features % 32 == 0, "psb_ring_store_palette: num_features must be a multiple of 32");
    return psb_launch_ring_store_palette(d_bitmap, d_row_off, d_payload, d_labels_in, d_worker_slot, n, d_X, d_labels,
                                         cap, num_features, (cudaStream_t)stream);
}

int psb_window_stats_update(const void* d_X, const int* d_change, int num_workers, int cap, int num_features,
                            double* d_sums, float* d_inv_std, void* stream) {
    PSB_REQUIRE(d_X && d_change && d_sums && d_inv_std, "psb_window_stats_update: null pointer");
    PSB_REQUIRE(num_features % 4 == 0 && cap >= 1, "psb_window_stats_update: num_features must be a multiple of 4");
    return psb_launch_window_stats(d_X, d_change, num_workers, cap, num_features, d_sums, d_inv_std,
                                   (cudaStream_t)stream);
}

// Several host -> device copies with one call (the ingestion path issues 2-4 small copies per batch; from Python
// each costs more in interpreter time than on the wire).  Sources should be pinned for the copies to be asynchronous.
int psb_h2d_copies(int n, void* const* d_dst, const void* const* h_src, const size_t* bytes, void* stream) {
    PSB_REQUIRE(n >= 0 && (n == 0 || (d_dst && h_src && bytes)), "psb_h2d_copies: null table");
    for (int i = 0; i < n; ++i) {
        if (bytes[i] == 0) continue;
        PSB_REQUIRE(d_dst[i] && h_src[i], "psb_h2d_copies: null pointer in entry %d", i);
        PSB_CHECK_CUDA(cudaMemcpyAsync(d_dst[i], h_src[i], bytes[i], cudaMemcpyHostToDevice, (cudaStream_t)stream));
    }
    return 0;
}

// One device -> host copy on `stream` (pinned destination for an asynchronous copy): control words and batch
// headers of the two-process transport.
int psb_d2h_copy(void* h_dst, const void* d_src, size_t bytes, void* stream) {
    PSB_REQUIRE(h_dst && d_src, "psb_d2h_copy: null pointer");
    PSB_CHECK_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return 0;
}
// device -> device copy on `stream` (snapshots of weights that live in another process' block)
int psb_d2d_copy(void* d_dst, const void* d_src, size_t bytes, void* stream) {
    PSB_REQUIRE(d_dst && d_src, "psb_d2d_copy: null pointer");
    PSB_CHECK_CUDA(cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}

// Random 32-byte row traffic (sharded sparse path): ask L2 to fetch 32-byte sectors instead of promoting every
// miss to 64/128 bytes.  A hint; returns the granularity the device reports afterwards.
int psb_set_l2_fetch_granularity(int bytes) {
    PSB_CHECK_CUDA(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes));
    size_t got = 0;
    PSB_CHECK_CUDA(cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity));
    return (int)got;
}

// ------------------------------------------------------------------ peer memory plumbing
int psb_malloc(void** d_ptr_out, size_t bytes) {
    PSB_REQUIRE(d_ptr_out && bytes > 0, "psb_malloc: bad arguments");
    PSB_CHECK_CUDA(cudaMalloc(d_ptr_out, bytes));
    return 0;
}
int psb_free(void* d_ptr) {
    PSB_CHECK_CUDA(cudaFree(d_ptr));
    return 0;
}
int psb_ipc_export(void* d_ptr, unsigned char handle_out[64]) {
    cudaIpcMemHandle_t h;
    PSB_CHECK_CUDA(cudaIpcGetMemHandle(&h, d_ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out, &h, 64);
    return 0;
}
int psb_ipc_import(const unsigned char handle[64], void** d_ptr_out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    PSB_CHECK_CUDA(cudaIpcOpenMemHandle(d_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int psb_ipc_close(void* d_ptr) {
    PSB_CHECK_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return 0;
}
int psb_enable_peer_access(int num_devices) {
    int cur = 0;
    PSB_CHECK_CUDA(cudaGetDevice(&cur));
    for (int d = 0; d < num_devices; ++d) {
        if (d == cur) continue;
        int can = 0;
        PSB_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, cur, d));
        if (!can) {
            psb_set_error("device %d cannot access peer %d", cur, d);
            return PSB_E_CUDA;
        }
        cudaError_t e = cudaDeviceEnablePeerAccess(d, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
            psb_set_error("cudaDeviceEnablePeerAccess(%d): %s", d, cudaGetErrorString(e));
            return PSB_E_CUDA;
        }
        cudaGetLastError();
    }
    return 0;
}

}  // extern "C"
