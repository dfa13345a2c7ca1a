/*
 * psb200 — C ABI of the Blackwell-native parameter server (libpsb200.so).
 *
 * Every entry point returns 0 on success or a negative PSB_E_* code; the text of
 * the last error on the calling thread is available from psb_last_error().
 * Pointers named d_* are device pointers, h_* are host pointers.  Kernels are
 * enqueued on the cudaStream_t passed as `void* stream` and never synchronise.
 *
 * Parameter layout (device): W[(F+1)][8] fp32 — feature-major, one 32-byte
 * sector per feature, row F = intercepts, class columns >= R are padding.
 * Reference key k <-> W[k / R][k % R] for k < R*F, intercept c <-> W[F][c]
 * (LogisticRegressionTaskSpark.java:98-140, 173).
 *
 * Input buffer (device): per worker a ring of `cap` rows, X[w][cap][F] fp16
 * dense + labels[w][cap] int32 + win[w] = {head, count}: the active rows are the
 * `count` most recent insertions, slots (head-count .. head-1) mod cap
 * (WorkerSamplingProcessor.java:50-113 => "last min(prev+1, target) rows").
 */
#ifndef PSB200_H
#define PSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PSB_CLASS_PAD 8
#define PSB_MAX_RANKS 8
#define PSB_HIST 4                 /* L-BFGS pairs kept on device (reference runs 2 iterations) */
#define PSB_SCRATCH_VECS (7 + 2 * PSB_HIST)   /* x g d xt gt tmp w0 + s[], y[] */

#define PSB_E_INVALID (-1)
#define PSB_E_CUDA (-2)
#define PSB_E_LIMIT (-3)
#define PSB_E_TIMEOUT (-4)

#define PSB_SOLVER_LBFGS 0
#define PSB_SOLVER_SGD 1

#define PSB_COMM_NONE 0            /* write delta to d_delta (host applies)                     */
#define PSB_COMM_SEQUENTIAL 1      /* mailbox push + last-arriver apply + replica broadcast      */
#define PSB_COMM_ASYNC 2           /* red.add lr*delta into owner shard + clock table (k or inf) */
#define PSB_COMM_WORKER 3          /* as ASYNC, but every WORKER owns a clock word, applies its own delta and is
                                      gated on its own: the reference's per-partition MessageTracker state
                                      (MessageTracker.java:42-88) evaluated by each worker CTA            */
#define PSB_MAX_WORKERS 4096       /* entries of the per-worker clock table in a peer-memory block          */

const char* psb_last_error(void);
int psb_version(void);

/* ---------------------------------------------------------------- model / solver */
typedef struct PsbSolver {
    int num_features;              /* F                                                    */
    int num_class_rows;            /* R = numClasses + 1  (<= 8)                             */
    int cap;                       /* ring capacity per worker (max_buffer_size)           */
    int solver;                    /* PSB_SOLVER_*                                          */
    int max_iter;                  /* L-BFGS iterations (reference: 2)                      */
    int standardize;               /* Spark standardization=true semantics                  */
    int center;                    /* centre coefficients / intercepts across classes       */
    float sgd_lr;
    const float* inv_std;          /* optional (device): precomputed 1/std per worker and feature, [num_workers][F],
                                      kept current by psb_window_stats_update; NULL: the kernel computes the
                                      statistics of the window itself, fused into its first pass              */
} PsbSolver;

/* Peer-memory block of one rank.  The same layout lives on every GPU; `peers[q]` is
 * rank q's block mapped into this process (cudaIpc) — peers[rank] is the local one.
 * Replaces WEIGHTS_TOPIC / GRADIENTS_TOPIC and the MessageTracker state
 * (BaseKafkaApp.java:27-33, MessageTracker.java:42-88). */
typedef struct PsbComm {
    int mode;                      /* PSB_COMM_*                                            */
    int rank, num_ranks;
    int workers_per_rank;          /* worker CTAs launched on every rank                    */
    int clock;                     /* vector clock of this step                             */
    int max_delay;                 /* bounded delay k (async mode; <0: eventual)            */
    float lr;                      /* server learning rate 1/numWorkers (ServerProcessor.java:36) */
    unsigned long long spin_limit; /* clock64 cycles before a wait gives up (watchdog)      */
    void* peers[PSB_MAX_RANKS];
    /* ---- PSB_COMM_WORKER only (leave zero otherwise) */
    int num_clocks;                /* clocks every worker of this launch runs back to back (0 counts as 1): each at
                                      its own pace in PSB_COMM_WORKER; in PSB_COMM_SEQUENTIAL the workers meet at
                                      every clock's version words, without a launch in between (all workers of all
                                      ranks must be resident).  The window of a worker is read once per launch */
    int worker_base;               /* global id of the launch's worker 0 (clock table index)                 */
    int total_workers;             /* workers of all ranks = live entries of the clock table (<= PSB_MAX_WORKERS) */
    int delay_worker;              /* straggler injection: this global worker id (-2: every worker, a different
                                      pseudo-random fraction each clock) sleeps delay_cycles before            */
    unsigned long long delay_cycles; /* every clock (0: off)                                                  */
    void* hist;                    /* optional PsbClockRecord[num_clocks][num_workers] of this launch (device) */
    int home_remote;               /* peers[rank] lives on another GPU (server process' block): the intra-rank
                                      reduction and ticket then use system-scope operations                   */
    int publish;                   /* sequential mode, one rank: the applier also writes the new weights into the
                                      block's published snapshot under a seqlock word (2v+1 while clock v's weights
                                      are being written, 2v+2 once complete) — what a server process that does
                                      not take part in the clocks reads (psb_comm_published)                  */
} PsbComm;

/* What a worker saw when it started clock `clock`: lag = clock - min over all workers' clocks at the moment the
 * gate let it through (<= max_delay under bounded delay), lead = max over all workers' clocks - clock. */
typedef struct PsbClockRecord {
    float loss;
    int clock, lag, lead;
} PsbClockRecord;

size_t psb_comm_block_bytes(int num_features, int num_ranks);
/* zero the block and write the initial weights into both replicas and the owned shard */
int psb_comm_init_block(void* d_block, const float* d_w0, int num_features, int num_ranks, int rank, void* stream);
/* host-visible views into a block (device pointers) */
float* psb_comm_replica(void* d_block, int num_features, int num_ranks, int parity);
float* psb_comm_shard_weights(void* d_block, int num_features, int num_ranks);
unsigned int* psb_comm_status(void* d_block, int num_features, int num_ranks);   /* [0]=error flag, [1..]=debug */
unsigned int* psb_comm_version_words(void* d_block, int num_features, int num_ranks);   /* [PSB_MAX_RANKS] */
unsigned int* psb_comm_rank_clocks(void* d_block, int num_features, int num_ranks);     /* [PSB_MAX_RANKS] */
unsigned int* psb_comm_worker_clocks(void* d_block, int num_features, int num_ranks);   /* [PSB_MAX_WORKERS] */
/* published snapshot ([(F+1)*8] floats) and its seqlock word (see PsbComm.publish) */
float* psb_comm_published(void* d_block, int num_features, int num_ranks);
unsigned int* psb_comm_published_seq(void* d_block, int num_features, int num_ranks);
/* after a resume: every version / rank-clock / worker-clock word of the block restarts at `clock` */
int psb_comm_set_clocks(void* d_block, int num_features, int num_ranks, unsigned int clock, void* stream);

/* Sequential mode: enqueue a (bounded) device-side wait until all shards of clock `clock` have been published into
 * this block's replica (version words >= clock); later work on `stream` sees the complete snapshot. */
int psb_comm_wait_version(void* d_block, int num_features, int num_ranks, unsigned int clock,
                          unsigned long long spin_limit, void* stream);
/* 0, or PSB_E_TIMEOUT when an in-kernel wait of this rank gave up (synchronises `stream`) */
int psb_comm_check(void* d_block, int num_features, int num_ranks, void* stream);

size_t psb_worker_scratch_bytes(int num_features);                               /* per worker */

/* One parameter-server iteration for `num_workers` local workers (one CTA each):
 * pull (wait + load weights) -> local solve on the buffer -> push/apply per comm.mode.
 * Replaces WorkerTrainingProcessor.process + ServerProcessor.process
 * (WorkerTrainingProcessor.java:63-98, ServerProcessor.java:143-183).
 *   d_w_in  : weights to start from when comm.mode == NONE (else the replica is used)
 *   d_delta : [num_workers][(F+1)*8] out when comm.mode == NONE (may be NULL otherwise)
 *   d_loss  : [num_workers] final train loss of the local solve
 *   d_info  : [num_workers][8] ints: evals, iterations (tcgen05 kernel: -site of the barrier wait that gave up when
 *             status is -2), status, active rows, then SM cycles / 64 spent in
 *             pull+statistics, local solve, delta+push, whole step
 *   impl    : 0 = SIMT fp32 kernel, 1 = tcgen05/TMA kernel                                */
int psb_worker_step(const PsbSolver* s, const PsbComm* comm, int num_workers, int impl,
                    const void* d_X, const int* d_labels, const int* d_win,
                    const float* d_w_in, float* d_delta, float* d_loss, int* d_info,
                    float* d_scratch, void* stream);

/* The same step on SPARSE-row buffers (ELL rings: d_idx [num_workers][cap][max_nnz] int32 feature ids, d_val the
 * fp16 values, d_nnz [num_workers][cap] non-zeros per slot): any num_features, no densification — the reference
 * fits on sparse vectors of any width (LogisticRegressionTaskSpark.java:146-184).  Same solver, same comm modes
 * (including PSB_COMM_WORKER and several clocks per launch); the L-BFGS memory is PSB_HIST - 1 pairs here. */
int psb_worker_step_csr(const PsbSolver* s, const PsbComm* comm, int num_workers, int max_nnz, const int* d_idx,
                        const void* d_val, const int* d_nnz, const int* d_labels, const int* d_win,
                        const float* d_w_in, float* d_delta, float* d_loss, int* d_info, float* d_scratch,
                        void* stream);
/* staged sparse rows (the record format of psb_ring_scatter) -> ELL ring slots; rows keep their first max_nnz non-zeros */
int psb_ring_store_ell(const uint32_t* d_stage, int n, int stride_words, int* d_idx, void* d_val, int* d_nnz,
                       int* d_labels, int cap, int max_nnz, int num_features, void* stream);

/* Worker CTAs of this shape that fit on the current device at the same time when `num_workers` are launched
 * together.  Bounded-delay PSB_COMM_WORKER launches spin on each other's clock words inside the kernel, so all
 * workers that can be in flight on a GPU together (all groups, all streams) must fit; psb_worker_step checks its
 * own launch and returns PSB_E_LIMIT otherwise. */
int psb_worker_max_resident(const PsbSolver* s, int impl, int num_workers);

/* The same step for callers that must not allocate device memory (a host process whose GPU memory
 * belongs to another framework's pool, a real-time loop that cannot afford cudaMalloc's device-wide
 * synchronisation, a plugin ABI that forbids allocation after start-up): the solver scratch is a fixed
 * array inside the library's device image instead of a caller buffer.  It has no counterpart in the
 * reference (the JVM allocates freely); it exists for embedders of this library.
 * Limit: psb_static_scratch_bytes() (about 1.9 MiB per GPU, held until exit) = 4 workers at 1024
 * features; larger requests return PSB_E_LIMIT and launch nothing.  All calls on a device share the
 * array (it keeps the previous call's contents; nothing in it needs zeroing), so calls that can overlap
 * on the GPU race — keep them on one stream.  The first call on each device resolves the array's
 * address and should be made outside stream capture.  See csrc/ps_noalloc.cu. */
size_t psb_static_scratch_bytes(void);
int psb_worker_step_static(const PsbSolver* s, const PsbComm* comm, int num_workers, int impl,
                           const void* d_X, const int* d_labels, const int* d_win,
                           const float* d_w_in, float* d_delta, float* d_loss, int* d_info, void* stream);

/* Server-side apply for comm.mode == NONE: w += lr * delta (ServerProcessor.java:148-151, 225-228);
 * skip_class >= 0 leaves that class' intercept untouched (quirk Q1 compatibility), -1 updates all. */
int psb_apply_delta(float* d_w, const float* d_delta, float lr, int num_features, int skip_class,
                    void* stream);

/* ---------------------------------------------------------------- sharded sparse path (2^24 features) */
/* Weights range-partitioned over the ranks (KeyRange, BaseMessage.java:24-27; the reference keeps a
 * single server partition, ServerApp.java:38).  shard[s]: (rows_per_shard + 1) x 8 floats on rank s,
 * the intercept row is the extra row of shard 0.  phase[q]: num_ranks counters on rank q. */
typedef struct PsbSparseComm {
    void* shard[PSB_MAX_RANKS];
    void* phase[PSB_MAX_RANKS];
    void* ticket;                  /* one zeroed uint32 on this rank */
    void* status;                  /* one uint32 error flag on this rank */
    int rank, num_ranks, rows_per_shard;
    unsigned long long spin_limit;
} PsbSparseComm;
/* One SGD step of a worker on a CSR batch.  mode bit 0 = forward (pull fused into the GEMV operand loads,
 * residuals -> d_residuals when given), bit 1 = scatter (push + apply fused: w += scale * v * r over P2P).
 * wait_phase / publish_phase (0 = none): device-side barrier over peer memory for sequential consistency.
 * scale = -eta / (batch * numWorkers).  d_loss: one float, the batch's summed loss is ADDED to it. */
int psb_sparse_step(const PsbSparseComm* comm, int mode, unsigned int wait_phase, unsigned int publish_phase,
                    const long long* d_indptr, const int* d_indices, const float* d_values, const int* d_labels,
                    int batch, int num_features, int num_class_rows, float scale, float* d_residuals, float* d_loss,
                    void* stream);
int psb_sparse_eval(const PsbSparseComm* comm, const long long* d_indptr, const int* d_indices, const float* d_values,
                    const int* d_labels, int n, int num_features, int num_class_rows, int* d_conf, void* stream);

/* Slab exchange variant of the sharded step: every NVLink transfer is a contiguous slab (request lists,
 * reply rows = the pull, update rows = the push) instead of one small request per non-zero.
 * req[s]: int[2][N][cap] on rank s; req_cnt[s]: uint[2][N] on rank s; rep[q] / upd[s]: float[N][cap][8];
 * cnt: uint[N] zeroed once; pos: int[nnz] local.  stage_mask bits: 1 bucket, 2 serve, 4 step, 8 apply
 * (15 = whole step; single bits let one device emulate several ranks in launch order). */
typedef struct PsbSlabComm {
    PsbSparseComm base;
    void* req[PSB_MAX_RANKS];
    void* req_cnt[PSB_MAX_RANKS];
    void* rep[PSB_MAX_RANKS];
    void* upd[PSB_MAX_RANKS];
    void* cnt;
    void* pos;
    int cap;
    int parity;
} PsbSlabComm;
int psb_sparse_step_slab(const PsbSlabComm* comm, int step_index, const long long* d_indptr, const int* d_indices,
                         const float* d_values, const int* d_labels, int batch, long long nnz, int num_features,
                         int num_class_rows, float scale, float* d_loss, int stage_mask, void* stream);

/* ---------------------------------------------------------------- evaluation */
/* argmax(X W + b) -> confusion matrix -> weighted F1, accuracy (Metrics.java:15-24).
 *   d_X fp16 dense [n][F]; d_conf: int[64] scratch (8x8, zeroed by the call);
 *   d_out: float[2] = {f1, accuracy}.  impl as above.                                     */
int psb_eval(const void* d_X, const int* d_labels, int n, const float* d_w, int num_features,
             int num_class_rows, int* d_conf, float* d_out, int impl, void* stream);

/* The same for `num_models` weight sets in ONE launch: model m is d_w + m * model_stride_floats, plus, element-wise,
 * d_w_add + m * model_stride_floats when d_w_add is given (a worker's locally fitted model = the weights it pulled +
 * its delta: the worker log row of WorkerTrainingProcessor.java:85-92 scores that model on the test set,
 * LogisticRegressionTaskSpark.java:186).  d_conf: int[num_models][64]; d_out: float[num_models][2]. */
int psb_eval_models(const void* d_X, const int* d_labels, int n, const float* d_w, const float* d_w_add,
                    size_t model_stride_floats, int num_models, int num_features, int num_class_rows, int* d_conf,
                    float* d_out, int impl, void* stream);

/* weighted F1 + accuracy from an 8x8 confusion matrix (shared by the dense and the sparse evaluation) */
int psb_eval_finalize(const int* d_conf, float* d_out, void* stream);

/* ---------------------------------------------------------------- ingestion */
/* Scatter a staged batch of sparse rows into the dense fp16 ring slots (K1).
 *   d_stage: packed rows [n][stride_words] (word0 = nnz, word1 = label, word2 = worker,
 *   word3 = slot, then nnz x (int32 index, fp32 value)); the slot row is zero-filled first. */
int psb_ring_scatter(const uint32_t* d_stage, int n, int stride_words, void* d_X, int* d_labels,
                     int cap, int num_features, void* stream);
/* dense rows [n][F] (fp32 when src_half == 0, fp16 otherwise) -> ring slots given (worker, slot) pairs */
int psb_ring_store_dense(const void* d_rows, int src_half, const int* d_labels_in, const int* d_worker_slot, int n,
                         void* d_X, int* d_labels, int cap, int num_features, void* stream);

/* Bitmap-packed rows -> ring slots: d_bitmap [n][F/32] presence words (bit b of word w = feature 32w+b),
 * d_row_off [n] = index of the row's first value in d_values (fp16, feature order).  The sparse wire format
 * of LabeledData (LabeledData.java:14-28) without the keys: 128 B + 2 B per non-zero at F = 1024. */
int psb_ring_store_bitmap(const uint32_t* d_bitmap, const int* d_row_off, const void* d_values,
                          const int* d_labels_in, const int* d_worker_slot, int n, void* d_X, int* d_labels,
                          int cap, int num_features, void* stream);

/* Bitmap + per-row palette rows -> ring slots (see psb_pack_palette_dense for the payload format):
 * d_row_off [n] = byte offset of the row's payload in d_payload. */
int psb_ring_store_palette(const uint32_t* d_bitmap, const int* d_row_off, const void* d_payload,
                           const int* d_labels_in, const int* d_worker_slot, int n, void* d_X, int* d_labels,
                           int cap, int num_features, void* stream);

/* Per-feature statistics of every worker's window, maintained incrementally at ingestion time (Spark recomputes
 * them on every fit, LogisticRegressionTaskSpark.java:179-184 with standardization=true; the window of a streaming
 * buffer changes by a few rows per iteration).  d_change [num_workers][4] ints: {old_head, old_count, rows_added,
 * new_count}; the new head is (old_head + rows_added) mod cap.  The kernel subtracts the rows that left the window
 * and adds the rows that entered it — both must still be in the ring, i.e. rows_added <= cap - old_count; a worker
 * whose change says rows_added < 0 (or violates that) is recomputed from its whole new window.  d_sums
 * [num_workers][F][2] doubles (sum, sum of squares; exact products of fp16 values), d_inv_std [num_workers][F]
 * floats out: 1/std (unbiased), 0 for constant columns. */
int psb_window_stats_update(const void* d_X, const int* d_change, int num_workers, int cap, int num_features,
                            double* d_sums, float* d_inv_std, void* stream);

/* n host -> device copies on `stream` with one call (pinned sources for asynchronous copies) */
int psb_h2d_copies(int n, void* const* d_dst, const void* const* h_src, const size_t* bytes, void* stream);

/* single device -> host / device -> device copies on `stream` (control words, weight snapshots) */
int psb_d2h_copy(void* h_dst, const void* d_src, size_t bytes, void* stream);
int psb_d2d_copy(void* d_dst, const void* d_src, size_t bytes, void* stream);

/* ---------------------------------------------------------------- host runtime (C++) */
/* Dense fp16 rows -> presence bitmap + per-row palette (the <= 16 distinct fp16 values of the row) + one 4-bit
 * code per non-zero; rows with more distinct values keep raw fp16 values.  Bit-exact.  out_row_off [n+1] are byte
 * offsets into out_payload (capacity in bytes; n * (2 F + 2) always suffices).  Returns the payload bytes, -1 on error. */
long long psb_pack_palette_dense(const uint16_t* rows, long long n_rows, int num_features, uint32_t* out_bitmap,
                                 int* out_row_off, uint8_t* out_payload, long long payload_capacity);
/* CSR (sorted or unsorted column indices, fp32 values) -> bitmap-packed rows with fp16 values (host side of
 * psb_ring_store_bitmap).  out_bitmap [n][F/32], out_row_off [n+1], out_values: uint16 fp16 bits, capacity
 * indptr[n].  Explicit zeros and out-of-range columns are dropped.  Returns the number of values written. */
long long psb_pack_bitmap_csr(const long long* indptr, const int* indices, const float* values, long long n_rows,
                              int num_features, uint32_t* out_bitmap, int* out_row_off, uint16_t* out_values);
/* the same from dense fp16 rows [n][F] (uint16 bit patterns; -0.0 counts as zero) */
long long psb_pack_bitmap_dense(const uint16_t* rows, long long n_rows, int num_features, uint32_t* out_bitmap,
                                int* out_row_off, uint16_t* out_values);

/* CSV -> CSR tokenizer (CsvProducer.java:46-58).  Two-pass: count, then fill. */
int psb_csv_count(const char* path, int has_header, long long* n_rows, long long* n_nnz, int* n_cols);
int psb_csv_parse(const char* path, int has_header, long long* indptr, int* indices, float* values,
                  int* labels, long long n_rows, long long n_nnz);
/* The same in one pass over the file, tokenised by `threads` threads (<= 0: up to 8): open, ask for the shape, fill
 * caller arrays indptr[rows + 1] / indices[nnz] / values[nnz] / labels[rows], close.  NULL when the file cannot be read. */
void* psb_csv_open(const char* path, int has_header, int threads);
int psb_csv_shape(void* csv, long long* n_rows, long long* n_nnz, int* n_cols);
int psb_csv_fill(void* csv, long long* indptr, int* indices, float* values, int* labels);
void psb_csv_close(void* csv);

/* Consistency tracker (MessageTracker.java) — returns replies as (worker, clock) pairs */
void* psb_tracker_create(int num_workers, int consistency_model);
void psb_tracker_destroy(void* t);
int psb_tracker_on_gradient(void* t, int worker, int clock, int* replies, int max_replies);
int psb_tracker_clocks(void* t, int* clocks);

/* Dynamic buffer policy (WorkerSamplingProcessor.java:115-135) */
void* psb_buffer_create(int min_size, int max_size, float bc, int window);
void psb_buffer_destroy(void* b);
int psb_buffer_on_event(void* b, double now_ms, int* head, int* count, int* target);
int psb_buffer_set_ring(void* b, int ring_slots);   /* physical ring slots >= max_size (default: max_size) */
long long psb_buffer_inserted(void* b);   /* rows ever inserted == numTuplesSeen of the worker log */
/* Checkpoint of a buffer's policy state: the inter-arrival window and the ring cursors.  state: double[window + 8];
 * get returns the number of doubles written (-1: capacity too small), set returns 0 or PSB_E_INVALID. */
int psb_buffer_get_state(void* b, double* state, int capacity);
int psb_buffer_set_state(void* b, const double* state, int n);
/* n events at once: fills worker_slot[n][2] and updates win[num_buffers][2] for the touched workers */
int psb_buffers_on_batch(void** buffers, int num_buffers, const int* workers, const double* now_ms, int n,
                         int* worker_slot, int* win);
/* the same, and optionally change[num_buffers][4] = {head before, count before, rows added, count after} per worker
 * (the input of psb_window_stats_update) and *max_rows_out = the most rows any worker received from the batch */
int psb_buffers_on_batch2(void** buffers, int num_buffers, const int* workers, const double* now_ms, int n,
                          int* worker_slot, int* win, int* change, int* max_rows_out);

/* Streaming producer thread (CsvProducer.java:36-87 + INPUT_DATA): tokenises the CSV on a host thread,
 * keeps the rows of partitions [worker_begin, worker_begin + worker_count), converts them to dense fp16
 * and fills a ring of `num_batches` caller-provided PINNED batches (h_rows_f16[num_batches][batch_rows][F],
 * h_labels / h_workers / h_times [num_batches][batch_rows]); -p pacing as in the reference.
 * poll: 1 = batch ready (*batch, *n) and handed to the caller until psb_producer_release (several may be held,
 * released in poll order), 0 = nothing; *eof = 1 once the file is exhausted and every batch was polled, 2 when it
 * could not be opened, 3 when it is a damaged rows file.
 * `path` may also name a PSBROWS1 rows file (tools/csv_to_rows.py: the same rows converted once into the palette
 * wire format; `has_header` is ignored): nothing is tokenised, a kept row is two copies into the batch. */
void* psb_producer_start(const char* path, int has_header, int num_workers_total, int worker_begin,
                         int worker_count, int num_features, int p_ms, int batch_rows, int num_batches,
                         void* h_rows_f16, int* h_labels, int* h_workers, double* h_times, int flush_ms);
/* the same, and every published batch is also bitmap-packed by the producer thread (psb_pack_bitmap_dense) into
 * h_bitmap [num_batches][batch_rows][F/32], h_row_off [num_batches][batch_rows + 1] (row_off[n] = number of
 * values), h_values_f16 [num_batches][batch_rows * F] — ready for psb_ring_store_bitmap after the H2D. */
void* psb_producer_start_packed(const char* path, int has_header, int num_workers_total, int worker_begin,
                                int worker_count, int num_features, int p_ms, int batch_rows, int num_batches,
                                void* h_rows_f16, int* h_labels, int* h_workers, double* h_times, int flush_ms,
                                uint32_t* h_bitmap, int* h_row_off, void* h_values_f16);
/* the same with palette packing (psb_pack_palette_dense): h_payload [num_batches][payload_stride] bytes,
 * payload_stride >= batch_rows * (2 F + 2); h_row_off [num_batches][batch_rows + 1] are byte offsets. */
void* psb_producer_start_palette(const char* path, int has_header, int num_workers_total, int worker_begin,
                                 int worker_count, int num_features, int p_ms, int batch_rows, int num_batches,
                                 void* h_rows_f16, int* h_labels, int* h_workers, double* h_times, int flush_ms,
                                 uint32_t* h_bitmap, int* h_row_off, uint8_t* h_payload, long long payload_stride);
int psb_producer_poll(void* h, int* batch, int* n, int* eof);
void psb_producer_release(void* h, int batch);
long long psb_producer_rows_sent(void* h);
void psb_producer_stop(void* h);

/* L2 fetch-granularity hint for random 32-byte row traffic (sparse path); returns the value in effect */
int psb_set_l2_fetch_granularity(int bytes);

/* Peer memory: raw cudaMalloc (IPC-exportable, unlike a caching-allocator block) and IPC export /
 * import of a device allocation (transport set-up; replaces ServerApp.createTopics, ServerApp.java:31-42) */
int psb_malloc(void** d_ptr_out, size_t bytes);
int psb_free(void* d_ptr);
int psb_ipc_export(void* d_ptr, unsigned char handle_out[64]);
int psb_ipc_import(const unsigned char handle[64], void** d_ptr_out);
int psb_ipc_close(void* d_ptr);
int psb_enable_peer_access(int num_devices);

#ifdef __cplusplus
}
#endif
#endif
