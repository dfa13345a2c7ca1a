// Streaming producer thread: CSV text or a PSBROWS1 rows file -> a ring of pinned batches.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda_fp16.h>

#include "../include/psb200.h"
#include "host_csv.hpp"

void psb_set_error(const char* fmt, ...);
using psb_host::parse_line;

// ------------------------------------------------------------------ streaming producer thread
// Replaces CsvProducer (CsvProducer.java:36-87) + the INPUT_DATA topic: a host thread tokenises the CSV,
// keeps the rows of this rank's partitions (row i -> worker i % numWorkers, :61), converts them to dense
// fp16 and fills a ring of PINNED batches that the engine ships with cudaMemcpyAsync on its ingestion
// stream.  Pacing is the reference's: the first numWorkers*128 rows unthrottled (:73), then a 1000 ms
// sleep every 1000/p rows (:75-82); p > 1000 sleeps p ms per row and p == 0 never sleeps (quirk Q7).

extern "C" long long psb_pack_palette_dense(const uint16_t* rows, long long n_rows, int F, uint32_t* out_bitmap,
                                            int* out_row_off, uint8_t* out_payload, long long payload_capacity);
extern "C" long long psb_pack_bitmap_dense(const uint16_t* rows, long long n_rows, int F, uint32_t* out_bitmap,
                                           int* out_row_off, uint16_t* out_values);
namespace {
struct Producer {
    std::string path;
    int has_header, num_workers, wbegin, wcount, F, p_ms, batch_rows, num_batches, flush_ms, parse_threads = 1;
    uint16_t* rows;
    int* labels;
    int* workers;
    double* times;
    uint32_t* pk_bitmap = nullptr;            // optional: every published batch is also bitmap-packed
    int* pk_row_off = nullptr;                //   [num_batches][batch_rows + 1]
    uint16_t* pk_values = nullptr;            //   [num_batches][batch_rows * F]
    uint8_t* pk_payload = nullptr;            // palette mode instead of pk_values: [num_batches][pk_stride] bytes
    long long pk_stride = 0;
    std::vector<std::atomic<int>> ready;      // -1 free, >=0 rows published
    std::atomic<long long> rows_sent{0};      // global rows "sent" (all partitions)
    std::atomic<int> eof{0}, stop{0};
    std::thread th;
    int read_cursor = 0;                      // consumer side: next batch to poll
    int held = 0;                             // consumer side: batches polled and not yet released
    Producer(int nb) : ready(nb) {
        for (auto& r : ready) r.store(-1);
    }
};

// A kept line waiting to be tokenised into row `o` of the ring arrays.
struct ParseJob {
    const char* lp;
    const char* le;
    size_t o;
};

inline void parse_job(Producer* P, const ParseJob& j) {
    uint16_t* dst = P->rows + j.o * (size_t)P->F;
    memset(dst, 0, (size_t)P->F * 2);
    int nc = 0;
    const int F = P->F;
    P->labels[j.o] = parse_line(j.lp, j.le, &nc, [&](int c, float v) {
        if (c < F) {
            const __half h = __float2half(v);
            memcpy(dst + c, &h, 2);
        }
    });
}

// Helper threads of one producer: tokenising text is the cost of a kept row (~3 us), cutting lines is not (~0.6 us),
// so the reading thread cuts the lines of a batch and the pool tokenises them side by side, each line straight into
// its row of the batch.  run() returns when every job is done; the reading thread takes part.  A helper polls for
// the next batch for 3 ms (longer than the reading thread needs to cut a batch or to read the next 4 MB block)
// before it sleeps on the condition variable, and the reading thread polls for the helpers to finish: a batch is
// ~1 ms of work, and a thread woken through the kernel starts late and often on the waker's core — sleeping
// between batches gave 1.4x from four threads where polling gives 3.8x.  Paced streams never reach the pool
// (fewer than 64 lines at a time are tokenised by the reading thread), so its helpers sleep.
struct ParsePool {
    Producer* P;
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv_start;
    const ParseJob* jobs = nullptr;
    size_t n = 0;
    std::atomic<size_t> next{0}, busy{0};
    std::atomic<unsigned long long> gen{0};
    std::atomic<bool> quit{false};
    ParsePool(Producer* P_, int helpers) : P(P_) {
        for (int i = 0; i < helpers; ++i) th.emplace_back([this] { worker(); });
    }
    ~ParsePool() {
        {
            std::lock_guard<std::mutex> l(m);
            quit.store(true);
        }
        cv_start.notify_all();
        for (auto& t : th) t.join();
    }
    void drain() {
        for (;;) {
            const size_t i = next.fetch_add(16);
            if (i >= n) return;
            for (size_t j = i; j < std::min(i + 16, n); ++j) parse_job(P, jobs[j]);
        }
    }
    void worker() {
        unsigned long long seen = 0;
        for (;;) {
            const auto t0 = std::chrono::steady_clock::now();
            for (int spins = 0; gen.load(std::memory_order_acquire) == seen && !quit.load(std::memory_order_relaxed); ++spins) {
                if ((spins & 255) == 255 && std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(3000)) {
                    std::unique_lock<std::mutex> l(m);
                    cv_start.wait(l, [&] { return quit.load() || gen.load() != seen; });
                    break;
                }
            }
            if (quit.load()) return;
            seen = gen.load(std::memory_order_acquire);      // jobs / n / next were written before gen moved
            drain();
            busy.fetch_sub(1, std::memory_order_release);
        }
    }
    void run(const ParseJob* j, size_t count) {
        if (th.empty() || count < 64) {
            for (size_t i = 0; i < count; ++i) parse_job(P, j[i]);
            return;
        }
        {
            std::lock_guard<std::mutex> l(m);                // (the lock orders gen against a helper about to sleep)
            jobs = j, n = count;
            next.store(0);
            busy.store(th.size());
            gen.fetch_add(1, std::memory_order_release);
        }
        cv_start.notify_all();
        drain();
        while (busy.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    }
};

// A PSBROWS1 file (tools/csv_to_rows.py): the rows of a CSV converted once into the palette wire format, mapped
// read-only.  64-byte header {magic[8], u32 F, u32 version, u64 rows, u64 payload bytes}, then — each section on a
// 64-byte boundary — u64 offsets[rows + 1] into the payload, i32 labels[rows], u32 bitmap[rows][F / 32], payload.
struct RowsFile {
    const uint8_t* base = nullptr;
    size_t size = 0;
    uint32_t F = 0;
    uint64_t n = 0;
    const uint64_t* off = nullptr;
    const int32_t* labels = nullptr;
    const uint32_t* bitmap = nullptr;
    const uint8_t* payload = nullptr;
    static size_t up64(size_t x) { return (x + 63) & ~(size_t)63; }
    bool open(FILE* f) {
        fseek(f, 0, SEEK_END);
        const long sz = ftell(f);
        if (sz < 64) return false;
        void* m = mmap(nullptr, (size_t)sz, PROT_READ, MAP_PRIVATE, fileno(f), 0);
        if (m == MAP_FAILED) return false;
        base = (const uint8_t*)m, size = (size_t)sz;
        uint32_t version;
        uint64_t payload_bytes;
        memcpy(&F, base + 8, 4), memcpy(&version, base + 12, 4), memcpy(&n, base + 16, 8), memcpy(&payload_bytes, base + 24, 8);
        if (version != 1 || F == 0 || F % 32 != 0 || n > ((uint64_t)1 << 40)) return false;
        size_t at = 64;
        off = (const uint64_t*)(base + at), at = up64(at + (size_t)(n + 1) * 8);
        labels = (const int32_t*)(base + at), at = up64(at + (size_t)n * 4);
        bitmap = (const uint32_t*)(base + at), at = up64(at + (size_t)n * (F / 32) * 4);
        payload = base + at;
        return at + payload_bytes <= size && (n == 0 || at <= size) && off[n] == payload_bytes;
    }
    // the payload of row i is what its bitmap says it must be (a damaged file must not make the store kernel read astray)
    bool row_ok(uint64_t i) const {
        if (off[i + 1] < off[i] || off[i + 1] - off[i] < 2 || off[i + 1] > off[n]) return false;
        const uint64_t len = off[i + 1] - off[i];
        int nnz = 0;
        for (uint32_t w = 0; w < F / 32; ++w) nnz += __builtin_popcount(bitmap[i * (F / 32) + w]);
        uint16_t npal;
        memcpy(&npal, payload + off[i], 2);
        if (npal > 16 || (off[i] & 1)) return false;
        const uint64_t want = npal ? 2 + 2 * (uint64_t)npal + ((((uint64_t)nnz + 1) / 2 + 1) & ~(uint64_t)1) : 2 + 2 * (uint64_t)nnz;
        return len == want;
    }
    void close() {
        if (base) munmap((void*)base, size);
        base = nullptr;
    }
};

// Expands row i of a rows file into dense fp16 (consumers that did not ask for palette batches).
inline void rows_file_expand(const RowsFile& R, uint64_t i, uint16_t* dst) {
    memset(dst, 0, (size_t)R.F * 2);
    const uint8_t* p = R.payload + R.off[i];
    uint16_t npal;
    memcpy(&npal, p, 2);
    const uint8_t* codes = p + 2 + 2 * npal;
    int k = 0;
    for (uint32_t w = 0; w < R.F / 32; ++w)
        for (uint32_t m = R.bitmap[i * (R.F / 32) + w]; m; m &= m - 1, ++k) {
            uint16_t v;
            if (npal) {
                const int c = (codes[k >> 1] >> ((k & 1) * 4)) & 15;
                memcpy(&v, p + 2 + 2 * c, 2);
            } else {
                memcpy(&v, p + 2 + 2 * k, 2);
            }
            dst[w * 32 + __builtin_ctz(m)] = v;
        }
}

void producer_main(Producer* P) {
    FILE* f = fopen(P->path.c_str(), "rb");
    if (!f) {
        P->eof.store(2);
        return;
    }
    const auto t0 = std::chrono::steady_clock::now();
    auto now_ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    long long row_count = 0;
    const long long warm = (long long)P->num_workers * 128;
    int batch = 0, fill = 0;
    double batch_t0 = 0.0;
    bool prepacked = false;                    // rows file + palette consumer: the batch's packed form is written directly
    long long pre_bytes = 0;                   //   payload bytes of the batch so far
    ParsePool pool(P, P->parse_threads - 1);
    std::vector<ParseJob> jobs;                // lines of the batch being filled, cut and not yet tokenised
    jobs.reserve((size_t)P->batch_rows);
    double ms_parse = 0.0, ms_pack = 0.0, ms_wait = 0.0;       // where the thread's time went (PSB_PRODUCER_TRACE)
    auto flush_jobs = [&] {                    // before the text they point into moves, and before a batch is published
        if (jobs.empty()) return;
        const double t = now_ms();
        pool.run(jobs.data(), jobs.size());
        jobs.clear();
        ms_parse += now_ms() - t;
    };
    auto publish = [&] {
        flush_jobs();
        if (fill == 0) return;
        const double t_pack = now_ms();
        if (prepacked)
            P->pk_row_off[(size_t)batch * (P->batch_rows + 1) + fill] = (int)pre_bytes;
        else if (P->pk_bitmap && P->pk_payload)
            psb_pack_palette_dense(P->rows + (size_t)batch * P->batch_rows * P->F, fill, P->F,
                                   P->pk_bitmap + (size_t)batch * P->batch_rows * (P->F / 32),
                                   P->pk_row_off + (size_t)batch * (P->batch_rows + 1),
                                   P->pk_payload + (size_t)batch * P->pk_stride, P->pk_stride);
        else if (P->pk_bitmap)
            psb_pack_bitmap_dense(P->rows + (size_t)batch * P->batch_rows * P->F, fill, P->F,
                                  P->pk_bitmap + (size_t)batch * P->batch_rows * (P->F / 32),
                                  P->pk_row_off + (size_t)batch * (P->batch_rows + 1),
                                  P->pk_values + (size_t)batch * P->batch_rows * P->F);
        ms_pack += now_ms() - t_pack;
        P->ready[batch].store(fill, std::memory_order_release);
        batch = (batch + 1) % P->num_batches, fill = 0, pre_bytes = 0;
    };
    auto wait_free = [&]() -> bool {
        if (P->ready[batch].load(std::memory_order_acquire) == -1) return true;
        const double t = now_ms();
        while (P->ready[batch].load(std::memory_order_acquire) != -1) {
            if (P->stop.load()) return false;
            std::this_thread::sleep_for(std::chrono::microseconds(100));
        }
        ms_wait += now_ms() - t;
        return true;
    };
    // Row `row_count` of the stream goes to partition row_count % numWorkers (CsvProducer.java:61).  claim(): the slot
    // of the ring arrays for it when the partition is this rank's (-1: not ours, -2: stopped while waiting for a free
    // batch), with worker and arrival time filled in; the caller writes the row, then row_done(), then next_row().
    double t_row = -1.0;                       // one clock read per kept row; every 64th row of other partitions
    auto claim = [&]() -> long long {
        const int part = (int)(row_count % P->num_workers);
        t_row = -1.0;
        if (part < P->wbegin || part >= P->wbegin + P->wcount) {
            if ((row_count & 63) == 0) t_row = now_ms();
            return -1;
        }
        if (fill == 0) {
            if (!wait_free()) return -2;
            batch_t0 = now_ms();
        }
        const size_t o = (size_t)batch * P->batch_rows + fill;
        t_row = now_ms();
        P->workers[o] = part - P->wbegin, P->times[o] = t_row;
        return (long long)o;
    };
    auto row_done = [&] {
        if (++fill == P->batch_rows) publish();
    };
    auto next_row = [&] {                      // flush timer and -p pacing (CsvProducer.java:73-82, quirk Q7)
        ++row_count;
        P->rows_sent.store(row_count, std::memory_order_relaxed);
        if (fill && t_row >= 0.0 && t_row - batch_t0 > P->flush_ms) publish();
        if (P->p_ms > 0 && row_count >= warm) {
            int sleep_ms = 0;
            if (P->p_ms <= 1000) {
                if (row_count % (1000 / P->p_ms) == 0) sleep_ms = 1000;
            } else {
                sleep_ms = P->p_ms;
            }
            if (sleep_ms) {
                publish();
                for (int s = 0; s < sleep_ms && !P->stop.load(); s += 10)
                    std::this_thread::sleep_for(std::chrono::milliseconds(std::min(10, sleep_ms - s)));
            }
        }
    };
    auto finish = [&](int eof_code) {
        publish();
        fclose(f);
        if (getenv("PSB_PRODUCER_TRACE"))
            fprintf(stderr, "psb producer: %lld rows read in %.1f ms with %d tokenising thread(s): tokenise %.1f ms, pack %.1f ms, "
                            "waiting for a free batch %.1f ms\n", row_count, now_ms(), P->parse_threads, ms_parse, ms_pack, ms_wait);
        P->eof.store(eof_code);
    };

    char magic[8] = {0};
    const size_t got_magic = fread(magic, 1, 8, f);
    if (got_magic == 8 && memcmp(magic, "PSBROWS1", 8) == 0) {
        // ---- converted rows: nothing to tokenise; a kept row is two copies (bitmap, payload) or one expansion
        RowsFile R;
        if (!R.open(f) || (int)R.F != P->F) {
            R.close();
            finish(3);                         // damaged, or written for another feature count
            return;
        }
        prepacked = P->pk_bitmap && P->pk_payload;
        const size_t words = R.F / 32;
        bool bad = false;
        for (uint64_t i = 0; i < R.n && !P->stop.load(std::memory_order_relaxed); ++i) {
            const long long o = claim();
            if (o == -2) break;
            if (o >= 0) {
                if (!R.row_ok(i)) {
                    bad = true;
                    break;
                }
                P->labels[o] = R.labels[i];
                if (prepacked) {
                    const size_t len = (size_t)(R.off[i + 1] - R.off[i]);
                    memcpy(P->pk_bitmap + (size_t)o * words, R.bitmap + i * words, words * 4);
                    P->pk_row_off[(size_t)batch * (P->batch_rows + 1) + fill] = (int)pre_bytes;
                    memcpy(P->pk_payload + (size_t)batch * P->pk_stride + pre_bytes, R.payload + R.off[i], len);
                    pre_bytes += (long long)len;
                } else {
                    rows_file_expand(R, i, P->rows + (size_t)o * P->F);
                }
                row_done();
            }
            next_row();
        }
        R.close();
        finish(bad ? 3 : 1);
        return;
    }
    fseek(f, 0, SEEK_SET);

    // ---- CSV text.  The file is read in 4 MB blocks and lines are cut with memchr: a row of another rank's
    // partition costs one scan for its newline, nothing is copied per line.  A line longer than the block grows it.
    std::vector<char> buf((size_t)4 << 20);
    size_t have = 0, pos = 0;                  // buf[pos, have) is unread
    bool file_done = false;
    // next line without its terminator in [*lp, *le); false at the end of the file
    auto next_line = [&](const char** lp, const char** le) -> bool {
        for (;;) {
            const char* nl = have > pos ? (const char*)memchr(buf.data() + pos, '\n', have - pos) : nullptr;
            if (nl) {
                *lp = buf.data() + pos, *le = nl;
                pos = (size_t)(nl - buf.data()) + 1;
                return true;
            }
            if (file_done) {
                if (have == pos) return false;
                *lp = buf.data() + pos, *le = buf.data() + have;      // last line without a newline
                pos = have;
                return true;
            }
            flush_jobs();
            if (pos) {
                memmove(buf.data(), buf.data() + pos, have - pos);
                have -= pos, pos = 0;
            }
            if (have == buf.size()) buf.resize(buf.size() * 2);
            const size_t got = fread(buf.data() + have, 1, buf.size() - have, f);
            have += got;
            if (got == 0) file_done = true;
        }
    };
    bool first = true;
    const char *lp, *le;
    while (!P->stop.load(std::memory_order_relaxed) && next_line(&lp, &le)) {
        while (le > lp && (le[-1] == '\r' || le[-1] == ' ')) --le;
        if (first && P->has_header) {
            first = false;
            continue;
        }
        first = false;
        if (le == lp) continue;
        const long long o = claim();
        if (o == -2) break;
        if (o >= 0) {
            jobs.push_back(ParseJob{lp, le, (size_t)o});
            if (P->parse_threads == 1 || (P->p_ms > 0 && row_count >= warm)) flush_jobs();     // paced rows go out one by one
            row_done();
        }
        next_row();
    }
    finish(1);
}
}  // namespace

extern "C" void* psb_producer_start(const char* path, int has_header, int num_workers_total, int worker_begin,
                                    int worker_count, int num_features, int p_ms, int batch_rows, int num_batches,
                                    void* h_rows_f16, int* h_labels, int* h_workers, double* h_times, int flush_ms) {
    return psb_producer_start_packed(path, has_header, num_workers_total, worker_begin, worker_count, num_features, p_ms,
                                     batch_rows, num_batches, h_rows_f16, h_labels, h_workers, h_times, flush_ms, nullptr,
                                     nullptr, nullptr);
}
static void* producer_start_impl(const char* path, int has_header, int num_workers_total, int worker_begin,
                                 int worker_count, int num_features, int p_ms, int batch_rows, int num_batches,
                                 void* h_rows_f16, int* h_labels, int* h_workers, double* h_times, int flush_ms,
                                 uint32_t* h_bitmap, int* h_row_off, void* h_values_f16, uint8_t* h_payload,
                                 long long payload_stride);
extern "C" void* psb_producer_start_packed(const char* path, int has_header, int num_workers_total, int worker_begin,
                                           int worker_count, int num_features, int p_ms, int batch_rows,
                                           int num_batches, void* h_rows_f16, int* h_labels, int* h_workers,
                                           double* h_times, int flush_ms, uint32_t* h_bitmap, int* h_row_off,
                                           void* h_values_f16) {
    return producer_start_impl(path, has_header, num_workers_total, worker_begin, worker_count, num_features, p_ms,
                               batch_rows, num_batches, h_rows_f16, h_labels, h_workers, h_times, flush_ms, h_bitmap,
                               h_row_off, h_values_f16, nullptr, 0);
}
extern "C" void* psb_producer_start_palette(const char* path, int has_header, int num_workers_total, int worker_begin,
                                            int worker_count, int num_features, int p_ms, int batch_rows,
                                            int num_batches, void* h_rows_f16, int* h_labels, int* h_workers,
                                            double* h_times, int flush_ms, uint32_t* h_bitmap, int* h_row_off,
                                            uint8_t* h_payload, long long payload_stride) {
    if (!h_bitmap || !h_row_off || !h_payload || payload_stride < (long long)batch_rows * (2 * num_features + 2)) {
        psb_set_error("psb_producer_start_palette: needs bitmap, offsets and a payload of >= batch_rows * (2 F + 2) bytes per batch");
        return nullptr;
    }
    return producer_start_impl(path, has_header, num_workers_total, worker_begin, worker_count, num_features, p_ms,
                               batch_rows, num_batches, h_rows_f16, h_labels, h_workers, h_times, flush_ms, h_bitmap,
                               h_row_off, h_payload, h_payload, payload_stride);
}
static void* producer_start_impl(const char* path, int has_header, int num_workers_total, int worker_begin,
                                 int worker_count, int num_features, int p_ms, int batch_rows, int num_batches,
                                 void* h_rows_f16, int* h_labels, int* h_workers, double* h_times, int flush_ms,
                                 uint32_t* h_bitmap, int* h_row_off, void* h_values_f16, uint8_t* h_payload,
                                 long long payload_stride) {
    if ((h_bitmap || h_row_off || h_values_f16) && !(h_bitmap && h_row_off && h_values_f16 && num_features % 32 == 0)) {
        psb_set_error("psb_producer_start_packed: packing needs all three buffers and num_features %% 32 == 0");
        return nullptr;
    }
    if (!path || num_workers_total < 1 || worker_count < 1 || batch_rows < 1 || num_batches < 2 || !h_rows_f16) {
        psb_set_error("psb_producer_start: bad arguments");
        return nullptr;
    }
    Producer* P = new Producer(num_batches);
    P->path = path, P->has_header = has_header, P->num_workers = num_workers_total, P->wbegin = worker_begin;
    P->wcount = worker_count, P->F = num_features, P->p_ms = p_ms, P->batch_rows = batch_rows;
    P->num_batches = num_batches, P->flush_ms = flush_ms > 0 ? flush_ms : 50;
    P->rows = (uint16_t*)h_rows_f16, P->labels = h_labels, P->workers = h_workers, P->times = h_times;
    P->pk_bitmap = h_bitmap, P->pk_row_off = h_row_off, P->pk_values = (uint16_t*)h_values_f16;
    P->pk_payload = h_payload, P->pk_stride = payload_stride;
    // threads that tokenise (the reading thread included): PSB_PRODUCER_THREADS, else a sixteenth of the cores (one
    // process per GPU on an 8-GPU node, each with its Python threads beside this one), at most 4
    int threads = (int)std::min(4u, std::max(1u, std::thread::hardware_concurrency() / 16u));
    if (const char* e = getenv("PSB_PRODUCER_THREADS")) threads = atoi(e);
    P->parse_threads = std::max(1, std::min(threads, 16));
    P->th = std::thread(producer_main, P);
    return P;
}
// 1: *batch holds *n rows and now belongs to the caller until it is released (several batches may be held at
// once, e.g. while their host->device copies are in flight; release them in the order they were polled).
// 0: nothing ready.  *eof: 1 when the file is done AND every batch was polled ("CSV Producer sent all
// records!"), 2 when the file could not be opened.
extern "C" int psb_producer_poll(void* h, int* batch, int* n, int* eof) {
    Producer* P = (Producer*)h;
    if (P->held < P->num_batches) {
        const int r = P->ready[P->read_cursor].load(std::memory_order_acquire);
        if (r >= 0) {
            *batch = P->read_cursor, *n = r, *eof = 0;
            P->read_cursor = (P->read_cursor + 1) % P->num_batches;
            ++P->held;
            return 1;
        }
    }
    const int e = P->eof.load();
    *eof = (e && P->held < P->num_batches && P->ready[P->read_cursor].load(std::memory_order_acquire) < 0) ? e : 0;
    return 0;
}
extern "C" void psb_producer_release(void* h, int batch) {
    Producer* P = (Producer*)h;
    P->ready[batch].store(-1, std::memory_order_release);
    if (P->held > 0) --P->held;
}
extern "C" long long psb_producer_rows_sent(void* h) { return ((Producer*)h)->rows_sent.load(); }
extern "C" void psb_producer_stop(void* h) {
    Producer* P = (Producer*)h;
    P->stop.store(1);
    if (P->th.joinable()) P->th.join();
    delete P;
}

