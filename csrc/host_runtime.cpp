// Host runtime in C++: CSV bulk loader, consistency tracker, dynamic-buffer policy.
// These are the native counterparts of psb200/data/csv_io.py, psb200/consistency/tracker.py and
// psb200/engine/buffer.py (which stay as the readable statement + test oracle).  The streaming producer thread is
// host_producer.cpp, the row packers are host_pack.cpp, the tokenizer both readers share is host_csv.hpp.
//
// Reference: CsvProducer.java:46-58 (split on ',', drop zeros, last column = label),
// MessageTracker.java:10-88 + ServerProcessor.java:95-134, WorkerSamplingProcessor.java:115-135.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../include/psb200.h"
#include "host_csv.hpp"

void psb_set_error(const char* fmt, ...);
using psb_host::parse_line;

// ------------------------------------------------------------------ CSV
namespace {
struct FileBuf {
    std::vector<char> data;
    bool load(const char* path) {
        FILE* f = fopen(path, "rb");
        if (!f) return false;
        fseek(f, 0, SEEK_END);
        long n = ftell(f);
        fseek(f, 0, SEEK_SET);
        data.resize((size_t)n + 1);
        size_t got = fread(data.data(), 1, (size_t)n, f);
        fclose(f);
        data[got] = '\n';
        data.resize(got + 1);
        return true;
    }
};

}  // namespace

extern "C" int psb_csv_count(const char* path, int has_header, long long* n_rows, long long* n_nnz, int* n_cols) {
    FileBuf fb;
    if (!fb.load(path)) {
        psb_set_error("cannot open %s", path);
        return PSB_E_INVALID;
    }
    const char* p = fb.data.data();
    const char* end = p + fb.data.size();
    long long rows = 0, nnz = 0;
    int cols = 0;
    bool first = true;
    while (p < end) {
        const char* e = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!e) e = end;
        const char* le = e;
        while (le > p && (le[-1] == '\r' || le[-1] == ' ')) --le;
        if (first && has_header) {
            first = false;
        } else if (le > p) {
            first = false;
            int nc = 0;
            parse_line(p, le, &nc, [&](int, float) { ++nnz; });
            if (nc - 1 > cols) cols = nc - 1;
            ++rows;
        }
        p = e + 1;
    }
    *n_rows = rows, *n_nnz = nnz, *n_cols = cols;
    return 0;
}

extern "C" int psb_csv_parse(const char* path, int has_header, long long* indptr, int* indices, float* values,
                             int* labels, long long n_rows, long long n_nnz) {
    FileBuf fb;
    if (!fb.load(path)) {
        psb_set_error("cannot open %s", path);
        return PSB_E_INVALID;
    }
    const char* p = fb.data.data();
    const char* end = p + fb.data.size();
    long long row = 0, k = 0;
    bool first = true;
    indptr[0] = 0;
    while (p < end && row < n_rows) {
        const char* e = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!e) e = end;
        const char* le = e;
        while (le > p && (le[-1] == '\r' || le[-1] == ' ')) --le;
        if (first && has_header) {
            first = false;
        } else if (le > p) {
            first = false;
            int nc = 0;
            bool overflow = false;
            labels[row] = parse_line(p, le, &nc, [&](int c, float v) {
                if (k < n_nnz) indices[k] = c, values[k] = v, ++k;
                else overflow = true;
            });
            if (overflow) {
                psb_set_error("psb_csv_parse: more non-zeros than counted");
                return PSB_E_INVALID;
            }
            indptr[++row] = k;
        }
        p = e + 1;
    }
    return 0;
}

// One pass, several threads: the file is cut into chunks at line ends, every chunk is tokenised into its own
// vectors, psb_csv_fill lays them out as one CSR.  (psb_csv_count + psb_csv_parse read and tokenise the file twice.)
namespace {
struct CsvChunk {
    std::vector<long long> row_end;            // non-zeros of the chunk up to and including each row
    std::vector<int> indices, labels;
    std::vector<float> values;
    int cols = 0;
};
struct CsvFile {
    std::vector<CsvChunk> chunks;
    long long rows = 0, nnz = 0;
    int cols = 0;
};
void csv_parse_chunk(const char* p, const char* end, CsvChunk* out) {
    out->indices.reserve((size_t)(end - p) / 32), out->values.reserve((size_t)(end - p) / 32);
    out->labels.reserve((size_t)(end - p) / 1024), out->row_end.reserve((size_t)(end - p) / 1024);
    while (p < end) {
        const char* e = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!e) e = end;
        const char* le = e;
        while (le > p && (le[-1] == '\r' || le[-1] == ' ')) --le;
        if (le > p) {
            int nc = 0;
            out->labels.push_back(parse_line(p, le, &nc, [&](int c, float v) {
                out->indices.push_back(c), out->values.push_back(v);
            }));
            out->row_end.push_back((long long)out->indices.size());
            if (nc - 1 > out->cols) out->cols = nc - 1;
        }
        p = e + 1;
    }
}
}  // namespace

extern "C" void* psb_csv_open(const char* path, int has_header, int threads) {
    // mapped, not read: the threads fault their own chunks in, nothing is copied or zero-filled first
    FILE* fp = path ? fopen(path, "rb") : nullptr;
    if (!fp) {
        psb_set_error("cannot open %s", path ? path : "(null)");
        return nullptr;
    }
    fseek(fp, 0, SEEK_END);
    const long size = ftell(fp);
    if (size <= 0) {
        fclose(fp);
        return new CsvFile();
    }
    void* map = mmap(nullptr, (size_t)size, PROT_READ, MAP_PRIVATE, fileno(fp), 0);
    fclose(fp);
    if (map == MAP_FAILED) {
        psb_set_error("cannot map %s", path);
        return nullptr;
    }
    struct Unmap {
        void* m;
        size_t n;
        ~Unmap() { munmap(m, n); }
    } unmap{map, (size_t)size};
    const char* p = (const char*)map;
    const char* end = p + size;
    if (has_header) {
        const char* e = (const char*)memchr(p, '\n', (size_t)(end - p));
        p = e ? e + 1 : end;
    }
    const size_t bytes = (size_t)(end - p);
    if (threads <= 0) threads = (int)std::min(8u, std::max(1u, std::thread::hardware_concurrency()));
    threads = (int)std::max<size_t>(1, std::min<size_t>((size_t)threads, bytes >> 20));      // >= 1 MB of text per thread
    std::vector<const char*> cut((size_t)threads + 1, end);
    cut[0] = p;
    for (int t = 1; t < threads; ++t) {
        const char* q = p + bytes * (size_t)t / (size_t)threads;
        if (q < cut[(size_t)t - 1]) q = cut[(size_t)t - 1];
        const char* e = q < end ? (const char*)memchr(q, '\n', (size_t)(end - q)) : nullptr;
        cut[(size_t)t] = e ? e + 1 : end;
    }
    CsvFile* f = new CsvFile();
    f->chunks.resize((size_t)threads);
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t)
        th.emplace_back(csv_parse_chunk, cut[(size_t)t], cut[(size_t)t + 1], &f->chunks[(size_t)t]);
    csv_parse_chunk(cut[0], cut[1], &f->chunks[0]);
    for (auto& t : th) t.join();
    for (const auto& c : f->chunks) {
        f->rows += (long long)c.labels.size(), f->nnz += (long long)c.indices.size();
        f->cols = std::max(f->cols, c.cols);
    }
    return f;
}
extern "C" int psb_csv_shape(void* h, long long* n_rows, long long* n_nnz, int* n_cols) {
    const CsvFile* f = (const CsvFile*)h;
    if (!f) return PSB_E_INVALID;
    *n_rows = f->rows, *n_nnz = f->nnz, *n_cols = f->cols;
    return 0;
}
// indptr[rows + 1], indices / values [nnz], labels [rows]
extern "C" int psb_csv_fill(void* h, long long* indptr, int* indices, float* values, int* labels) {
    const CsvFile* f = (const CsvFile*)h;
    if (!f || !indptr || !indices || !values || !labels) return PSB_E_INVALID;
    long long row = 0, k = 0;
    indptr[0] = 0;
    for (const auto& c : f->chunks) {
        for (size_t r = 0; r < c.labels.size(); ++r) indptr[row + 1 + (long long)r] = k + c.row_end[r];
        if (!c.labels.empty()) memcpy(labels + row, c.labels.data(), c.labels.size() * sizeof(int));
        if (!c.indices.empty()) {
            memcpy(indices + k, c.indices.data(), c.indices.size() * sizeof(int));
            memcpy(values + k, c.values.data(), c.values.size() * sizeof(float));
        }
        row += (long long)c.labels.size(), k += (long long)c.indices.size();
    }
    return 0;
}
extern "C" void psb_csv_close(void* h) { delete (CsvFile*)h; }

// ------------------------------------------------------------------ consistency tracker
namespace {
struct Tracker {
    int n, model;
    std::vector<int> clock;
    std::vector<char> sent;
    Tracker(int n_, int m) : n(n_), model(m), clock(n_, 0), sent(n_, 1) {}
    int min_clock() const { return *std::min_element(clock.begin(), clock.end()); }
};
}  // namespace

extern "C" void* psb_tracker_create(int num_workers, int consistency_model) {
    if (num_workers < 1 || consistency_model < -1) return nullptr;
    return new Tracker(num_workers, consistency_model);
}
extern "C" void psb_tracker_destroy(void* t) { delete (Tracker*)t; }

extern "C" int psb_tracker_on_gradient(void* tp, int worker, int clock, int* replies, int max_replies) {
    Tracker* t = (Tracker*)tp;
    if (!t || worker < 0 || worker >= t->n) {
        psb_set_error("tracker: bad worker %d", worker);
        return PSB_E_INVALID;
    }
    if (t->clock[worker] != clock) {
        psb_set_error("Expected value %d, actual value %d", t->clock[worker], clock);
        return PSB_E_INVALID;
    }
    t->clock[worker]++;
    t->sent[worker] = 0;
    int nrep = 0;
    auto reply = [&](int w, int c) {
        if (nrep < max_replies) replies[2 * nrep] = w, replies[2 * nrep + 1] = c;
        ++nrep;
        t->sent[w] = 1;
    };
    if (t->model == -1) {
        reply(worker, clock + 1);
    } else if (t->model == 0) {
        if (t->min_clock() >= clock + 1)
            for (int w = 0; w < t->n; ++w) reply(w, clock + 1);
    } else {
        const int lo = t->min_clock();
        for (int w = 0; w < t->n; ++w)
            if (!t->sent[w] && lo >= t->clock[w] - t->model) reply(w, t->clock[w]);
    }
    return nrep;
}
extern "C" int psb_tracker_clocks(void* tp, int* clocks) {
    Tracker* t = (Tracker*)tp;
    for (int w = 0; w < t->n; ++w) clocks[w] = t->clock[w];
    return t->n;
}

// ------------------------------------------------------------------ dynamic buffer policy
namespace {
struct Buffer {
    int min_size, max_size, window;
    int ring = 0;                        // physical slots of the device ring (>= max_size; head wraps here)
    float bc;
    std::vector<double> deltas;          // circular: the last `window` inter-arrival times
    int n_deltas = 0, pos = 0;
    double total = 0.0, last = -1.0;
    bool has_last = false;
    int head = 0, count = 0;
    long long inserted = 0;
};
}  // namespace

extern "C" void* psb_buffer_create(int min_size, int max_size, float bc, int window) {
    if (min_size < 1 || max_size < min_size) return nullptr;
    Buffer* b = new Buffer();
    b->min_size = min_size, b->max_size = max_size, b->bc = bc, b->window = window > 0 ? window : 500;
    b->ring = max_size;
    b->deltas.assign((size_t)b->window, 0.0);
    return b;
}
extern "C" void psb_buffer_destroy(void* b) { delete (Buffer*)b; }
// The device ring may hold more slots than the largest window (spare slots let the next batch be stored while
// a step still trains on the current window): head then wraps at ring_slots instead of max_size.
extern "C" int psb_buffer_set_ring(void* bp, int ring_slots) {
    Buffer* b = (Buffer*)bp;
    if (!b || ring_slots < b->max_size || b->inserted != 0) {
        psb_set_error("psb_buffer_set_ring: ring must hold at least max_size slots and be set before the first event");
        return PSB_E_INVALID;
    }
    b->ring = ring_slots;
    return 0;
}

// Returns the slot to write; head/count describe the window after the event.
static inline int buffer_on_event(Buffer* b, double now_ms, int* head, int* count, int* target) {
    if (!b->has_last) {
        b->has_last = true;
    } else {
        const double d = now_ms - b->last;
        if (b->n_deltas == b->window) b->total -= b->deltas[b->pos];
        else ++b->n_deltas;
        b->deltas[b->pos] = d;
        b->total += d;
        if (++b->pos == b->window) b->pos = 0;
    }
    b->last = now_ms;
    int calc;
    if (b->n_deltas == 0) {
        calc = (int)floor((double)b->bc * 60.0 + 0.5);      // no history yet: mean inter-arrival taken as 1000 ms
    } else if (!(b->total > 0.0)) {
        calc = b->max_size;                       // Q8: infinite rate clamps to max (reference: min)
    } else {
        const double v = (double)b->bc * (60000.0 / (b->total / (double)b->n_deltas));
        calc = v < 2.0e9 ? (int)(v + 0.5) : b->max_size;     // v >= 0 here: truncation == floor; NaN / inf -> max
    }
    const int tgt = std::max(b->min_size, std::min(b->max_size, calc));
    const int slot = b->head;
    if (++b->head == b->ring) b->head = 0;
    b->count = std::min(b->count + 1, tgt);
    b->inserted++;
    *head = b->head, *count = b->count, *target = tgt;
    return slot;
}
extern "C" int psb_buffer_on_event(void* bp, double now_ms, int* head, int* count, int* target) {
    return buffer_on_event((Buffer*)bp, now_ms, head, count, target);
}

extern "C" long long psb_buffer_inserted(void* bp) { return ((Buffer*)bp)->inserted; }

// state layout: [window, n_deltas, pos, total, last, has_last, head, count, inserted(lo 2^32 units), deltas...]
extern "C" int psb_buffer_get_state(void* bp, double* state, int capacity) {
    const Buffer* b = (const Buffer*)bp;
    if (!b || !state) return -1;
    const int n = 9 + b->window;
    if (capacity < n) return -1;
    state[0] = b->window, state[1] = b->n_deltas, state[2] = b->pos, state[3] = b->total, state[4] = b->last;
    state[5] = b->has_last ? 1.0 : 0.0, state[6] = b->head, state[7] = b->count, state[8] = (double)b->inserted;
    for (int i = 0; i < b->window; ++i) state[9 + i] = b->deltas[(size_t)i];
    return n;
}
extern "C" int psb_buffer_set_state(void* bp, const double* state, int n) {
    Buffer* b = (Buffer*)bp;
    if (!b || !state || n < 9 || (int)state[0] != b->window || n != 9 + b->window) {
        psb_set_error("psb_buffer_set_state: state does not fit this buffer (window %d)", b ? b->window : -1);
        return PSB_E_INVALID;
    }
    b->n_deltas = (int)state[1], b->pos = (int)state[2], b->total = state[3], b->last = state[4];
    b->has_last = state[5] != 0.0, b->head = (int)state[6], b->count = (int)state[7];
    b->inserted = (long long)state[8];
    if (b->head < 0 || b->head >= b->ring || b->count < 0 || b->count > b->max_size || b->pos < 0 ||
        b->pos >= b->window || b->n_deltas < 0 || b->n_deltas > b->window) {
        psb_set_error("psb_buffer_set_state: cursors out of range");
        return PSB_E_INVALID;
    }
    for (int i = 0; i < b->window; ++i) b->deltas[(size_t)i] = state[9 + i];
    return 0;
}

// Batch form used by the ingestion hot path: n events -> (worker, slot) table + the new (head, count)
// window of every worker that received a row.  `buffers` is an array of psb_buffer_create handles.
// (Splitting the workers of a batch over a small thread pool was measured: 0.33 -> 0.8-1.2 ms for 37,888 events —
//  the wake-ups cost more than the two divisions per event they spread.)
// change (optional, [num_buffers][4]): {head before, count before, rows added, count after} per worker — what
// psb_window_stats_update needs; max_rows_out (optional): the most rows any worker received.
extern "C" int psb_buffers_on_batch2(void** buffers, int num_buffers, const int* workers, const double* now_ms, int n,
                                     int* worker_slot, int* win, int* change, int* max_rows_out) {
    static thread_local std::vector<int> added_tls;
    added_tls.assign((size_t)num_buffers, 0);
    int* const added = added_tls.data();                 // (no thread-local lookup inside the loops)
    if (change)
        for (int w = 0; w < num_buffers; ++w) {
            const Buffer* b = (const Buffer*)buffers[w];
            change[4 * w] = b->head, change[4 * w + 1] = b->count, change[4 * w + 2] = 0, change[4 * w + 3] = b->count;
        }
    // The events are grouped by worker first (stable counting sort: a worker's events keep their order, and different
    // workers' buffers are independent), then every worker's events are processed in one tight loop with the
    // buffer's state in locals — the same arithmetic as buffer_on_event, ~3x faster than visiting the buffers in
    // arrival order (whose state the compiler has to reload after every store to the output tables).
    static thread_local std::vector<int> order_tls, start_tls;
    order_tls.resize((size_t)n), start_tls.assign((size_t)num_buffers + 1, 0);
    int* const order = order_tls.data();
    int* const start = start_tls.data();
    static thread_local std::vector<int> first_tls;
    first_tls.assign((size_t)num_buffers, -1);
    int* const first = first_tls.data();
    int runs = 0, distinct = 0;
    for (int i = 0, prev = -1; i < n; ++i) {
        const int w = workers[i];
        if (w < 0 || w >= num_buffers) {
            psb_set_error("psb_buffers_on_batch: worker %d out of range", w);
            return PSB_E_INVALID;
        }
        if (w != prev) {
            ++runs, prev = w;
            if (first[w] < 0) first[w] = i, ++distinct;
        }
        ++start[w + 1];
    }
    const bool grouped = runs == distinct;               // every worker's events are already contiguous: no permutation
    for (int w = 0; w < num_buffers; ++w) added[w] = start[w + 1], start[w + 1] += start[w];
    if (!grouped)
        for (int i = 0; i < n; ++i) order[start[workers[i]]++] = i;      // start[w] now = end of worker w's events
    else
        for (int w = 0; w < num_buffers; ++w) start[w] = start[w + 1];
    for (int w = 0, lo = 0; w < num_buffers; lo = start[w], ++w) {
        const int hi = start[w];
        if (hi == lo) continue;
        const int shift = grouped ? first[w] - lo : 0;   // grouped: event k of the sorted order is event k + shift
        Buffer* b = (Buffer*)buffers[w];
        const int min_size = b->min_size, max_size = b->max_size, ring = b->ring, window = b->window;
        const double bc = (double)b->bc;
        double* const deltas = b->deltas.data();
        int n_deltas = b->n_deltas, pos = b->pos, head = b->head, count = b->count;
        double total = b->total, last = b->last;
        bool has_last = b->has_last;
        for (int k = lo; k < hi; ++k) {
            const int j = grouped ? k + shift : order[k];
            const double now = now_ms[j];
            if (!has_last) {
                has_last = true;
            } else {
                const double d = now - last;
                if (n_deltas == window) total -= deltas[pos];
                else ++n_deltas;
                deltas[pos] = d;
                total += d;
                if (++pos == window) pos = 0;
            }
            last = now;
            int calc;
            if (n_deltas == 0) {
                calc = (int)floor(bc * 60.0 + 0.5);
            } else if (!(total > 0.0)) {
                calc = max_size;
            } else {
                const double v = bc * (60000.0 / (total / (double)n_deltas));
                calc = v < 2.0e9 ? (int)(v + 0.5) : max_size;
            }
            const int tgt = std::max(min_size, std::min(max_size, calc));
            worker_slot[2 * j] = w, worker_slot[2 * j + 1] = head;
            if (++head == ring) head = 0;
            count = std::min(count + 1, tgt);
        }
        b->n_deltas = n_deltas, b->pos = pos, b->head = head, b->count = count, b->total = total, b->last = last;
        b->has_last = has_last, b->inserted += hi - lo;
        win[2 * w] = head, win[2 * w + 1] = count;
    }
    // A batch that carries more rows for one worker than its ring holds writes some slots twice.  The store
    // kernels handle the rows of a batch in parallel, so only the LAST row of each (worker, slot) may be
    // written: earlier ones get slot -1 and are skipped on the device (the ring then holds the last
    // ring-size rows, as a row-by-row replay would leave it).
    int max_rows = 0;
    bool over = false;
    for (int w = 0; w < num_buffers; ++w) {
        const Buffer* b = (const Buffer*)buffers[w];
        max_rows = std::max(max_rows, added[w]);
        over = over || added[w] > b->ring;
        if (change) change[4 * w + 2] = added[w], change[4 * w + 3] = b->count;
    }
    if (over) {
        std::vector<int> kept((size_t)num_buffers, 0);
        for (int i = n - 1; i >= 0; --i) {
            const int w = worker_slot[2 * i];
            if (kept[w] >= ((Buffer*)buffers[w])->ring) worker_slot[2 * i + 1] = -1;
            else ++kept[w];
        }
    }
    if (max_rows_out) *max_rows_out = max_rows;
    return 0;
}
extern "C" int psb_buffers_on_batch(void** buffers, int num_buffers, const int* workers, const double* now_ms, int n,
                                    int* worker_slot, int* win) {
    return psb_buffers_on_batch2(buffers, num_buffers, workers, now_ms, n, worker_slot, win, nullptr, nullptr);
}

