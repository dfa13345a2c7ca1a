// Host runtime in C++: CSV tokenizer, consistency tracker, dynamic-buffer policy.
// These are the native counterparts of psb200/data/csv_io.py, psb200/consistency/tracker.py and
// psb200/engine/buffer.py (which stay as the readable statement + test oracle).
//
// Reference: CsvProducer.java:46-58 (split on ',', drop zeros, last column = label),
// MessageTracker.java:10-88 + ServerProcessor.java:95-134, WorkerSamplingProcessor.java:115-135.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../include/psb200.h"

void psb_set_error(const char* fmt, ...);

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

// Eight ASCII digits (first character in the lowest byte; bytes already reduced to 0..9) -> their integer.
inline uint64_t digits8(uint64_t x) {
    x = (x * 2561) >> 8;
    x = ((x & 0x00FF00FF00FF00FFULL) * 6553601) >> 16;
    return ((x & 0x0000FFFF0000FFFFULL) * 42949672960001ULL) >> 32;
}
// High bit of every byte of `t` (= text ^ "00000000") that is NOT a digit.
inline uint64_t non_digits(uint64_t t) {
    return (((t & 0x7f7f7f7f7f7f7f7fULL) + 0x7676767676767676ULL) | t) & 0x8080808080808080ULL;
}
// The common field "[-]d.ffffff," (one integer digit, up to 15 fraction digits): the fraction digits are found and
// converted eight at a time instead of one character per loop trip.  On success *v is the value, computed exactly
// as the general path computes it (integer mantissa / exact power of ten), and the ',' is returned; nullptr sends
// the field to the general path.
inline const char* fast_decimal(const char* s, const char* e, const double* pow10, float* v) {
    const bool neg = *s == '-';
    s += neg;
    if (e - s < 18) return nullptr;                    // two 8-byte loads behind "d."
    const unsigned d0 = (unsigned)(unsigned char)s[0] - (unsigned)'0';
    if (d0 > 9u || s[1] != '.') return nullptr;
    uint64_t a, b;
    memcpy(&a, s + 2, 8);
    a ^= 0x3030303030303030ULL;
    uint64_t nd = non_digits(a), mant;
    int n;
    if (nd) {
        n = __builtin_ctzll(nd) >> 3;                  // 0..7 fraction digits
        if (s[2 + n] != ',') return nullptr;
        mant = d0;
        if (n) mant = mant * (uint64_t)pow10[n] + digits8(a << (8 * (8 - n)));
    } else {
        memcpy(&b, s + 10, 8);
        b ^= 0x3030303030303030ULL;
        nd = non_digits(b);
        if (!nd) return nullptr;                       // 16 or more fraction digits
        const int n2 = __builtin_ctzll(nd) >> 3;       // 0..7 more
        if (s[10 + n2] != ',') return nullptr;
        mant = d0 * 100000000ULL + digits8(a);
        if (n2) mant = mant * (uint64_t)pow10[n2] + digits8(b << (8 * (8 - n2)));
        n = 8 + n2;
    }
    const float r = (float)((double)mant / pow10[n]);
    *v = neg ? -r : r;
    return s + 2 + n;
}

// Parses one line [p, e): calls emit(col, value) for every non-zero feature, returns the label.
// One pass per field: "0," and "0.0," (nearly every field of a hashed bag-of-words row) are skipped outright,
// plain decimals accumulate an integer mantissa that is divided by an exact power of ten, anything else
// (exponents, more than 18 digits, inf / nan) goes through strtof.
template <class Emit>
inline int parse_line(const char* p, const char* e, int* ncols, Emit emit) {
    static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                      1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    int col = 0;
    float last = 0.f;
    while (p < e) {
        if (*p == '0') {
            while (e - p >= 8) {               // runs of "0,0,0,0,": four absent features per 64-bit compare
                uint64_t x;
                memcpy(&x, p, 8);
                x ^= 0x2c302c302c302c30ULL;
                if (x) {                       // jump over the "0," pairs in front of the first byte that differs
                    const int k = (__builtin_ctzll(x) >> 3) & ~1;
                    p += k, col += k >> 1;
                    break;
                }
                p += 8, col += 4;
            }
            if (p >= e) break;
            if (*p != '0') continue;
            if (p + 1 < e && p[1] == ',') {
                p += 2, ++col;
                continue;
            }
            if (p + 3 < e && p[1] == '.' && p[2] == '0' && p[3] == ',') {
                p += 4, ++col;
                continue;
            }
        }
        float v;
        const char* s = fast_decimal(p, e, kPow10, &v);
        if (s) {                   // (always a feature column: a ',' follows)
            if (v != 0.f) emit(col, v);
            ++col;
            p = s + 1;
            continue;
        }
        s = p;
        bool neg = false, simple = true, frac = false;
        if (*s == '-' || *s == '+') neg = *s == '-', ++s;
        uint64_t mant = 0;
        int digits = 0, frac_digits = 0;
        for (; s < e; ++s) {
            const unsigned d = (unsigned)(unsigned char)*s - (unsigned)'0';
            if (d <= 9u) {
                mant = mant * 10u + d;
                ++digits, frac_digits += frac;
            } else if (*s == ',') {
                break;
            } else if (*s == '.' && !frac) {
                frac = true;
            } else if (*s == ' ' || *s == '\r') {
            } else {
                simple = false;
                break;
            }
        }
        if (simple && digits <= 18 && frac_digits <= 22) {
            v = (float)((double)mant / kPow10[frac_digits]);
            if (neg) v = -v;
        } else {
            const char* q = (const char*)memchr(p, ',', (size_t)(e - p));
            if (!q) q = e;
            v = strtof(std::string(p, q).c_str(), nullptr);
            s = q;
        }
        if (s < e) {               // a feature column
            if (v != 0.f) emit(col, v);
        } else {
            last = v;              // last column: the label
        }
        ++col;
        p = s + 1;
    }
    *ncols = col;
    return last >= -2147483648.f && last < 2147483648.f ? (int)last : 0;      // (nan / inf / huge labels: 0, not UB)
}
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

// ------------------------------------------------------------------ streaming producer thread
// Replaces CsvProducer (CsvProducer.java:36-87) + the INPUT_DATA topic: a host thread tokenises the CSV,
// keeps the rows of this rank's partitions (row i -> worker i % numWorkers, :61), converts them to dense
// fp16 and fills a ring of PINNED batches that the engine ships with cudaMemcpyAsync on its ingestion
// stream.  Pacing is the reference's: the first numWorkers*128 rows unthrottled (:73), then a 1000 ms
// sleep every 1000/p rows (:75-82); p > 1000 sleeps p ms per row and p == 0 never sleeps (quirk Q7).
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>

#include <cuda_fp16.h>
#include <sys/mman.h>
#if defined(__SSE2__) && !defined(PSB_NO_SIMD)
#include <emmintrin.h>
#endif

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

// ------------------------------------------------------------------ bitmap packing (host side of psb_ring_store_bitmap)
namespace {
// Bit b of the result: fp16 value p[b] (b < 32) is non-zero; +0.0 and -0.0 are both "absent".
#if defined(__SSE2__) && !defined(PSB_NO_SIMD)
inline uint32_t nonzero_mask32(const uint16_t* p) {
    const __m128i mag = _mm_set1_epi16(0x7fff), zero = _mm_setzero_si128();
    const __m128i a = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p)), mag), zero);
    const __m128i b = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p + 8)), mag), zero);
    const __m128i c = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p + 16)), mag), zero);
    const __m128i d = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p + 24)), mag), zero);
    const uint32_t lo = (uint32_t)_mm_movemask_epi8(_mm_packs_epi16(a, b));      // bit set = value is zero
    const uint32_t hi = (uint32_t)_mm_movemask_epi8(_mm_packs_epi16(c, d));
    return ~(lo | (hi << 16));
}
#else
// Four values per 64-bit word: rows are ~95 % zeros, so most words cost one load and one test.
inline uint32_t nonzero_mask32(const uint16_t* p) {
    uint32_t bits = 0;
    for (int q = 0; q < 8; ++q) {
        uint64_t x;
        memcpy(&x, p + 4 * q, 8);
        x &= 0x7fff7fff7fff7fffULL;
        if (x) {
            // lane <= 0x7fff, so lane + 0x7fff sets bit 15 of the lane iff lane != 0 and never carries over;
            // the multiply gathers bits 0 / 16 / 32 / 48 into bits 48..51
            const uint64_t t = ((x + 0x7fff7fff7fff7fffULL) >> 15) & 0x0001000100010001ULL;
            bits |= (uint32_t)(((t * 0x0001000200040008ULL) >> 48) & 0xFu) << (4 * q);
        }
    }
    return bits;
}
#endif

// Appends the values of src[0..32) whose bit is set in `bits` to out[k...] in order and returns the new k.  The first
// four go without a data-dependent branch (a 32-value group of a ~5 % dense row holds 0-3 non-zeros, so a loop over
// the set bits mispredicts its exit on nearly every group); a step with no bit left stores a dummy at out[k], which
// is inside the caller's buffer: such a group has at least 28 absent values, so k <= capacity - 28.
inline long long gather32(const uint16_t* src, uint32_t bits, uint16_t* out, long long k) {
    uint32_t m = bits;
    for (int i = 0; i < 4; ++i) {
        out[k] = src[__builtin_ctz(m | 0x80000000u)];
        k += m != 0;
        m &= m - 1;
    }
    for (; m; m &= m - 1) out[k++] = src[__builtin_ctz(m)];
    return k;
}
}  // namespace

extern "C" long long psb_pack_bitmap_dense(const uint16_t* rows, long long n_rows, int F, uint32_t* out_bitmap,
                                           int* out_row_off, uint16_t* out_values) {
    if (!rows || !out_bitmap || !out_row_off || !out_values || F % 32 != 0) {
        psb_set_error("psb_pack_bitmap_dense: null pointer or num_features not a multiple of 32");
        return -1;
    }
    const int words = F / 32;
    long long k = 0;
    for (long long r = 0; r < n_rows; ++r) {
        const uint16_t* row = rows + (size_t)r * F;
        out_row_off[r] = (int)k;
        for (int w = 0; w < words; ++w) {
            const uint32_t bits = nonzero_mask32(row + w * 32);
            k = gather32(row + w * 32, bits, out_values, k);
            out_bitmap[(size_t)r * words + w] = bits;
        }
    }
    out_row_off[n_rows] = (int)k;
    return k;
}

extern "C" long long psb_pack_bitmap_csr(const long long* indptr, const int* indices, const float* values,
                                         long long n_rows, int F, uint32_t* out_bitmap, int* out_row_off,
                                         uint16_t* out_values) {
    if (!indptr || !indices || !values || !out_bitmap || !out_row_off || !out_values || F % 32 != 0) {
        psb_set_error("psb_pack_bitmap_csr: null pointer or num_features not a multiple of 32");
        return -1;
    }
    const int words = F / 32;
    std::vector<uint16_t> dense((size_t)F, (uint16_t)0);
    std::vector<uint32_t> touched((size_t)words, 0u);
    long long k = 0;
    for (long long r = 0; r < n_rows; ++r) {
        // column order of the source is not trusted (and a column may repeat: the last entry wins): go through a
        // dense scratch row, of which only the 32-value groups this row touched are read and cleared again
        for (long long e = indptr[r]; e < indptr[r + 1]; ++e) {
            const int c = indices[e];
            if (c < 0 || c >= F) continue;
            const __half h = __float2half(values[e]);
            memcpy(&dense[c], &h, 2);
            touched[c >> 5] = 1u;
        }
        out_row_off[r] = (int)k;
        for (int w = 0; w < words; ++w) {
            uint32_t bits = 0;
            if (touched[w]) {
                bits = nonzero_mask32(dense.data() + w * 32);
                k = gather32(dense.data() + w * 32, bits, out_values, k);
                memset(dense.data() + w * 32, 0, 64);
                touched[w] = 0u;
            }
            out_bitmap[(size_t)r * words + w] = bits;
        }
    }
    out_row_off[n_rows] = (int)k;
    return k;
}

// Bitmap + per-row palette: hashed bag-of-words rows hold a handful of distinct values (count / norm, signed), so
// a row ships as its presence bitmap, the distinct fp16 values it uses (<= 16) and one 4-bit code per non-zero —
// bit-exact, a quarter of the value bytes.  A row with more than 16 distinct values ships its fp16 values raw.
// Payload of a row (starts at the even byte offset out_row_off[r]):
//   u16 npal;  npal > 0: u16 palette[npal], then ceil(nnz / 2) code bytes (low nibble first), padded to even
//              npal == 0: u16 values[nnz]
extern "C" long long psb_pack_palette_dense(const uint16_t* rows, long long n_rows, int F, uint32_t* out_bitmap,
                                            int* out_row_off, uint8_t* out_payload, long long payload_capacity) {
    if (!rows || !out_bitmap || !out_row_off || !out_payload || F % 32 != 0) {
        psb_set_error("psb_pack_palette_dense: null pointer or num_features not a multiple of 32");
        return -1;
    }
    const int words = F / 32;
    long long off = 0;
    std::vector<uint8_t> codes((size_t)F + 1);
    std::vector<uint16_t> vals((size_t)F);
    for (long long r = 0; r < n_rows; ++r) {
        const uint16_t* row = rows + (size_t)r * F;
        if (off + 2 + 2LL * F > payload_capacity) {
            psb_set_error("psb_pack_palette_dense: payload buffer too small (row %lld)", r);
            return -1;
        }
        out_row_off[r] = (int)off;
        uint32_t* bitmap = out_bitmap + (size_t)r * words;
        long long nz = 0;                                   // the row's non-zeros, in feature order
        for (int w = 0; w < words; ++w) {
            const uint32_t bits = nonzero_mask32(row + w * 32);
            bitmap[w] = bits;
            nz = gather32(row + w * 32, bits, vals.data(), nz);
        }
        const int nnz = (int)nz;
        alignas(16) uint16_t pal[16] = {0};                 // 0 never equals a non-zero value: free entries do not match
        int npal = 0;
        bool overflow = false;
#if defined(__SSE2__) && !defined(PSB_NO_SIMD)
        __m128i p0 = _mm_setzero_si128(), p1 = _mm_setzero_si128();
        for (int i = 0; i < nnz; ++i) {
            const __m128i h = _mm_set1_epi16((short)vals[i]);
            const int hit = _mm_movemask_epi8(_mm_packs_epi16(_mm_cmpeq_epi16(p0, h), _mm_cmpeq_epi16(p1, h)));
            int c;
            if (hit) {
                c = __builtin_ctz((unsigned)hit);
            } else {
                if (npal == 16) {
                    overflow = true;
                    break;
                }
                pal[npal] = vals[i];
                c = npal++;
                p0 = _mm_load_si128((const __m128i*)pal), p1 = _mm_load_si128((const __m128i*)(pal + 8));
            }
            codes[i] = (uint8_t)c;
        }
#else
        for (int i = 0; i < nnz && !overflow; ++i) {
            const uint16_t h = vals[i];
            int c = 0;
            while (c < npal && pal[c] != h) ++c;
            if (c == npal) {
                if (npal == 16) overflow = true;
                else pal[npal++] = h;
            }
            codes[i] = (uint8_t)c;
        }
#endif
        uint8_t* p = out_payload + off;
        if (overflow || nnz == 0) {
            const uint16_t zero = 0;
            memcpy(p, &zero, 2);
            if (nnz) memcpy(p + 2, vals.data(), (size_t)nnz * 2);
            off += 2 + 2LL * nnz;
        } else {
            const uint16_t np = (uint16_t)npal;
            memcpy(p, &np, 2);
            memcpy(p + 2, pal, (size_t)npal * 2);
            uint8_t* c = p + 2 + 2 * npal;
            const int nb = (nnz + 1) / 2;
            codes[nnz] = 0;                                   // high nibble of the last byte when the count is odd
            for (int i = 0; i < nb; ++i) c[i] = (uint8_t)(codes[2 * i] | (codes[2 * i + 1] << 4));
            off += 2 + 2LL * npal + ((nb + 1) & ~1);
        }
    }
    out_row_off[n_rows] = (int)off;
    return off;
}
