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
#include <deque>
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

// Parses one line [p, e): calls emit(col, value) for every non-zero feature, returns the label.
// Fast path for plain decimals; strtof for anything with an exponent.
template <class Emit>
inline int parse_line(const char* p, const char* e, int* ncols, Emit emit) {
    int col = 0;
    float last = 0.f;
    while (p < e) {
        const char* q = p;
        while (q < e && *q != ',') ++q;
        float v;
        {
            const char* s = p;
            bool neg = false, simple = true;
            if (s < q && (*s == '-' || *s == '+')) neg = *s == '-', ++s;
            double acc = 0.0, scale = 1.0;
            bool frac = false;
            for (; s < q; ++s) {
                const char c = *s;
                if (c >= '0' && c <= '9') {
                    acc = acc * 10.0 + (c - '0');
                    if (frac) scale *= 10.0;
                } else if (c == '.' && !frac) {
                    frac = true;
                } else if (c == ' ' || c == '\r') {
                } else {
                    simple = false;
                    break;
                }
            }
            if (simple) {
                v = (float)(acc / scale);
                if (neg) v = -v;
            } else {
                v = strtof(std::string(p, q).c_str(), nullptr);
            }
        }
        if (q < e) {               // a feature column
            if (v != 0.f) emit(col, v);
        } else {
            last = v;              // last column: the label
        }
        ++col;
        p = q + 1;
    }
    *ncols = col;
    return (int)last;
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
    float bc;
    std::deque<double> deltas;
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
    return b;
}
extern "C" void psb_buffer_destroy(void* b) { delete (Buffer*)b; }

// Returns the slot to write; head/count describe the window after the event.
extern "C" int psb_buffer_on_event(void* bp, double now_ms, int* head, int* count, int* target) {
    Buffer* b = (Buffer*)bp;
    if (!b->has_last) {
        b->has_last = true;
    } else {
        const double d = now_ms - b->last;
        b->deltas.push_back(d);
        b->total += d;
        if ((int)b->deltas.size() > b->window) {
            b->total -= b->deltas.front();
            b->deltas.pop_front();
        }
    }
    b->last = now_ms;
    const double mean = b->deltas.empty() ? 1000.0 : b->total / (double)b->deltas.size();
    int calc;
    if (mean <= 0.0) {
        calc = b->max_size;                       // Q8: infinite rate clamps to max (reference: min)
    } else {
        const double v = (double)b->bc * (60000.0 / mean);
        calc = std::isfinite(v) ? (int)floor(v + 0.5) : b->max_size;
    }
    const int tgt = std::max(b->min_size, std::min(b->max_size, calc));
    const int slot = b->head;
    b->head = (b->head + 1) % b->max_size;
    b->count = std::min(b->count + 1, tgt);
    b->inserted++;
    *head = b->head, *count = b->count, *target = tgt;
    return slot;
}

// Batch form used by the ingestion hot path: n events -> (worker, slot) table + the new (head, count)
// window of every worker that received a row.  `buffers` is an array of psb_buffer_create handles.
extern "C" int psb_buffers_on_batch(void** buffers, int num_buffers, const int* workers, const double* now_ms, int n,
                                    int* worker_slot, int* win) {
    for (int i = 0; i < n; ++i) {
        const int w = workers[i];
        if (w < 0 || w >= num_buffers) {
            psb_set_error("psb_buffers_on_batch: worker %d out of range", w);
            return PSB_E_INVALID;
        }
        int head, count, target;
        const int slot = psb_buffer_on_event(buffers[w], now_ms[i], &head, &count, &target);
        worker_slot[2 * i] = w, worker_slot[2 * i + 1] = slot;
        win[2 * w] = head, win[2 * w + 1] = count;
    }
    return 0;
}
