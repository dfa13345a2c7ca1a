// AddressSanitizer / UBSan driver for the host packers, the CSV tokenizer and the producer thread: buffers of exactly the documented size,
// rows from empty to fully dense, so that any store past a buffer's end is caught.  `make asan`.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../include/psb200.h"

static void expect(bool ok, const char* what) {
    if (!ok) {
        fprintf(stderr, "asan driver: %s\n", what);
        exit(1);
    }
}

// asan_packers <rows file> <num_features> <rows>: stream a PSBROWS1 file through the producer thread into buffers of
// exactly the documented size, as palette batches and as dense batches (tests/test_sanitizers.py writes the file).
static int stream_rows_file(const char* path, int F, long long want_rows) {
    for (int palette = 0; palette < 2; ++palette) {
        const int batch_rows = 96, nb = 3;
        uint16_t* rows = (uint16_t*)malloc((size_t)nb * batch_rows * F * 2);
        int* labels = (int*)malloc((size_t)nb * batch_rows * 4);
        int* workers = (int*)malloc((size_t)nb * batch_rows * 4);
        double* times = (double*)malloc((size_t)nb * batch_rows * 8);
        uint32_t* bitmap = (uint32_t*)malloc((size_t)nb * batch_rows * (F / 32) * 4);
        int* row_off = (int*)malloc((size_t)nb * (batch_rows + 1) * 4);
        const long long stride = (long long)batch_rows * (2 * F + 2);
        uint8_t* payload = (uint8_t*)malloc((size_t)nb * stride);
        void* p = palette ? psb_producer_start_palette(path, 1, 3, 0, 3, F, 0, batch_rows, nb, rows, labels, workers, times,
                                                       5, bitmap, row_off, payload, stride)
                          : psb_producer_start(path, 1, 3, 0, 3, F, 0, batch_rows, nb, rows, labels, workers, times, 5);
        expect(p != nullptr, "producer_start on a rows file");
        long long got = 0, checksum = 0;
        int eof = 0;
        for (;;) {
            int b = 0, n = 0;
            if (psb_producer_poll(p, &b, &n, &eof)) {
                if (palette) {
                    const int* off = row_off + (size_t)b * (batch_rows + 1);
                    expect(off[0] == 0 && off[n] <= stride, "rows file: payload offsets of a batch");
                    for (int i = 0; i < off[n]; ++i) checksum += payload[(size_t)b * stride + i];
                } else {
                    for (size_t i = 0; i < (size_t)n * F; ++i) checksum += rows[(size_t)b * batch_rows * F + i];
                }
                got += n;
                psb_producer_release(p, b);
            } else if (eof) {
                break;
            }
        }
        expect(eof == 1 && got == want_rows, "rows file: wrong number of rows");
        psb_producer_stop(p);
        free(rows), free(labels), free(workers), free(times), free(bitmap), free(row_off), free(payload);
        printf("asan driver: rows file as %s batches: %lld rows, checksum %lld\n", palette ? "palette" : "dense", got, checksum);
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc == 4) return stream_rows_file(argv[1], atoi(argv[2]), atoll(argv[3]));
    std::mt19937 rng(7);
    const double densities[] = {0.0, 0.02, 0.05, 0.3, 0.9, 1.0};
    long long checked = 0;
    for (int F : {32, 96, 1024})
        for (double dens : densities)
            for (int distinct : {3, 16, 17, 400}) {
                const int n = 9;
                uint16_t* rows = (uint16_t*)malloc((size_t)n * F * 2);
                for (int i = 0; i < n * F; ++i) {
                    const bool on = (rng() % 10000) < dens * 10000;
                    uint16_t v = (uint16_t)(0x3c00 + (rng() % distinct));
                    if (rng() % 2) v |= 0x8000;
                    rows[i] = on ? v : (uint16_t)((rng() % 2) ? 0x8000 : 0);      // -0.0 counts as absent
                }
                uint32_t* bitmap = (uint32_t*)malloc((size_t)n * (F / 32) * 4);
                int* row_off = (int*)malloc((size_t)(n + 1) * 4);
                uint16_t* values = (uint16_t*)malloc((size_t)n * F * 2);
                const long long cap = (long long)n * (2 * F + 2);
                uint8_t* payload = (uint8_t*)malloc((size_t)cap);
                const long long nnz = psb_pack_bitmap_dense(rows, n, F, bitmap, row_off, values);
                long long want = 0;
                for (int i = 0; i < n * F; ++i) want += (rows[i] & 0x7fff) != 0;
                expect(nnz == want && row_off[n] == want, "bitmap packer: wrong number of non-zeros");
                long long k = 0;
                for (int i = 0; i < n * F; ++i)
                    if (rows[i] & 0x7fff) expect(values[k++] == rows[i], "bitmap packer: wrong value order");
                const long long bytes = psb_pack_palette_dense(rows, n, F, bitmap, row_off, payload, cap);
                expect(bytes >= 2 * n && bytes <= cap && row_off[n] == bytes, "palette packer: bad payload size");
                for (int r = 0; r < n; ++r) {                         // decode and compare
                    const uint8_t* p = payload + row_off[r];
                    uint16_t npal;
                    memcpy(&npal, p, 2);
                    int j = 0;
                    for (int f = 0; f < F; ++f) {
                        const uint16_t h = rows[(size_t)r * F + f];
                        const bool present = (bitmap[(size_t)r * (F / 32) + f / 32] >> (f % 32)) & 1u;
                        expect(present == ((h & 0x7fff) != 0), "palette packer: bitmap bit");
                        if (!present) continue;
                        uint16_t got;
                        if (npal == 0) {
                            memcpy(&got, p + 2 + 2 * j, 2);
                        } else {
                            const uint8_t c = p[2 + 2 * npal + j / 2];
                            memcpy(&got, p + 2 + 2 * ((j & 1) ? (c >> 4) : (c & 15)), 2);
                        }
                        expect(got == h, "palette packer: value");
                        ++j;
                    }
                }
                // the CSR packer: same rows as (unsorted) triplets
                std::vector<long long> indptr(1, 0);
                std::vector<int> idx;
                std::vector<float> val;
                for (int r = 0; r < n; ++r) {
                    for (int f = F - 1; f >= 0; --f)
                        if (rows[(size_t)r * F + f] & 0x7fff) idx.push_back(f), val.push_back((float)(1 + f % 7) * 0.25f);
                    indptr.push_back((long long)idx.size());
                }
                idx.push_back(0), val.push_back(0.f);                 // keeps data() non-null for empty inputs
                const long long nnz2 = psb_pack_bitmap_csr(indptr.data(), idx.data(), val.data(), n, F, bitmap, row_off, values);
                expect(nnz2 == want, "csr packer: wrong number of non-zeros");
                checked += n;
                free(rows), free(bitmap), free(row_off), free(values), free(payload);
            }
    // tokenizer: lines that end inside the 8-byte zero-run compare, with and without a final newline
    for (int cols = 1; cols <= 40; ++cols)
        for (int tail = 0; tail < 2; ++tail) {
            char path[] = "/tmp/psb_asan_XXXXXX";
            const int fd = mkstemp(path);
            expect(fd >= 0, "mkstemp");
            FILE* f = fdopen(fd, "w");
            for (int r = 0; r < 3; ++r) {
                for (int c = 0; c < cols; ++c) fputs(c == cols / 2 && r == 1 ? "0.5," : "0,", f);
                fputs(r == 2 && !tail ? "3" : "3\n", f);
            }
            fclose(f);
            long long n_rows = 0, n_nnz = 0;
            int n_cols = 0;
            expect(psb_csv_count(path, 0, &n_rows, &n_nnz, &n_cols) == 0, "csv_count");
            expect(n_rows == 3 && n_nnz == 1 && n_cols == cols, "csv_count: wrong shape");
            long long* indptr = (long long*)malloc(4 * 8);
            int* indices = (int*)malloc(4);
            float* values = (float*)malloc(4);
            int* labels = (int*)malloc(3 * 4);
            expect(psb_csv_parse(path, 0, indptr, indices, values, labels, n_rows, n_nnz) == 0, "csv_parse");
            expect(indices[0] == cols / 2 && values[0] == 0.5f && labels[2] == 3 && indptr[3] == 1, "csv_parse: wrong row");
            free(indptr), free(indices), free(values), free(labels);
            remove(path);
        }
    // the one-pass loader on ~3.3 MB of text: three chunks cut at line ends, tokenised side by side
    {
        char path[] = "/tmp/psb_asan_big_XXXXXX";
        const int fd = mkstemp(path);
        expect(fd >= 0, "mkstemp");
        FILE* f = fdopen(fd, "w");
        const int cols = 500, lines = 3300;
        for (int r = 0; r < lines; ++r) {
            for (int c = 0; c < cols; ++c) fputs(c == r % cols ? "-0.125," : "0,", f);
            fprintf(f, "%d%s", r % 6, r + 1 < lines ? "\n" : "");
        }
        fclose(f);
        for (int threads : {1, 3, 64}) {
            void* h = psb_csv_open(path, 0, threads);
            expect(h != nullptr, "csv_open");
            long long n_rows = 0, n_nnz = 0;
            int n_cols = 0;
            expect(psb_csv_shape(h, &n_rows, &n_nnz, &n_cols) == 0 && n_rows == lines && n_nnz == lines && n_cols == cols,
                   "csv_open: wrong shape");
            long long* indptr = (long long*)malloc((size_t)(n_rows + 1) * 8);
            int* indices = (int*)malloc((size_t)n_nnz * 4);
            float* values = (float*)malloc((size_t)n_nnz * 4);
            int* labels = (int*)malloc((size_t)n_rows * 4);
            expect(psb_csv_fill(h, indptr, indices, values, labels) == 0, "csv_fill");
            for (int r = 0; r < lines; ++r)
                expect(indptr[r + 1] == r + 1 && indices[r] == r % cols && values[r] == -0.125f && labels[r] == r % 6,
                       "csv_fill: wrong row");
            psb_csv_close(h);
            free(indptr), free(indices), free(values), free(labels);
        }
        remove(path);
        expect(psb_csv_open("/tmp/psb_asan_no_such_file.csv", 1, 0) == nullptr, "csv_open of a missing file");
    }
    // producer thread on awkward text: an empty line, CRLF, a line longer than the 4 MB read block (the block grows),
    // more columns than features, junk fields, no newline at the end; with and without tokenising helper threads
    for (const char* threads : {"1", "3"}) {
        setenv("PSB_PRODUCER_THREADS", threads, 1);
        char path[] = "/tmp/psb_asan_prod_XXXXXX";
        const int fd = mkstemp(path);
        expect(fd >= 0, "mkstemp");
        FILE* f = fdopen(fd, "w");
        const int F = 32, rows_ok = 300;
        for (int c = 0; c < F; ++c) fprintf(f, "%d,", c);
        fputs("Score\n", f);
        for (int r = 0; r < rows_ok; ++r) {
            for (int c = 0; c < F; ++c) fputs(c == r % F ? "0.25," : "0,", f);
            fprintf(f, "%d%s", 1 + r % 5, r % 7 == 0 ? "\r\n" : "\n");
            if (r == 10) fputs("\n", f);                                    // an empty line is not a row
        }
        for (int c = 0; c < 2600000; ++c) fputs(c == 5 ? "1.5," : "0,", f);  // 5.2 MB of text in one line
        fputs("2\n", f);
        fputs("abc,-,+,.,0x10,1e400,nan,-0.000,", f);                         // junk in the first columns
        for (int c = 8; c < F; ++c) fputs("0,", f);
        fputs("4", f);                                                        // last line, no newline
        fclose(f);
        const int batch_rows = 128, nb = 3;
        uint16_t* rows = (uint16_t*)malloc((size_t)nb * batch_rows * F * 2);
        int* labels = (int*)malloc((size_t)nb * batch_rows * 4);
        int* workers = (int*)malloc((size_t)nb * batch_rows * 4);
        double* times = (double*)malloc((size_t)nb * batch_rows * 8);
        uint32_t* bitmap = (uint32_t*)malloc((size_t)nb * batch_rows * (F / 32) * 4);
        int* row_off = (int*)malloc((size_t)nb * (batch_rows + 1) * 4);
        const long long stride = (long long)batch_rows * (2 * F + 2);
        uint8_t* payload = (uint8_t*)malloc((size_t)nb * stride);
        void* p = psb_producer_start_palette(path, 1, 2, 0, 2, F, 0, batch_rows, nb, rows, labels, workers, times, 5,
                                             bitmap, row_off, payload, stride);
        expect(p != nullptr, "producer_start");
        long long got = 0, long_line_ok = 0, junk_ok = 0;
        int eof = 0;
        for (;;) {
            int b = 0, n = 0;
            if (psb_producer_poll(p, &b, &n, &eof)) {
                for (int i = 0; i < n; ++i) {
                    const size_t o = (size_t)b * batch_rows + i;
                    const long long g = got + i;
                    if (g < rows_ok) {
                        expect(labels[o] == 1 + g % 5 && workers[o] == g % 2, "producer: label / partition of a plain row");
                        expect(rows[o * F + g % F] == 0x3400, "producer: 0.25 in its column");       // fp16(0.25)
                    } else if (g == rows_ok) {
                        long_line_ok = labels[o] == 2 && rows[o * F + 5] == 0x3e00;                   // fp16(1.5)
                    } else {
                        // abc -> 0, "-" "+" "." -> 0, 0x10 -> 16, 1e400 -> inf, nan -> nan, -0.000 -> -0 (absent)
                        junk_ok = labels[o] == 4 && rows[o * F] == 0 && rows[o * F + 4] == 0x4c00 &&
                                  rows[o * F + 5] == 0x7c00 && (rows[o * F + 6] & 0x7c00) == 0x7c00 &&
                                  (rows[o * F + 7] & 0x7fff) == 0;
                    }
                }
                got += n;
                psb_producer_release(p, b);
            } else if (eof) {
                break;
            }
        }
        expect(eof == 1 && got == rows_ok + 2, "producer: wrong number of rows");
        expect(long_line_ok && junk_ok, "producer: the long line or the junk line came out wrong");
        expect(psb_producer_rows_sent(p) == rows_ok + 2, "producer: rows_sent");
        psb_producer_stop(p);
        free(rows), free(labels), free(workers), free(times), free(bitmap), free(row_off), free(payload);
        remove(path);
    }
    // buffer policy: tables of exactly n / W entries; a batch with more rows than a ring (the superseded-rows pass),
    // an out-of-range worker (refused before anything is written), state save / restore with a wrong size
    {
        const int W = 3, n = 50;
        void* bufs[W];
        for (auto& b : bufs) b = psb_buffer_create(4, 8, 0.3f, 16);
        expect(psb_buffer_set_ring(bufs[0], 12) == 0 && psb_buffer_set_ring(bufs[1], 4) != 0, "set_ring");
        int* workers = (int*)malloc(n * 4);
        double* now = (double*)malloc(n * 8);
        int* ws = (int*)malloc(2 * n * 4);
        int* win = (int*)malloc(2 * W * 4);
        int* change = (int*)malloc(4 * W * 4);
        for (int i = 0; i < n; ++i) workers[i] = i % 5 == 0 ? 1 : 0, now[i] = 10.0 * i;
        int most = 0;
        expect(psb_buffers_on_batch2(bufs, W, workers, now, n, ws, win, change, &most) == 0 && most == 40, "on_batch2");
        int kept0 = 0, kept1 = 0;
        for (int i = 0; i < n; ++i) {
            expect(ws[2 * i] == workers[i] && ws[2 * i + 1] >= -1 && ws[2 * i + 1] < 12, "on_batch2: slot range");
            if (ws[2 * i + 1] >= 0) (workers[i] ? kept1 : kept0)++;
        }
        expect(kept0 == 12 && kept1 == 8 && change[4 * 2 + 2] == 0, "on_batch2: rows kept per ring");
        workers[7] = W;
        expect(psb_buffers_on_batch2(bufs, W, workers, now, n, ws, win, change, &most) != 0, "on_batch2: bad worker");
        double* state = (double*)malloc((9 + 16) * 8);
        expect(psb_buffer_get_state(bufs[0], state, 9 + 15) < 0 && psb_buffer_get_state(bufs[0], state, 9 + 16) == 25, "get_state");
        expect(psb_buffer_set_state(bufs[2], state, 24) != 0, "set_state: wrong size");
        void* twin = psb_buffer_create(4, 8, 0.3f, 16);
        expect(psb_buffer_set_ring(twin, 12) == 0 && psb_buffer_set_state(twin, state, 25) == 0, "set_state");
        int h1, c1, t1, h2, c2, t2;
        expect(psb_buffer_on_event(bufs[0], 1e4, &h1, &c1, &t1) == psb_buffer_on_event(twin, 1e4, &h2, &c2, &t2) &&
                   h1 == h2 && c1 == c2 && t1 == t2, "restored buffer continues like the original");
        void* tr = psb_tracker_create(2, 1);
        int replies[4];
        expect(psb_tracker_on_gradient(tr, 0, 0, replies, 2) >= 0 && psb_tracker_on_gradient(tr, 0, 5, replies, 2) < 0 &&
                   psb_tracker_on_gradient(tr, 9, 0, replies, 2) < 0, "tracker: protocol violations are refused");
        psb_tracker_destroy(tr);
        for (auto& b : bufs) psb_buffer_destroy(b);
        psb_buffer_destroy(twin);
        free(workers), free(now), free(ws), free(win), free(change), free(state);
    }
    printf("asan driver: %lld packed rows, 80 csv files, 2 awkward producer streams, buffer policy and tracker checked\n", checked);
    return 0;
}
