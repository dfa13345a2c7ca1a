// ThreadSanitizer driver for the C++ host runtime (csrc/host_*.cpp): the streaming producer thread filling its
// ring of batches (cut lines, tokenise them with two helper threads, palette-pack, publish) against a consumer that polls, holds several batches at a
// time and releases them in order — the access pattern of apps/node.py and apps/p2p_apps.py — plus the buffer
// policy and tracker used from a second thread the way the engine's slot-assignment helper thread uses them.
// Build + run: make tsan   (host-only: no GPU, no libcudart calls are made)
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../include/psb200.h"

int main() {
    const int F = 64, W = 4, rows = 3000, batch_rows = 256, nb = 4;
    const std::string path = "/tmp/psb_tsan_train.csv";
    {
        FILE* f = fopen(path.c_str(), "w");
        for (int c = 0; c < F; ++c) fprintf(f, "%d,", c);
        fprintf(f, "Score\n");
        unsigned s = 12345;
        for (int r = 0; r < rows; ++r) {
            for (int c = 0; c < F; ++c) {
                s = s * 1664525u + 1013904223u;
                fprintf(f, "%s,", (s >> 28) < 4 ? "0.25" : "0");
            }
            fprintf(f, "%d\n", 1 + (int)(s % 5));
        }
        fclose(f);
    }
    std::vector<uint16_t> h_rows((size_t)nb * batch_rows * F);
    std::vector<int> h_labels((size_t)nb * batch_rows), h_workers((size_t)nb * batch_rows), h_off((size_t)nb * (batch_rows + 1));
    std::vector<double> h_times((size_t)nb * batch_rows);
    std::vector<uint32_t> h_bitmap((size_t)nb * batch_rows * (F / 32));
    const long long stride = (long long)batch_rows * (2 * F + 2);
    std::vector<uint8_t> h_payload((size_t)nb * stride);
    setenv("PSB_PRODUCER_THREADS", "3", 1);          // the reading thread + two tokenising helpers
    void* p = psb_producer_start_palette(path.c_str(), 1, W, 0, W, F, 0, batch_rows, nb, h_rows.data(), h_labels.data(),
                                         h_workers.data(), h_times.data(), 5, h_bitmap.data(), h_off.data(),
                                         h_payload.data(), stride);
    if (!p) {
        fprintf(stderr, "producer_start failed: %s\n", psb_last_error());
        return 1;
    }
    // consumer: hold up to 3 batches, hand each to a helper thread (slot assignment) before releasing it
    std::vector<void*> buffers(W);
    for (auto& b : buffers) b = psb_buffer_create(16, 128, 0.3f, 500);
    void* tracker = psb_tracker_create(W, 2);
    long long got = 0, checksum = 0;
    std::vector<int> held;
    int eof = 0;
    while (true) {
        int b = 0, n = 0;
        if (psb_producer_poll(p, &b, &n, &eof)) {
            held.push_back(b);
            std::vector<int> ws(2 * (size_t)n), win(2 * (size_t)W);
            std::thread helper([&] {
                psb_buffers_on_batch(buffers.data(), W, h_workers.data() + (size_t)b * batch_rows,
                                     h_times.data() + (size_t)b * batch_rows, n, ws.data(), win.data());
                for (int i = 0; i < n; ++i) checksum += h_labels[(size_t)b * batch_rows + i] + h_payload[(size_t)b * stride + i];
            });
            helper.join();
            got += n;
            if (held.size() == 3) {
                psb_producer_release(p, held.front());
                held.erase(held.begin());
            }
        } else if (eof) {
            break;
        } else {
            std::this_thread::yield();
        }
    }
    for (int b : held) psb_producer_release(p, b);
    int replies[16];
    for (int c = 0; c < 3; ++c)
        for (int w = 0; w < W; ++w) psb_tracker_on_gradient(tracker, w, c, replies, 8);
    const long long sent = psb_producer_rows_sent(p);
    psb_producer_stop(p);
    for (auto b : buffers) psb_buffer_destroy(b);
    psb_tracker_destroy(tracker);
    printf("tsan driver: %lld rows consumed, %lld sent, checksum %lld, eof %d\n", got, sent, checksum, eof);
    return got == rows && eof == 1 ? 0 : 2;
}
