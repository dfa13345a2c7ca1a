// psb_worker_step_static: the worker step for callers that are not allowed to allocate device memory.
//
// Who needs this.  Some hosts embed libpsb200.so inside a process whose device memory is owned by
// somebody else: an inference server or training framework that reserved (nearly) the whole GPU for its
// own pool at start-up, a real-time ingestion loop that may not take the implicit device-wide
// synchronisation cudaMalloc / cudaFree imply, or a plugin interface that simply forbids allocation
// calls after initialisation.  Such a caller can hand us its inputs and outputs, but it has nowhere to
// get the solver's work vectors (psb_worker_scratch_bytes() per worker) from.
//
// What it does.  Exactly psb_worker_step, except that the solver scratch is not a caller buffer but a
// fixed array that is part of this library's device image: PSB_STATIC_SCRATCH_FLOATS floats
// (about 1.9 MiB), enough for PSB_STATIC_MAX_WORKERS workers at the reference model size (F = 1024),
// or more workers of a smaller model.  Every GPU that loads the library carries that array until the
// process exits.  Requests that need more are refused with PSB_E_LIMIT and a message that says how
// much was asked for and how much there is; nothing is launched in that case.
//
// Rules of use.
//   * The array is shared by every call on a device and keeps whatever the previous call left in it.
//     Nothing in it has to start at zero: each work vector (x, g, d, trial point, trial gradient, delta,
//     start weights, L-BFGS pairs) is fully written by a step before that step reads it, so no clearing
//     pass is needed or done.
//   * Two calls whose kernels can be on the GPU at the same time (different streams, or different host
//     threads without ordering between them) would both write the array: that is a data race.  Keep all
//     calls for one device on one stream, or order them with events.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress) with that
//     device current and remembers it.  Make that first call outside of any stream capture.
//   * No cudaMalloc, cudaFree or other allocating routine is called on this path, here or in the
//     launchers it shares with psb_worker_step.
//
// The name `g_static_scratch` below is only a handle in host code — it is NOT a device pointer there and
// must never be passed to a kernel directly; the address that kernels may use comes from
// cudaGetSymbolAddress.  (Device code in this file could name the array directly.)
#include <mutex>

#include "common.cuh"

#define PSB_STATIC_MAX_WORKERS 4
#define PSB_STATIC_SCRATCH_FLOATS (PSB_STATIC_MAX_WORKERS * PSB_SCRATCH_VECS * (1024 + 1) * PSB_CLASS_PAD)

static __device__ __align__(256) float g_static_scratch[PSB_STATIC_SCRATCH_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
float* g_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_scratch_mutex;

// Device address of the static array on the current device (looked up once per device).
int static_scratch(float** out) {
    int dev = 0;
    PSB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        psb_set_error("psb_worker_step_static: device ordinal %d is beyond the %d devices supported", dev, kMaxDevices);
        return PSB_E_LIMIT;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        PSB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        g_scratch_addr[dev] = (float*)p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}
}  // namespace

extern "C" size_t psb_static_scratch_bytes(void) { return (size_t)PSB_STATIC_SCRATCH_FLOATS * sizeof(float); }

extern "C" int psb_worker_step_static(const PsbSolver* s, const PsbComm* comm, int num_workers, int impl,
                                      const void* d_X, const int* d_labels, const int* d_win, const float* d_w_in,
                                      float* d_delta, float* d_loss, int* d_info, void* stream) {
    PSB_REQUIRE(s && comm, "psb_worker_step_static: null solver/comm");
    PSB_REQUIRE(num_workers >= 1 && s->num_features >= 2, "psb_worker_step_static: bad num_workers/num_features");
    const size_t need = (size_t)num_workers * psb_worker_scratch_bytes(s->num_features);
    if (need > psb_static_scratch_bytes()) {
        psb_set_error("psb_worker_step_static: %d workers x %d features need %zu bytes of solver scratch but the "
                      "built-in array holds %zu bytes (%d workers at 1024 features); use psb_worker_step with a "
                      "caller-provided scratch buffer for larger jobs",
                      num_workers, s->num_features, need, psb_static_scratch_bytes(), PSB_STATIC_MAX_WORKERS);
        return PSB_E_LIMIT;
    }
    float* scratch = nullptr;
    int rc = static_scratch(&scratch);
    if (rc) return rc;
    return psb_worker_step(s, comm, num_workers, impl, d_X, d_labels, d_win, d_w_in, d_delta, d_loss, d_info, scratch,
                           stream);
}
