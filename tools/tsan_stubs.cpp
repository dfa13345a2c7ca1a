// psb_set_error lives in csrc/api.cu (a CUDA translation unit); the host-only ThreadSanitizer build supplies its own.
#include <cstdarg>
#include <cstdio>
static thread_local char g_err[512] = "";
void psb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char* psb_last_error(void) { return g_err; }
