#include <cstdarg>
#include <cstdint>
#include <cstdio>
struct igpu_ctx;
static thread_local char g_err[512];
int32_t igpu_fail(igpu_ctx*, int32_t status, const char* fmt, ...) { va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap); return status; }
extern "C" const char* igpu_last_error(const igpu_ctx*) { return g_err; }
