// Host-side runtime glue shared by every translation unit: the thread-local error
// string behind the C ABI's integer return codes, and version/feature queries.
#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace {
thread_local char g_err[1024] = {0};
}

extern "C" {

void b200_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* b200_last_error(void) { return g_err; }

int b200_version(void) { return 100; }

}  // extern "C"
