// Error string + version of liboflow_b200.
#include "common.cuh"

static thread_local char g_err[512] = "";

void of_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;
void of_count_launch(void) { __atomic_fetch_add(&g_launches, 1ull, __ATOMIC_RELAXED); }

extern "C" {
unsigned long long of_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }
int of_version(void) { return 100; }
const char* of_last_error(void) { return g_err; }
}
