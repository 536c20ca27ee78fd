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

// ---- kernel-variant options (of_set_option): a tiny process-wide table, empty by default --------------------
#include <string.h>
#define OF_MAX_OPTIONS 16
static struct { char name[32]; char value[32]; } g_opt[OF_MAX_OPTIONS];
static int g_nopt = 0;

const char* of_option_get(const char* name) {
  for (int i = 0; i < g_nopt; ++i)
    if (g_opt[i].value[0] && strcmp(g_opt[i].name, name) == 0) return g_opt[i].value;
  return nullptr;
}

extern "C" int of_set_option(const char* name, const char* value) {
  OF_REQUIRE(name && strlen(name) < sizeof(g_opt[0].name) && (!value || strlen(value) < sizeof(g_opt[0].value)),
             OF_ERR_ARG, "of_set_option: bad name / value");
  int slot = -1;
  for (int i = 0; i < g_nopt; ++i)
    if (strcmp(g_opt[i].name, name) == 0) slot = i;
  if (slot < 0) {
    if (!value) return 0;
    OF_REQUIRE(g_nopt < OF_MAX_OPTIONS, OF_ERR_STATE, "of_set_option: table full");
    slot = g_nopt++;
    strcpy(g_opt[slot].name, name);
  }
  strcpy(g_opt[slot].value, value ? value : "");
  return 0;
}

extern "C" {
unsigned long long of_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }
int of_version(void) { return 100; }
const char* of_last_error(void) { return g_err; }
}
