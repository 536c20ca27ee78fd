// Error string + version of liboflow_b200.
#include "common.cuh"
#include <string.h>

static thread_local char g_err[512] = "";

void of_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;
void of_count_launch(void) { __atomic_fetch_add(&g_launches, 1ull, __ATOMIC_RELAXED); }

// ---- per-launch timeline (of_prof_*): one CUDA event after every launch of this library on the profiled stream;
// the duration of launch k is event[k] - event[k-1] (one stream, in order), so bench.py can report the live,
// in-step time of each kernel class without ncu.  Off by default: of_prof_mark() is one predictable branch.
#define OF_PROF_MAX 8192
static bool g_prof_on = false;
static cudaStream_t g_prof_stream = nullptr;
static cudaEvent_t g_prof_ev[OF_PROF_MAX + 1];
static int g_prof_nev_created = 0;
static int g_prof_n = 0;
static struct { char name[40]; double work; } g_prof_rec[OF_PROF_MAX];
static double g_prof_pending_work = 0.0;

void of_prof_note(double work) { if (g_prof_on) g_prof_pending_work = work; }

void of_prof_mark(const char* what) {
  if (!g_prof_on || g_prof_n >= OF_PROF_MAX) return;
  const int k = g_prof_n + 1;
  if (k >= g_prof_nev_created) {
    if (cudaEventCreate(&g_prof_ev[k]) != cudaSuccess) { g_prof_on = false; return; }
    g_prof_nev_created = k + 1;
  }
  if (cudaEventRecord(g_prof_ev[k], g_prof_stream) != cudaSuccess) { cudaGetLastError(); return; }   // e.g. capturing
  strncpy(g_prof_rec[g_prof_n].name, what, sizeof(g_prof_rec[0].name) - 1);
  g_prof_rec[g_prof_n].name[sizeof(g_prof_rec[0].name) - 1] = 0;
  g_prof_rec[g_prof_n].work = g_prof_pending_work;
  g_prof_pending_work = 0.0;
  ++g_prof_n;
}

extern "C" int of_prof_enable(int on, void* stream) {
  g_prof_on = false;
  g_prof_n = 0;
  g_prof_pending_work = 0.0;
  if (!on) return 0;
  g_prof_stream = (cudaStream_t)stream;
  if (g_prof_nev_created == 0) {
    OF_CUDA(cudaEventCreate(&g_prof_ev[0]));
    g_prof_nev_created = 1;
  }
  OF_CUDA(cudaEventRecord(g_prof_ev[0], g_prof_stream));
  g_prof_on = true;
  return 0;
}

// a named boundary in the timeline with no launch of its own (bench.py: absorbs the GPU-side delay it queues in front
// of every profiled step so that the host, not the device, is ahead for the whole step)
extern "C" int of_prof_marker(const char* name) {
  of_prof_mark(name ? name : "marker");
  return 0;
}

extern "C" int of_prof_collect(of_prof_rec* out, int max_recs) {
  const bool was_on = g_prof_on;
  g_prof_on = false;
  if (g_prof_n > 0) OF_CUDA(cudaEventSynchronize(g_prof_ev[g_prof_n]));
  int n = g_prof_n < max_recs ? g_prof_n : max_recs;
  for (int i = 0; i < n; ++i) {
    float ms = 0.f;
    OF_CUDA(cudaEventElapsedTime(&ms, g_prof_ev[i], g_prof_ev[i + 1]));
    memcpy(out[i].name, g_prof_rec[i].name, sizeof(out[i].name));
    out[i].work = g_prof_rec[i].work;
    out[i].ms = ms;
  }
  g_prof_on = was_on;
  return n;
}

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
