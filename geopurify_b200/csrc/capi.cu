// Error state and version of libgeopurify_b200.so.
#include <stdarg.h>
#include "common.cuh"

static thread_local char g_err[512] = "";

void gp_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* gp_last_error(void) { return g_err; }
extern "C" int gp_version(void) { return 100; }

// ---- optional per-kernel timing (off by default): CUDA events recorded on the launching stream around the launches
// of the dominant kernels, so that bench.py can report their average launch duration without a profiler.
#define GP_TIMING_SLOTS 4
#define GP_TIMING_MAX 2048
static int g_timing_on = 0;
static cudaEvent_t g_ev[GP_TIMING_SLOTS][GP_TIMING_MAX][2];
static int g_ev_made[GP_TIMING_SLOTS] = {0, 0, 0, 0};
static int g_ev_used[GP_TIMING_SLOTS] = {0, 0, 0, 0};

extern "C" int gp_timing_enable(int on) {
  g_timing_on = on;
  for (int k = 0; k < GP_TIMING_SLOTS; ++k) g_ev_used[k] = 0;
  return GP_OK;
}

bool gp_timing_begin(int which, cudaStream_t st) {
  if (!g_timing_on || which < 0 || which >= GP_TIMING_SLOTS || g_ev_used[which] >= GP_TIMING_MAX) return false;
  int i = g_ev_used[which];
  if (i >= g_ev_made[which]) {
    if (cudaEventCreate(&g_ev[which][i][0]) != cudaSuccess || cudaEventCreate(&g_ev[which][i][1]) != cudaSuccess)
      return false;
    g_ev_made[which] = i + 1;
  }
  cudaEventRecord(g_ev[which][i][0], st);
  return true;
}

void gp_timing_end(int which, cudaStream_t st) {
  cudaEventRecord(g_ev[which][g_ev_used[which]][1], st);
  ++g_ev_used[which];
}

extern "C" int gp_timing_read(int which, double* total_ms, int* launches) {
  GP_CHECK_ARG(which >= 0 && which < GP_TIMING_SLOTS && total_ms && launches, "gp_timing_read: bad arguments");
  double tot = 0.0;
  for (int i = 0; i < g_ev_used[which]; ++i) {
    GP_CUDA(cudaEventSynchronize(g_ev[which][i][1]));
    float ms = 0.f;
    GP_CUDA(cudaEventElapsedTime(&ms, g_ev[which][i][0], g_ev[which][i][1]));
    tot += ms;
  }
  *total_ms = tot;
  *launches = g_ev_used[which];
  return GP_OK;
}
