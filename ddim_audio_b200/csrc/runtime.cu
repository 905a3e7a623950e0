// Library runtime: error text, kernel-launch counter, and an opt-in per-launch profiler.
// The profiler brackets every C-ABI call with two CUDA events ON THE CALLER'S STREAM (the stream the kernels are
// launched on) and aggregates elapsed time, algorithmic bytes and algorithmic FLOPs per kernel class; bench.py
// turns it on for its roofline leg.  It is off by default and costs one relaxed load per call when off.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ddim_b200.h"
#include "common.cuh"

static thread_local char g_err[512] = "";
void ddimb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* ddimb_last_error(void) { return g_err; }
extern "C" int ddimb_version(void) { return 101; }

static std::atomic<long long> g_launches{0};
void ddimb_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long long ddimb_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

namespace {
struct ProfRec {
  cudaEvent_t beg, end;
  char name[40];
  double bytes, flops;
  long long launches0;
  int nlaunch;
};
std::atomic<int> g_prof_on{0};
std::mutex g_prof_mu;
std::vector<ProfRec> g_recs;
std::vector<cudaEvent_t> g_free_events;

cudaEvent_t get_event() {
  if (!g_free_events.empty()) {
    cudaEvent_t e = g_free_events.back();
    g_free_events.pop_back();
    return e;
  }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

DdimbProfScope::DdimbProfScope(cudaStream_t s, const char* name, int p0, int p1, double bytes, double flops) : slot(-1), stream(s) {
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r;
  r.beg = get_event();
  r.end = get_event();
  snprintf(r.name, sizeof(r.name), "%s[%d,%d]", name, p0, p1);
  r.bytes = bytes;
  r.flops = flops;
  r.launches0 = g_launches.load();
  r.nlaunch = 0;
  cudaEventRecord(r.beg, s);
  slot = (int)g_recs.size();
  g_recs.push_back(r);
}
DdimbProfScope::~DdimbProfScope() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec& r = g_recs[slot];
  r.nlaunch = (int)(g_launches.load() - r.launches0);
  cudaEventRecord(r.end, stream);
}

extern "C" int ddimb_prof_enable(int on) {
  g_prof_on.store(on ? 1 : 0);
  return DDIMB_OK;
}

// Synchronises the device, writes a JSON array of {name, calls, launches, ms, bytes, flops} (totals per class) into buf
// and clears the records.  Returns the number of bytes written, or DDIMB_ERR_ARG when buf is too small.
extern "C" int ddimb_prof_report(char* buf, int buflen) {
  DDIMB_CHECK_ARG(buf && buflen > 2);
  DDIMB_CUDA(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_prof_mu);
  struct Agg { long long calls = 0, launches = 0; double ms = 0, bytes = 0, flops = 0; };
  std::map<std::string, Agg> agg;
  for (ProfRec& r : g_recs) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.beg, r.end) != cudaSuccess) ms = 0.f;
    Agg& a = agg[r.name];
    a.calls += 1; a.launches += r.nlaunch; a.ms += ms; a.bytes += r.bytes; a.flops += r.flops;
    g_free_events.push_back(r.beg);
    g_free_events.push_back(r.end);
  }
  g_recs.clear();
  std::string out = "[";
  bool first = true;
  for (auto& kv : agg) {
    char line[256];
    snprintf(line, sizeof(line), "%s{\"name\":\"%s\",\"calls\":%lld,\"launches\":%lld,\"ms\":%.6f,\"bytes\":%.6e,\"flops\":%.6e}", first ? "" : ",",
             kv.first.c_str(), kv.second.calls, kv.second.launches, kv.second.ms, kv.second.bytes, kv.second.flops);
    out += line;
    first = false;
  }
  out += "]";
  if ((int)out.size() + 1 > buflen) {
    ddimb_set_error("ddimb_prof_report: buffer of %d bytes too small for %zu", buflen, out.size() + 1);
    return DDIMB_ERR_ARG;
  }
  memcpy(buf, out.c_str(), out.size() + 1);
  return (int)out.size();
}
