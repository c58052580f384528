// Per-kernel CUDA-event profiler behind dlwp_profile_begin / _end / _entry.
//
// While profiling is on, every launch site records one event on its stream right after the
// launch; the device time attributed to launch i is event[i] - event[i-1] (kernels of one
// stream run back to back, so this is the kernel's duration plus its launch gap).  bench.py
// uses it to time the dominant kernel live for the roofline figure; the hot path itself runs
// with profiling off and never touches events.
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace dlwp {

bool g_prof_on = false;

namespace {
struct Entry {
  std::string name;
  double total_ms = 0.0;
  int64_t launches = 0;
};
std::vector<cudaEvent_t> g_events;   // pool, grown on demand
std::vector<const char*> g_names;    // g_names[i] belongs to g_events[i + 1]
size_t g_used = 0;                   // events recorded since profile_begin
std::vector<Entry> g_entries;

cudaEvent_t next_event() {
  if (g_used == g_events.size()) {
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
    g_events.push_back(e);
  }
  return g_events[g_used++];
}
}  // namespace

void prof_mark(const char* name, cudaStream_t st) {
  cudaEvent_t e = next_event();
  if (!e) return;
  cudaEventRecord(e, st);
  g_names.push_back(name);
}

}  // namespace dlwp

using namespace dlwp;

extern "C" {

int dlwp_profile_begin(void* stream) {
  g_used = 0;
  g_names.clear();
  g_entries.clear();
  cudaEvent_t e = next_event();
  if (!e) return fail(DLWP_ERR_CUDA, "cudaEventCreate failed");
  DLWP_CUDA(cudaEventRecord(e, (cudaStream_t)stream));
  g_prof_on = true;
  return DLWP_OK;
}

int dlwp_profile_end(void* stream) {
  if (!g_prof_on) return fail(DLWP_ERR_STATE, "dlwp_profile_end without dlwp_profile_begin");
  g_prof_on = false;
  DLWP_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  std::map<std::string, size_t> index;
  for (size_t i = 0; i + 1 < g_used; ++i) {
    float ms = 0.f;
    DLWP_CUDA(cudaEventSynchronize(g_events[i + 1]));
    DLWP_CUDA(cudaEventElapsedTime(&ms, g_events[i], g_events[i + 1]));
    const std::string name = g_names[i];
    auto it = index.find(name);
    if (it == index.end()) {
      it = index.emplace(name, g_entries.size()).first;
      g_entries.push_back(Entry{name, 0.0, 0});
    }
    g_entries[it->second].total_ms += ms;
    g_entries[it->second].launches += 1;
  }
  return DLWP_OK;
}

int dlwp_profile_count(void) { return (int)g_entries.size(); }

int dlwp_profile_entry(int i, const char** name, double* total_ms, int64_t* launches) {
  if (i < 0 || i >= (int)g_entries.size() || !name || !total_ms || !launches)
    return fail(DLWP_ERR_INVALID, "dlwp_profile_entry: index %d out of range or null argument", i);
  *name = g_entries[i].name.c_str();
  *total_ms = g_entries[i].total_ms;
  *launches = g_entries[i].launches;
  return DLWP_OK;
}

}  // extern "C"
