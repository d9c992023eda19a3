// Optional per-stage device timing (CUDA events on the caller's stream around each launch group).
// Off by default; bench.py switches it on to measure each kernel's live duration inside the timed
// region (roofline.achieved) and to count launches.  Stage names mirror the reference's
// record_function ranges where one exists (efgpnd.py:61-280).
#include "common.cuh"
#include <map>
#include <mutex>

namespace efgp {

namespace {
struct Rec {
  std::string name;
  cudaEvent_t a, b;
  int launches;
  long long units;   // points (or grid entries) the scope processed: the roofline's unit count
};
std::mutex g_tmu;
bool g_on = false;
std::vector<Rec> g_recs;
std::vector<cudaEvent_t> g_pool;
std::map<std::string, std::pair<long long, long long>> g_counts;  // name -> (scopes, launches), always kept

cudaEvent_t get_event() {
  if (!g_pool.empty()) {
    cudaEvent_t e = g_pool.back();
    g_pool.pop_back();
    return e;
  }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

bool timing_on() { return g_on; }

void timing_enable(int on) {
  std::lock_guard<std::mutex> lk(g_tmu);
  for (auto& r : g_recs) {
    g_pool.push_back(r.a);
    g_pool.push_back(r.b);
  }
  g_recs.clear();
  g_counts.clear();
  g_on = on != 0;
}

ScopedStage::ScopedStage(const char* name, cudaStream_t st, int launches, long long units) : st_(st), idx_(-1) {
  if (!g_on) return;
  std::lock_guard<std::mutex> lk(g_tmu);
  Rec r;
  r.name = name;
  r.a = get_event();
  r.b = get_event();
  r.launches = launches;
  r.units = units;
  cudaEventRecord(r.a, st);
  g_recs.push_back(r);
  idx_ = (long long)g_recs.size() - 1;
}

ScopedStage::~ScopedStage() {
  if (idx_ < 0) return;
  std::lock_guard<std::mutex> lk(g_tmu);
  if (idx_ < (long long)g_recs.size()) cudaEventRecord(g_recs[idx_].b, st_);
}

void ScopedStage::add_launches(int n) {
  if (idx_ < 0) return;
  std::lock_guard<std::mutex> lk(g_tmu);
  if (idx_ < (long long)g_recs.size()) g_recs[idx_].launches += n;
}

// "name,scopes,launches,units,total_ms\n" per stage, in first-seen order.  Synchronises the device.
std::string timing_report() {
  std::lock_guard<std::mutex> lk(g_tmu);
  cudaDeviceSynchronize();
  std::vector<std::string> order;
  std::map<std::string, double> ms;
  std::map<std::string, std::pair<long long, long long>> cnt;
  std::map<std::string, long long> units;
  for (auto& r : g_recs) {
    float t = 0;
    if (cudaEventElapsedTime(&t, r.a, r.b) != cudaSuccess) {
      cudaGetLastError();
      t = 0;
    }
    if (!ms.count(r.name)) order.push_back(r.name);
    ms[r.name] += t;
    cnt[r.name].first += 1;
    cnt[r.name].second += r.launches;
    units[r.name] += r.units;
  }
  std::string out;
  char buf[256];
  for (auto& n : order) {
    snprintf(buf, sizeof(buf), "%s,%lld,%lld,%lld,%.6f\n", n.c_str(), cnt[n].first, cnt[n].second, units[n], ms[n]);
    out += buf;
  }
  return out;
}

}  // namespace efgp
