// Library-wide plumbing: thread-local error text, the kernel-launch counter and the optional
// per-kernel event profiler behind hcg_profile_* (include/hcg.h).  No compute here.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "hcg_common.cuh"

namespace hcg {

char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int set_err(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

static std::atomic<uint64_t> g_launches{0};
static std::atomic<int> g_prof_on{0};

struct Mark {
  const char* name;  // string literal; nullptr = region start (not attributed)
  cudaEvent_t ev;
  cudaStream_t st;   // a mark is timed against the previous mark of the SAME stream (hcg_train_step uses two)
};

struct Profiler {
  std::mutex mu;
  std::vector<cudaEvent_t> pool;
  std::vector<Mark> marks;
  size_t next = 0;
  static constexpr size_t kMaxMarks = 1 << 16;

  void mark(const char* name, cudaStream_t st) {
    std::lock_guard<std::mutex> lk(mu);
    if (marks.size() >= kMaxMarks) return;
    if (next == pool.size()) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) return;
      pool.push_back(e);
    }
    cudaEvent_t e = pool[next++];
    cudaEventRecord(e, st);
    marks.push_back({name, e, st});
  }
  void reset() {
    std::lock_guard<std::mutex> lk(mu);
    marks.clear();
    next = 0;
  }
};

static Profiler& prof() {
  static Profiler p;
  return p;
}

void note_launch(const char* name, cudaStream_t st) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (g_prof_on.load(std::memory_order_relaxed)) prof().mark(name, st);
}

void note_region(cudaStream_t st) {
  if (g_prof_on.load(std::memory_order_relaxed)) prof().mark(nullptr, st);
}

void note_other(const char* name, cudaStream_t st) {
  if (g_prof_on.load(std::memory_order_relaxed)) prof().mark(name, st);
}

}  // namespace hcg

using namespace hcg;

extern "C" const char* hcg_last_error(void) { return err_buf(); }
extern "C" const char* hcg_version(void) { return "hcg 0.2 (sm_100a)"; }
extern "C" uint64_t hcg_launch_count(void) { return g_launches.load(); }

extern "C" int hcg_profile_enable(int on) {
  if (on) prof().reset();
  g_prof_on.store(on ? 1 : 0);
  return HCG_OK;
}

// Aggregates the marks recorded since hcg_profile_enable(1): for every kernel name the number of
// launches and the summed device time between the previous mark on the stream and the kernel's
// own end mark.  Writes up to `cap` rows; names are copied into name_buf as cap x 64 bytes.
extern "C" int hcg_profile_read(char* name_buf, float* total_ms, int64_t* counts, int cap) {
  Profiler& p = prof();
  std::lock_guard<std::mutex> lk(p.mu);
  if (!p.marks.empty()) HCG_CUDA_OK(cudaEventSynchronize(p.marks.back().ev));
  std::map<std::string, std::pair<double, int64_t>> agg;
  std::vector<std::string> order;
  if (!p.marks.empty()) HCG_CUDA_OK(cudaDeviceSynchronize());   // (marks of every stream)
  std::map<cudaStream_t, size_t> last;   // index of the previous mark of each stream
  for (size_t i = 0; i < p.marks.size(); ++i) {
    const auto prev = last.find(p.marks[i].st);
    const bool have_prev = prev != last.end();
    const size_t j = have_prev ? prev->second : 0;
    last[p.marks[i].st] = i;
    if (p.marks[i].name == nullptr || !have_prev) continue;
    float ms = 0.f;
    HCG_CUDA_OK(cudaEventElapsedTime(&ms, p.marks[j].ev, p.marks[i].ev));
    auto it = agg.find(p.marks[i].name);
    if (it == agg.end()) {
      order.push_back(p.marks[i].name);
      agg[p.marks[i].name] = {ms, 1};
    } else {
      it->second.first += ms;
      it->second.second += 1;
    }
  }
  int n = 0;
  for (const std::string& k : order) {
    if (n >= cap) break;
    strncpy(name_buf + (size_t)n * 64, k.c_str(), 63);
    name_buf[(size_t)n * 64 + 63] = 0;
    total_ms[n] = (float)agg[k].first;
    counts[n] = agg[k].second;
    ++n;
  }
  return n < 0 ? 0 : -n;  // non-positive: -(rows written); positive values are error codes
}

static std::atomic<int> g_cfg[HCG_CFG_COUNT_];

namespace hcg {
int config_get(int key) { return (key >= 0 && key < HCG_CFG_COUNT_) ? g_cfg[key].load() : -1; }
}  // namespace hcg

extern "C" int hcg_config_set(int key, int value) {
  if (key < 0 || key >= HCG_CFG_COUNT_) return -1;
  return g_cfg[key].exchange(value);
}
extern "C" int hcg_config_get(int key) { return hcg::config_get(key); }

// struct sizes, so that a foreign-language binding can verify its mirror of include/hcg.h
extern "C" size_t hcg_sizeof_graph(void) { return sizeof(HcgGraph); }
extern "C" size_t hcg_sizeof_params(void) { return sizeof(HcgParams); }
extern "C" size_t hcg_sizeof_grads(void) { return sizeof(HcgGrads); }
extern "C" size_t hcg_sizeof_saved(void) { return sizeof(HcgSaved); }
extern "C" size_t hcg_sizeof_ensemble(void) { return sizeof(HcgEnsemble); }
