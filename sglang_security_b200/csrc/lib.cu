// Library-level C-ABI entry points: version, error text, launch counter.
#include "common.cuh"

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace b200 {

static thread_local char g_error[1024] = "";
std::atomic<int64_t> g_launch_count{0};

static int pdl_default() {
  const char* v = std::getenv("B200_PDL");
  if (!(v && *v)) return 0xff;
  const int m = std::atoi(v);  // bit mask over PdlFamily; 1 is read as "all"
  return m == 1 ? 0xff : m;
}
std::atomic<int> g_pdl_on{pdl_default()};

int pdl_max_tokens() {
  static int v = [] {
    const char* e = std::getenv("B200_PDL_MAX_M");
    // 512: re-measured in round 2 on the whole-model decode step (profiles/r02_bench_pdl.json): batch
    // 64 x 4k 8.11 ms with PDL against 8.25 ms without; batch 128 / 256 @ 1k 5.99 / 9.62 ms against
    // 6.20 / 9.84 ms with a limit of 64 tokens (round 1 had measured a loss from 64 tokens on)
    return (e && *e) ? std::atoi(e) : 512;
  }();
  return v;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

static thread_local char g_variant[256] = "";
void note_variant(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_variant, sizeof(g_variant), fmt, ap);
  va_end(ap);
}
const char* last_variant() { return g_variant; }

int device_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- per-kernel timing registry -----------------------------------------------------------------
struct TimingRecord {
  std::string name;
  cudaEvent_t start, stop;
};
static std::atomic<int> g_timing_on{0};
static std::mutex g_timing_mu;
static std::vector<TimingRecord> g_timing;

ScopedKernelTimer::ScopedKernelTimer(const char* name, cudaStream_t st) : name_(name), st_(st) {
  if (!g_timing_on.load(std::memory_order_relaxed)) return;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) return;
  if (cudaEventCreate(&start_) != cudaSuccess) {
    start_ = nullptr;
    return;
  }
  cudaEventRecord(start_, st_);
}

ScopedKernelTimer::~ScopedKernelTimer() {
  if (!start_) return;
  cudaEvent_t stop = nullptr;
  if (cudaEventCreate(&stop) != cudaSuccess) {
    cudaEventDestroy(start_);
    return;
  }
  cudaEventRecord(stop, st_);
  std::lock_guard<std::mutex> lk(g_timing_mu);
  g_timing.push_back(TimingRecord{name_, start_, stop});
}

}  // namespace b200

extern "C" {

int b200_timing_enable(int on) {
  b200::g_timing_on.store(on ? 1 : 0);
  return 0;
}

// Sums (and clears) the recorded launches of kernel `name` ("" or NULL = all kernels).
int b200_timing_read(const char* name, int64_t* launches, double* total_ms) {
  using namespace b200;
  std::lock_guard<std::mutex> lk(g_timing_mu);
  int64_t n = 0;
  double ms = 0.0;
  std::vector<TimingRecord> keep;
  for (auto& r : g_timing) {
    if (name && name[0] && r.name != name) {
      keep.push_back(r);
      continue;
    }
    float t = 0.f;
    if (cudaEventSynchronize(r.stop) == cudaSuccess &&
        cudaEventElapsedTime(&t, r.start, r.stop) == cudaSuccess) {
      ms += t;
      ++n;
    }
    cudaEventDestroy(r.start);
    cudaEventDestroy(r.stop);
  }
  g_timing.swap(keep);
  if (launches) *launches = n;
  if (total_ms) *total_ms = ms;
  return 0;
}

int b200_version(void) { return 2; }

int b200_set_pdl(int on) {
  const int prev = b200::g_pdl_on.exchange(on == 1 ? 0xff : on);
  return prev;
}

const char* b200_last_error(void) { return b200::g_error; }

const char* b200_last_variant(void) { return b200::last_variant(); }

int64_t b200_launch_count(void) { return b200::g_launch_count.load(); }

}  // extern "C"
