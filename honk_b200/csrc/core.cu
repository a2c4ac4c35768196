// Error plumbing, device check and launch accounting for libhonk_b200.so.
#include "common.cuh"

#include <mutex>
#include <vector>

namespace honk {

std::atomic<long long> g_launches{0};

std::string& last_error_ref() {
  static thread_local std::string msg;
  return msg;
}

int set_error(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return code;
}

int num_sms(int device) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
  return n;
}

namespace {
struct ProfEntry {
  std::string name;
  std::vector<cudaEvent_t> begin, end;
};
std::mutex g_prof_mu;
std::vector<ProfEntry> g_prof;
std::atomic<int> g_prof_on{0};
}  // namespace

ProfSpan::ProfSpan(const char* name, cudaStream_t s) : stream(s) {
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int idx = -1;
  for (size_t i = 0; i < g_prof.size(); ++i)
    if (g_prof[i].name == name) idx = static_cast<int>(i);
  if (idx < 0) {
    g_prof.push_back(ProfEntry{name, {}, {}});
    idx = static_cast<int>(g_prof.size()) - 1;
  }
  cudaEvent_t b, e;
  if (cudaEventCreate(&b) != cudaSuccess || cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(b, s);
  g_prof[idx].begin.push_back(b);
  g_prof[idx].end.push_back(e);
  slot = idx;
}

ProfSpan::~ProfSpan() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEventRecord(g_prof[slot].end.back(), stream);
}

}  // namespace honk

using namespace honk;

extern "C" int honk_profile_enable(int on) {
  g_prof_on.store(on ? 1 : 0);
  return HONK_OK;
}

extern "C" int honk_profile_reset(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& p : g_prof) {
    for (auto e : p.begin) cudaEventDestroy(e);
    for (auto e : p.end) cudaEventDestroy(e);
  }
  g_prof.clear();
  return HONK_OK;
}

extern "C" int honk_profile_get(const char* span, double* ms_total, int64_t* n_spans) {
  HONK_CHECK(span && ms_total && n_spans, HONK_ERR_INVALID, "null argument");
  std::lock_guard<std::mutex> lk(g_prof_mu);
  *ms_total = 0;
  *n_spans = 0;
  for (auto& p : g_prof) {
    if (p.name != span) continue;
    for (size_t i = 0; i < p.begin.size(); ++i) {
      HONK_CUDA(cudaEventSynchronize(p.end[i]));
      float ms = 0;
      HONK_CUDA(cudaEventElapsedTime(&ms, p.begin[i], p.end[i]));
      *ms_total += ms;
    }
    *n_spans = static_cast<int64_t>(p.begin.size());
  }
  return HONK_OK;
}

extern "C" int honk_abi_version(void) { return HONK_B200_ABI_VERSION; }

extern "C" const char* honk_last_error(void) { return last_error_ref().c_str(); }

extern "C" int64_t honk_launch_count(void) { return g_launches.load(); }

extern "C" int honk_device_check(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return set_error(HONK_ERR_DEVICE, "no CUDA device (%s); honk_b200 has no CPU fallback",
                     e == cudaSuccess ? "count=0" : cudaGetErrorString(e));
  HONK_CHECK(device >= 0 && device < n, HONK_ERR_DEVICE, "device %d out of range (%d devices)", device, n);
  int major = 0, minor = 0;
  HONK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  HONK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  HONK_CHECK(major == 10, HONK_ERR_DEVICE,
             "device %d is sm_%d%d; libhonk_b200 is built for sm_100a (B200) only", device, major, minor);
  return HONK_OK;
}
