// Error plumbing, device check and launch accounting for libhonk_b200.so.
#include "common.cuh"

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

}  // namespace honk

using namespace honk;

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
