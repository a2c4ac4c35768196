// Shared helpers for libhonk_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/honk_b200.h"

namespace honk {

// ---- error plumbing (thread-local message, integer status across the C ABI) ----------------
std::string& last_error_ref();
int set_error(int code, const char* fmt, ...);

#define HONK_CUDA(call)                                                                         \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess)                                                                     \
      return ::honk::set_error(HONK_ERR_CUDA, "%s failed: %s (%s:%d)", #call,                   \
                               cudaGetErrorString(e__), __FILE__, __LINE__);                    \
  } while (0)

#define HONK_CHECK(cond, code, ...)                                                             \
  do {                                                                                          \
    if (!(cond)) return ::honk::set_error(code, __VA_ARGS__);                                   \
  } while (0)

#define HONK_TRY(expr)                                                                          \
  do {                                                                                          \
    int rc__ = (expr);                                                                          \
    if (rc__ != 0) return rc__;                                                                 \
  } while (0)

// ---- launch accounting (bench.py reports gpu_launches from this) ---------------------------
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define HONK_LAUNCH_CHECK()                                                                     \
  do {                                                                                          \
    ::honk::count_launch();                                                                     \
    cudaError_t e__ = cudaGetLastError();                                                       \
    if (e__ != cudaSuccess)                                                                     \
      return ::honk::set_error(HONK_ERR_CUDA, "kernel launch failed: %s (%s:%d)",               \
                               cudaGetErrorString(e__), __FILE__, __LINE__);                    \
  } while (0)

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Bump allocator over the caller-provided workspace.
struct Arena {
  char* base;
  int64_t size;
  int64_t off = 0;
  bool dry;  // size-only pass
  Arena(void* b, int64_t s) : base(static_cast<char*>(b)), size(s), dry(b == nullptr) {}
  template <typename T>
  T* take(int64_t n) {
    off = align_up(off, 256);
    T* p = dry ? nullptr : reinterpret_cast<T*>(base + off);
    off += n * static_cast<int64_t>(sizeof(T));
    return p;
  }
  bool ok() const { return dry || off <= size; }
};

int num_sms(int device);

// Device-time spans (enabled by honk_profile_enable); no-ops otherwise.
struct ProfSpan {
  int slot = -1;
  cudaStream_t stream;
  ProfSpan(const char* name, cudaStream_t s);
  ~ProfSpan();
};

}  // namespace honk
