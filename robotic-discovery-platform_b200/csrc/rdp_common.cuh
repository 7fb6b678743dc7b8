// Shared host/device helpers for the rdp_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rdp_b200.h"

namespace rdp {

// Thread-local error text returned by rdp_last_error().
void set_error(const char* fmt, ...);

#define RDP_CUDA_TRY(expr)                                                                          \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess) {                                                                        \
      ::rdp::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return RDP_ERR_CUDA;                                                                          \
    }                                                                                               \
  } while (0)

#define RDP_REQUIRE(cond, code, ...) \
  do {                               \
    if (!(cond)) {                   \
      ::rdp::set_error(__VA_ARGS__); \
      return (code);                 \
    }                                \
  } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over a caller-provided workspace.
struct Carver {
  char* base;
  size_t off;
  explicit Carver(void* p) : base(static_cast<char*>(p)), off(0) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
  size_t total() const { return align_up(off, 256); }
};

}  // namespace rdp
