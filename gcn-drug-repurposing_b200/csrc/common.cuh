// common.cuh — shared helpers for libgdr.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/gdr.h"

namespace gdr {

// ---- error plumbing ---------------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launch(int n = 1);

#define GDR_CHECK_ARG(cond, code, ...)    \
  do {                                    \
    if (!(cond)) {                        \
      ::gdr::set_error(__VA_ARGS__);      \
      return (code);                      \
    }                                     \
  } while (0)

#define GDR_CUDA(expr)                                                          \
  do {                                                                          \
    cudaError_t e__ = (expr);                                                   \
    if (e__ != cudaSuccess) {                                                   \
      ::gdr::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                       __FILE__, __LINE__);                                     \
      return (int)e__;                                                          \
    }                                                                           \
  } while (0)

// call after every kernel launch
#define GDR_LAUNCHED()                  \
  do {                                  \
    ::gdr::count_launch();              \
    GDR_CUDA(cudaPeekAtLastError());    \
  } while (0)

inline cudaStream_t as_stream(gdr_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;  // B200

template <typename T>
__host__ __device__ constexpr T ceil_div(T a, T b) {
  return (a + b - 1) / b;
}

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

// carve aligned slices out of a caller-owned workspace
struct Carver {
  char *base;
  int64_t off;
  explicit Carver(void *p) : base(static_cast<char *>(p)), off(0) {}
  template <typename T>
  T *take(int64_t count) {
    off = align_up(off, 256);
    T *p = reinterpret_cast<T *>(base + off);
    off += count * (int64_t)sizeof(T);
    return p;
  }
  int64_t used() const { return align_up(off, 256); }
};

// ---- device helpers ------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// 128-bit read-only gathers (LDG.E.128 through the non-coherent path)
__device__ __forceinline__ float4 ldg_f4(const float *p) {
  return __ldg(reinterpret_cast<const float4 *>(p));
}
__device__ __forceinline__ double2 ldg_d2(const double *p) {
  return __ldg(reinterpret_cast<const double2 *>(p));
}

}  // namespace gdr
