// peaks.cu — measurement utility, not part of the product ABI (not declared in include/gdr.h; bench.py binds it by
// name): the L2 -> SM read bandwidth of THIS device, the denominator of the MSI-scale SpMM's roofline (its operand X,
// 31 MB, lives in L2, so HBM bandwidth is not what bounds it).
//
// Every CTA streams the whole buffer `reps` times with 128-bit non-allocating loads, eight in flight per thread,
// starting at a CTA-specific offset so that the SMs are not all on the same lines at once.  The buffer is meant to be
// L2-sized (default 32 MiB): larger than every L1, smaller than the L2, so after the first touch all traffic is
// L2 -> SM.  Bytes moved = gridDim.x * bytes * reps.
#include "common.cuh"

namespace gdr {

constexpr int L2R_THREADS = 512, L2R_UNROLL = 8;

__global__ void __launch_bounds__(L2R_THREADS)
l2_read_kernel(const uint4 *__restrict__ buf, int64_t n16, int reps, unsigned int *__restrict__ sink) {
  const int64_t start = ((int64_t)blockIdx.x * 7919 * L2R_THREADS * L2R_UNROLL) % n16;
  unsigned int acc = 0;
  for (int r = 0; r < reps; ++r) {
    for (int64_t base = 0; base < n16; base += L2R_THREADS * L2R_UNROLL) {
      uint4 v[L2R_UNROLL];
#pragma unroll
      for (int u = 0; u < L2R_UNROLL; ++u) {
        int64_t i = start + base + u * L2R_THREADS + threadIdx.x;
        if (i >= n16) i -= n16;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                     : "l"(buf + i));
      }
#pragma unroll
      for (int u = 0; u < L2R_UNROLL; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
  }
  if (acc == 0x9e3779b9u && reps < 0) sink[0] = acc;  // keeps the loads alive, never taken
}

}  // namespace gdr

using namespace gdr;

extern "C" {

// buf: device, `bytes` a multiple of 65536; n_ctas <= 0: 4 per SM.  Returns the number of bytes the launch reads
// (n_ctas * bytes * reps) or a negative GDR_E_* code.
int64_t gdr_bench_l2_read(const void *buf, int64_t bytes, int reps, int n_ctas, void *sink, gdr_stream_t stream) {
  GDR_CHECK_ARG(buf && sink && reps > 0, GDR_E_BADARG, "bench_l2_read: bad args");
  GDR_CHECK_ARG(bytes > 0 && bytes % (16 * L2R_THREADS * L2R_UNROLL) == 0, GDR_E_SHAPE,
                "bench_l2_read: bytes must be a positive multiple of %d", 16 * L2R_THREADS * L2R_UNROLL);
  const int grid = n_ctas > 0 ? n_ctas : 4 * kNumSMs;
  l2_read_kernel<<<grid, L2R_THREADS, 0, as_stream(stream)>>>(static_cast<const uint4 *>(buf), bytes / 16, reps,
                                                            static_cast<unsigned int *>(sink));
  count_launch();
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    set_error("bench_l2_read launch failed: %s", cudaGetErrorString(e));
    return -(int64_t)e - 1000;
  }
  return (int64_t)grid * bytes * reps;
}

}  // extern "C"
