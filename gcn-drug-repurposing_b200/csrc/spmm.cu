// spmm.cu — CSR SpMM  Y = Â·X, the aggregation of the GCN and of the diffusion power
// iteration.  HBM/L2-bound gather: one warp per output row, every lane owns 16 bytes
// (float4 / double2) of each 512-byte slab of the feature row, so a gathered row is
// read with fully coalesced LDG.E.128.  Column ids and values of a row are fetched 32 at
// a time by the whole warp and broadcast with shuffles; the gather loop is unrolled 4x
// to keep >= 4 (x V) independent 128-bit loads in flight per lane.
//
// Replaces: torch.sparse.mm(adj, features)           modules/model.py:163
//           torch.sparse.mm(adj, torch.mul(Ax, x))   modules/model.py:168-169 (mul folded in)
//           autograd of both (Âᵀ·G on the transposed CSR)
//           x * M                                     multiscale/diff_prof/diffusion_profiles.py:85
#include "common.cuh"

namespace gdr {

template <typename T>
struct VecOf;
template <>
struct VecOf<float> {
  using type = float4;
  static constexpr int N = 4;
};
template <>
struct VecOf<double> {
  using type = double2;
  static constexpr int N = 2;
};

__device__ __forceinline__ void vfma(float4 &a, float s, const float4 &x) {
  a.x = fmaf(s, x.x, a.x);
  a.y = fmaf(s, x.y, a.y);
  a.z = fmaf(s, x.z, a.z);
  a.w = fmaf(s, x.w, a.w);
}
__device__ __forceinline__ void vfma(double2 &a, double s, const double2 &x) {
  a.x = fma(s, x.x, a.x);
  a.y = fma(s, x.y, a.y);
}
__device__ __forceinline__ float4 vmul(const float4 &a, const float4 &b) {
  return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w);
}
__device__ __forceinline__ double2 vmul(const double2 &a, const double2 &b) {
  return make_double2(a.x * b.x, a.y * b.y);
}
__device__ __forceinline__ void vzero(float4 &a) { a = make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void vzero(double2 &a) { a = make_double2(0.0, 0.0); }
__device__ __forceinline__ void vadd(float4 &a, const float4 &b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
}
__device__ __forceinline__ void vadd(double2 &a, const double2 &b) {
  a.x += b.x; a.y += b.y;
}

template <typename T>
__device__ __forceinline__ typename VecOf<T>::type vload(const T *p) {
  return __ldg(reinterpret_cast<const typename VecOf<T>::type *>(p));
}

// One warp per row.  Row width d == 32 * VecOf<T>::N * V exactly (V slabs per lane).
template <typename T, int V, bool MUL, bool ADD>
__global__ void __launch_bounds__(256)
spmm_warp_row_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                     const T *__restrict__ val, int64_t n_rows, const T *__restrict__ x,
                     const T *__restrict__ mul, const T *__restrict__ e1, const T *__restrict__ e2,
                     T *__restrict__ y) {
  using Vec = typename VecOf<T>::type;
  constexpr int EPL = VecOf<T>::N;      // elements per 16-byte lane load
  constexpr int SLAB = 32 * EPL;        // elements per warp-wide load
  constexpr int64_t D = (int64_t)SLAB * V;
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n_rows) return;
  const int32_t s = rowptr[row], e = rowptr[row + 1];
  const int loff = lane * EPL;

  Vec acc[V];
#pragma unroll
  for (int v = 0; v < V; ++v) vzero(acc[v]);

  for (int32_t base = s; base < e; base += 32) {
    const int32_t idx = base + lane;
    int32_t c = 0;
    T w = (T)0;
    if (idx < e) {
      c = __ldg(col + idx);
      w = __ldg(val + idx);
    }
    const int cnt = min(32, e - base);
    int j = 0;
    for (; j + 4 <= cnt; j += 4) {
      int32_t cj[4];
      T wj[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        cj[u] = __shfl_sync(0xffffffffu, c, j + u);
        wj[u] = __shfl_sync(0xffffffffu, w, j + u);
      }
      Vec xv[4][V];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const T *xr = x + (int64_t)cj[u] * D + loff;
#pragma unroll
        for (int v = 0; v < V; ++v) xv[u][v] = vload<T>(xr + v * SLAB);
      }
      if (MUL) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const T *mr = mul + (int64_t)cj[u] * D + loff;
#pragma unroll
          for (int v = 0; v < V; ++v) xv[u][v] = vmul(xv[u][v], vload<T>(mr + v * SLAB));
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
#pragma unroll
        for (int v = 0; v < V; ++v) vfma(acc[v], wj[u], xv[u][v]);
      }
    }
    for (; j < cnt; ++j) {
      const int32_t c1 = __shfl_sync(0xffffffffu, c, j);
      const T w1 = __shfl_sync(0xffffffffu, w, j);
      const T *xr = x + (int64_t)c1 * D + loff;
#pragma unroll
      for (int v = 0; v < V; ++v) {
        Vec t = vload<T>(xr + v * SLAB);
        if (MUL) t = vmul(t, vload<T>(mul + (int64_t)c1 * D + loff + v * SLAB));
        vfma(acc[v], w1, t);
      }
    }
  }

  T *yr = y + row * D + loff;
#pragma unroll
  for (int v = 0; v < V; ++v) {
    if (ADD) {
      Vec a = vload<T>(e1 + row * D + loff + v * SLAB);
      Vec b = vload<T>(e2 + row * D + loff + v * SLAB);
      vadd(acc[v], vmul(a, b));
    }
    *reinterpret_cast<Vec *>(yr + v * SLAB) = acc[v];
  }
}

// Any d that is a multiple of the 16-byte vector: lanes stride over the row one slab at
// a time (re-walking the row's entries per slab; only used for unusual widths).
template <typename T, bool MUL, bool ADD>
__global__ void __launch_bounds__(256)
spmm_warp_row_generic_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                             const T *__restrict__ val, int64_t n_rows, int64_t d, const T *__restrict__ x,
                             const T *__restrict__ mul, const T *__restrict__ e1, const T *__restrict__ e2,
                             T *__restrict__ y) {
  using Vec = typename VecOf<T>::type;
  constexpr int EPL = VecOf<T>::N;
  constexpr int SLAB = 32 * EPL;
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n_rows) return;
  const int32_t s = rowptr[row], e = rowptr[row + 1];
  for (int64_t c0 = 0; c0 < d; c0 += SLAB) {
    const int64_t cc = c0 + lane * EPL;
    const bool live = cc < d;
    Vec acc;
    vzero(acc);
    for (int32_t base = s; base < e; base += 32) {
      const int32_t idx = base + lane;
      int32_t c = 0;
      T w = (T)0;
      if (idx < e) {
        c = __ldg(col + idx);
        w = __ldg(val + idx);
      }
      const int cnt = min(32, e - base);
      for (int j = 0; j < cnt; ++j) {
        const int32_t c1 = __shfl_sync(0xffffffffu, c, j);
        const T w1 = __shfl_sync(0xffffffffu, w, j);
        if (live) {
          Vec t = vload<T>(x + (int64_t)c1 * d + cc);
          if (MUL) t = vmul(t, vload<T>(mul + (int64_t)c1 * d + cc));
          vfma(acc, w1, t);
        }
      }
    }
    if (live) {
      if (ADD) vadd(acc, vmul(vload<T>(e1 + row * d + cc), vload<T>(e2 + row * d + cc)));
      *reinterpret_cast<Vec *>(y + row * d + cc) = acc;
    }
  }
}

template <typename T, bool MUL, bool ADD>
static int launch_spmm(const int32_t *rowptr, const int32_t *col, const T *val, int64_t n_rows, int64_t d,
                       const T *x, const T *mul, const T *e1, const T *e2, T *y, cudaStream_t st) {
  constexpr int SLAB = 32 * VecOf<T>::N;
  const int T_ = 256;
  const unsigned grid = (unsigned)ceil_div<int64_t>(n_rows * 32, T_);
  if (d == SLAB)
    spmm_warp_row_kernel<T, 1, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, x, mul, e1, e2, y);
  else if (d == 2 * SLAB)
    spmm_warp_row_kernel<T, 2, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, x, mul, e1, e2, y);
  else if (d == 4 * SLAB)
    spmm_warp_row_kernel<T, 4, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, x, mul, e1, e2, y);
  else
    spmm_warp_row_generic_kernel<T, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, d, x, mul, e1, e2,
                                                                  y);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int gdr_spmm_csr_f32(const int32_t *rowptr, const int32_t *col, const float *val, int64_t n_rows, int64_t d,
                     const float *x, const float *mul, const float *e1, const float *e2, float *y,
                     gdr_stream_t stream) {
  GDR_CHECK_ARG(n_rows >= 0 && d > 0, GDR_E_BADARG, "spmm_f32: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "spmm_f32: d=%lld must be a multiple of 4", (long long)d);
  GDR_CHECK_ARG(rowptr && x && y, GDR_E_BADARG, "spmm_f32: null pointer");
  GDR_CHECK_ARG((e1 == nullptr) == (e2 == nullptr), GDR_E_BADARG, "spmm_f32: e1/e2 must come together");
  GDR_CHECK_ARG((((uintptr_t)x | (uintptr_t)y | (uintptr_t)mul | (uintptr_t)e1 | (uintptr_t)e2) & 15) == 0,
                GDR_E_SHAPE, "spmm_f32: pointers must be 16-byte aligned");
  if (n_rows == 0) return GDR_OK;
  cudaStream_t st = as_stream(stream);
  if (mul && e1) return launch_spmm<float, true, true>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, st);
  if (mul) return launch_spmm<float, true, false>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, st);
  if (e1) return launch_spmm<float, false, true>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, st);
  return launch_spmm<float, false, false>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, st);
}

int gdr_spmm_csr_f64(const int32_t *rowptr, const int32_t *col, const double *val, int64_t n_rows, int64_t d,
                     const double *x, double *y, gdr_stream_t stream) {
  GDR_CHECK_ARG(n_rows >= 0 && d > 0, GDR_E_BADARG, "spmm_f64: bad sizes");
  GDR_CHECK_ARG(d % 2 == 0, GDR_E_SHAPE, "spmm_f64: d=%lld must be a multiple of 2", (long long)d);
  GDR_CHECK_ARG(rowptr && x && y, GDR_E_BADARG, "spmm_f64: null pointer");
  GDR_CHECK_ARG((((uintptr_t)x | (uintptr_t)y) & 15) == 0, GDR_E_SHAPE,
                "spmm_f64: pointers must be 16-byte aligned");
  if (n_rows == 0) return GDR_OK;
  return launch_spmm<double, false, false>(rowptr, col, val, n_rows, d, x, nullptr, nullptr, nullptr, y,
                                           as_stream(stream));
}

}  // extern "C"
