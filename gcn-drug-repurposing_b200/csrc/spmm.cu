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
#include <limits.h>

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
// coherent 128-bit load: for operands that may alias the output (e1 == y, in-place accumulation) the
// read-only (ld.global.nc) path is not defined
template <typename T>
__device__ __forceinline__ typename VecOf<T>::type vload_coherent(const T *p) {
  return *reinterpret_cast<const typename VecOf<T>::type *>(p);
}

// acc += sum_{e in [s,e)} val[e] * x[col[e], :] (* mul[col[e], :]).  The warp fetches 32 (col,val) pairs
// at a time, broadcasts them with shuffles and keeps 4 gathered rows (4*V 128-bit loads per lane) in flight.
template <typename T, int V, bool MUL>
__device__ __forceinline__ void gather_range(const int32_t *__restrict__ col, const T *__restrict__ val,
                                             const T *__restrict__ x, const T *__restrict__ mul, int32_t s,
                                             int32_t e, int lane, typename VecOf<T>::type (&acc)[V]) {
  using Vec = typename VecOf<T>::type;
  constexpr int EPL = VecOf<T>::N;
  constexpr int SLAB = 32 * EPL;
  constexpr int64_t D = (int64_t)SLAB * V;
  const int loff = lane * EPL;
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
}

// Degree-aware schedule ("plan", built once per CSR by the host mirror, graph.py):
//   rows with more than hub_threshold entries (hubs) are cut into chunks of <= chunk entries;
//   chunk_info[c] = (hub index, first entry, end entry, 0), hub_info[h] = (row, first chunk, #chunks, 0).
// Warps [0, n_chunks) each reduce one hub chunk into partial[c][:]; the last warp to finish a hub (an
// atomic ticket per hub, self-resetting) adds the partials IN CHUNK ORDER, so the result does not depend on
// scheduling.  Warps [n_chunks, n_chunks + n_rows) own one non-hub row each.  Hub chunks come first in the
// grid so the long rows start at once instead of forming the tail of the launch.
struct SpmmPlan {
  int32_t hub_threshold;  // INT32_MAX: no hubs
  int32_t n_chunks;
  const int4 *chunk_info;
  const int4 *hub_info;
  int32_t *tickets;  // [n_hubs], zero on entry and on exit
  void *partial;     // [n_chunks, d]
  // optional second output of the same launch: y2 = y ⊙ m2 (row-aligned with y).  The layer needs both
  // P = Â·X and P ⊙ X (modules/model.py:163,168); writing the product here turns the next SpMM into a plain
  // one-operand gather (half the gather traffic of multiplying inside the gather).
  const void *m2;
  void *y2;
};

// Row width d == 32 * VecOf<T>::N * V exactly (V slabs per lane).
template <typename T, int V, bool MUL, bool ADD>
__global__ void __launch_bounds__(256)
spmm_warp_row_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                     const T *__restrict__ val, int64_t n_rows, const T *__restrict__ x,
                     const T *__restrict__ mul, const T *e1, const T *__restrict__ e2, T *y,
                     const SpmmPlan plan) {
  using Vec = typename VecOf<T>::type;
  constexpr int EPL = VecOf<T>::N;      // elements per 16-byte lane load
  constexpr int SLAB = 32 * EPL;        // elements per warp-wide load
  constexpr int64_t D = (int64_t)SLAB * V;
  const int lane = threadIdx.x & 31;
  const int loff = lane * EPL;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;

  Vec acc[V];
#pragma unroll
  for (int v = 0; v < V; ++v) vzero(acc[v]);

  int64_t row;
  if (warp < plan.n_chunks) {
    // ---- one chunk of a hub row -------------------------------------------------------------------
    const int4 ci = __ldg(plan.chunk_info + warp);
    gather_range<T, V, MUL>(col, val, x, mul, ci.y, ci.z, lane, acc);
    T *pp = static_cast<T *>(plan.partial) + warp * D + loff;
#pragma unroll
    for (int v = 0; v < V; ++v) __stcg(reinterpret_cast<Vec *>(pp + v * SLAB), acc[v]);
    __threadfence();
    const int4 hi = __ldg(plan.hub_info + ci.x);
    int ticket = 0;
    if (lane == 0) ticket = atomicAdd(plan.tickets + ci.x, 1);
    ticket = __shfl_sync(0xffffffffu, ticket, 0);
    if (ticket != hi.z - 1) return;
    __threadfence();
    if (lane == 0) plan.tickets[ci.x] = 0;  // ready for the next launch
#pragma unroll
    for (int v = 0; v < V; ++v) vzero(acc[v]);
    const T *p0 = static_cast<const T *>(plan.partial) + (int64_t)hi.y * D + loff;
    for (int c = 0; c < hi.z; ++c) {
#pragma unroll
      for (int v = 0; v < V; ++v) vadd(acc[v], __ldcg(reinterpret_cast<const Vec *>(p0 + (int64_t)c * D + v * SLAB)));
    }
    row = hi.x;
  } else {
    row = warp - plan.n_chunks;
    if (row >= n_rows) return;
    const int32_t s = rowptr[row], e = rowptr[row + 1];
    if (e - s > plan.hub_threshold) return;  // a hub: handled by the chunk warps
    gather_range<T, V, MUL>(col, val, x, mul, s, e, lane, acc);
  }

  T *yr = y + row * D + loff;
#pragma unroll
  for (int v = 0; v < V; ++v) {
    if (ADD) {
      Vec a = vload_coherent<T>(e1 + row * D + loff + v * SLAB);  // e1 may be y itself
      if (e2 != nullptr) a = vmul(a, vload<T>(e2 + row * D + loff + v * SLAB));
      vadd(acc[v], a);
    }
    *reinterpret_cast<Vec *>(yr + v * SLAB) = acc[v];
    if (plan.y2 != nullptr) {
      const Vec m = vload<T>(static_cast<const T *>(plan.m2) + row * D + loff + v * SLAB);
      *reinterpret_cast<Vec *>(static_cast<T *>(plan.y2) + row * D + loff + v * SLAB) = vmul(acc[v], m);
    }
  }
}

// Any d that is a multiple of the 16-byte vector: lanes stride over the row one slab at
// a time (re-walking the row's entries per slab; only used for unusual widths).
template <typename T, bool MUL, bool ADD>
__global__ void __launch_bounds__(256)
spmm_warp_row_generic_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                             const T *__restrict__ val, int64_t n_rows, int64_t d, const T *__restrict__ x,
                             const T *__restrict__ mul, const T *e1, const T *__restrict__ e2,
                             T *y, const T *__restrict__ m2, T *__restrict__ y2) {
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
      if (ADD) {
        Vec a = vload_coherent<T>(e1 + row * d + cc);  // e1 may be y itself
        if (e2 != nullptr) a = vmul(a, vload<T>(e2 + row * d + cc));
        vadd(acc, a);
      }
      *reinterpret_cast<Vec *>(y + row * d + cc) = acc;
      if (y2 != nullptr) *reinterpret_cast<Vec *>(y2 + row * d + cc) = vmul(acc, vload<T>(m2 + row * d + cc));
    }
  }
}

template <typename T, bool MUL, bool ADD>
static int launch_spmm(const int32_t *rowptr, const int32_t *col, const T *val, int64_t n_rows, int64_t d,
                       const T *x, const T *mul, const T *e1, const T *e2, T *y, SpmmPlan plan, cudaStream_t st) {
  constexpr int SLAB = 32 * VecOf<T>::N;
  const int T_ = 256;
  const bool fixed = (d == SLAB || d == 2 * SLAB || d == 4 * SLAB);
  if (!fixed) {  // unusual widths: no hub splitting (the generic kernel walks every row)
    plan.hub_threshold = INT32_MAX;
    plan.n_chunks = 0;
  }
  const unsigned grid = (unsigned)ceil_div<int64_t>((n_rows + plan.n_chunks) * 32, T_);
  if (d == SLAB)
    spmm_warp_row_kernel<T, 1, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, x, mul, e1, e2, y, plan);
  else if (d == 2 * SLAB)
    spmm_warp_row_kernel<T, 2, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, x, mul, e1, e2, y, plan);
  else if (d == 4 * SLAB)
    spmm_warp_row_kernel<T, 4, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, x, mul, e1, e2, y, plan);
  else
    spmm_warp_row_generic_kernel<T, MUL, ADD><<<grid, T_, 0, st>>>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y,
                                                                  static_cast<const T *>(plan.m2),
                                                                  static_cast<T *>(plan.y2));
  GDR_LAUNCHED();
  return GDR_OK;
}

static SpmmPlan no_plan() {
  SpmmPlan p{};
  p.hub_threshold = INT32_MAX;
  return p;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

static int spmm_f32_impl(const int32_t *rowptr, const int32_t *col, const float *val, int64_t n_rows, int64_t d,
                         const float *x, const float *mul, const float *e1, const float *e2, float *y,
                         const float *m2, float *y2, int32_t hub_threshold, int64_t n_hubs, const int32_t *hub_info,
                         int64_t n_chunks, const int32_t *chunk_info, int32_t *tickets, float *partial,
                         gdr_stream_t stream) {
  GDR_CHECK_ARG((m2 == nullptr) == (y2 == nullptr), GDR_E_BADARG, "spmm_f32: m2 and y2 come together");
  GDR_CHECK_ARG((((uintptr_t)m2 | (uintptr_t)y2) & 15) == 0, GDR_E_SHAPE, "spmm_f32: pointers must be 16-byte aligned");
  GDR_CHECK_ARG(n_rows >= 0 && d > 0, GDR_E_BADARG, "spmm_f32: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "spmm_f32: d=%lld must be a multiple of 4", (long long)d);
  GDR_CHECK_ARG(rowptr && x && y, GDR_E_BADARG, "spmm_f32: null pointer");
  GDR_CHECK_ARG(e1 != nullptr || e2 == nullptr, GDR_E_BADARG, "spmm_f32: e2 needs e1");
  GDR_CHECK_ARG((((uintptr_t)x | (uintptr_t)y | (uintptr_t)mul | (uintptr_t)e1 | (uintptr_t)e2) & 15) == 0,
                GDR_E_SHAPE, "spmm_f32: pointers must be 16-byte aligned");
  if (n_rows == 0) return GDR_OK;
  GDR_CHECK_ARG(n_hubs >= 0 && n_chunks >= 0 && n_chunks < (int64_t)INT32_MAX, GDR_E_BADARG,
                "spmm_f32: bad plan sizes");
  GDR_CHECK_ARG(n_chunks == 0 || (hub_threshold > 0 && n_hubs > 0 && hub_info && chunk_info && tickets && partial),
                GDR_E_BADARG, "spmm_f32: incomplete plan");
  cudaStream_t st = as_stream(stream);
  SpmmPlan plan = no_plan();
  if (n_chunks > 0) {
    plan.hub_threshold = hub_threshold;
    plan.n_chunks = (int32_t)n_chunks;
    plan.chunk_info = reinterpret_cast<const int4 *>(chunk_info);
    plan.hub_info = reinterpret_cast<const int4 *>(hub_info);
    plan.tickets = tickets;
    plan.partial = partial;
  }
  plan.m2 = m2;
  plan.y2 = y2;
  if (mul && e1) return launch_spmm<float, true, true>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, plan, st);
  if (mul) return launch_spmm<float, true, false>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, plan, st);
  if (e1) return launch_spmm<float, false, true>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, plan, st);
  return launch_spmm<float, false, false>(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, plan, st);
}

int gdr_spmm_csr_planned_f32(const int32_t *rowptr, const int32_t *col, const float *val, int64_t n_rows,
                             int64_t d, const float *x, const float *mul, const float *e1, const float *e2,
                             float *y, int32_t hub_threshold, int64_t n_hubs, const int32_t *hub_info,
                             int64_t n_chunks, const int32_t *chunk_info, int32_t *tickets, float *partial,
                             gdr_stream_t stream) {
  return spmm_f32_impl(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, nullptr, nullptr, hub_threshold, n_hubs,
                       hub_info, n_chunks, chunk_info, tickets, partial, stream);
}

// y = Â·x (+ e1 ⊙ e2) and, in the same launch, y2 = y ⊙ m2.  Plan arguments as above (n_chunks = 0: no plan).
int gdr_spmm_csr_dual_f32(const int32_t *rowptr, const int32_t *col, const float *val, int64_t n_rows, int64_t d,
                          const float *x, const float *e1, const float *e2, float *y, const float *m2, float *y2,
                          int32_t hub_threshold, int64_t n_hubs, const int32_t *hub_info, int64_t n_chunks,
                          const int32_t *chunk_info, int32_t *tickets, float *partial, gdr_stream_t stream) {
  GDR_CHECK_ARG(m2 && y2, GDR_E_BADARG, "spmm_dual: m2 / y2 missing");
  return spmm_f32_impl(rowptr, col, val, n_rows, d, x, nullptr, e1, e2, y, m2, y2, hub_threshold, n_hubs, hub_info,
                       n_chunks, chunk_info, tickets, partial, stream);
}

int gdr_spmm_csr_f32(const int32_t *rowptr, const int32_t *col, const float *val, int64_t n_rows, int64_t d,
                     const float *x, const float *mul, const float *e1, const float *e2, float *y,
                     gdr_stream_t stream) {
  return gdr_spmm_csr_planned_f32(rowptr, col, val, n_rows, d, x, mul, e1, e2, y, 0, 0, nullptr, 0, nullptr,
                                  nullptr, nullptr, stream);
}

int gdr_spmm_csr_f64(const int32_t *rowptr, const int32_t *col, const double *val, int64_t n_rows, int64_t d,
                     const double *x, double *y, gdr_stream_t stream) {
  GDR_CHECK_ARG(n_rows >= 0 && d > 0, GDR_E_BADARG, "spmm_f64: bad sizes");
  GDR_CHECK_ARG(d % 2 == 0, GDR_E_SHAPE, "spmm_f64: d=%lld must be a multiple of 2", (long long)d);
  GDR_CHECK_ARG(rowptr && x && y, GDR_E_BADARG, "spmm_f64: null pointer");
  GDR_CHECK_ARG((((uintptr_t)x | (uintptr_t)y) & 15) == 0, GDR_E_SHAPE,
                "spmm_f64: pointers must be 16-byte aligned");
  if (n_rows == 0) return GDR_OK;
  return launch_spmm<double, false, false>(rowptr, col, val, n_rows, d, x, nullptr, nullptr, nullptr, y, no_plan(),
                                           as_stream(stream));
}

}  // extern "C"
