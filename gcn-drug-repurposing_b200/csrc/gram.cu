// gram.cu — the two N×N Gram reductions of the reference, without ever storing N×N:
//   gdr_gram_topk_f64        gen_graph: x_sim = Xᵀ·X (fp64) + np.argpartition(x_sim, -k, 1)[:, -k:]
//                            helpers/helper.py:39,44
//   gdr_gram_percentile_f32  beta = np.percentile(flatten(H·Hᵀ), q)  (fp32 Gram, linear interpolation)
//                            train.py:165-167
#include <math.h>

#include "gemm_tc.cuh"

namespace gdr {

// =============================================================================================
// percentile: 3-pass radix select (11+11+10 bits) over recomputed Gram tiles + an optional
// "smallest key above" pass when the two neighbouring order statistics differ.
// =============================================================================================
__global__ void select_init_kernel(SelectState *st, unsigned long long rank, unsigned long long *hist) {
  if (threadIdx.x == 0) {
    st->prefix = 0u;
    st->mask = 0u;
    st->rank = rank;
    st->count = 0ull;
    st->min_gt = 0xffffffffu;
  }
  for (int i = threadIdx.x; i < GHIST_BINS; i += blockDim.x) hist[i] = 0ull;
}

// pick the bin holding `rank`, descend into it, clear the histogram for the next pass
__global__ void select_descend_kernel(SelectState *st, unsigned long long *hist, int shift, int bins) {
  if (threadIdx.x == 0) {
    unsigned long long cum = 0ull, rank = st->rank;
    int b = 0;
    for (; b < bins; ++b) {
      const unsigned long long c = hist[b];
      if (rank < cum + c) break;
      cum += c;
    }
    if (b == bins) b = bins - 1;  // cannot happen when rank < total
    st->prefix |= ((unsigned int)b) << shift;
    st->mask |= ((unsigned int)(bins - 1)) << shift;
    st->rank = rank - cum;
    st->count = hist[b];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < GHIST_BINS; i += blockDim.x) hist[i] = 0ull;
}

// =============================================================================================
// top-k: fp64 SIMT Gram, 64x64 tiles, one CTA owns 64 rows and sweeps all column tiles keeping a
// sorted top-k list per row in shared memory.
// =============================================================================================
constexpr int TK_BM = 64, TK_BN = 64, TK_BK = 16, TK_THREADS = 256, TK_PAD = 2, TK_MAXK = 32, TK_QCAP = 64;

__global__ void __launch_bounds__(TK_THREADS)
gram_topk_f64_kernel(const double *__restrict__ x, int64_t n, int64_t d, int k, int32_t *__restrict__ idx_out,
                     double *__restrict__ val_out) {
  __shared__ __align__(16) double As[TK_BK][TK_BM + TK_PAD];
  __shared__ __align__(16) double Bs[TK_BK][TK_BN + TK_PAD];
  __shared__ double thr[TK_BM];
  __shared__ int q_cnt[TK_BM];
  // dynamic shared memory (72 KB, opted in by the launcher): per-row top-k lists and candidate queues
  extern __shared__ __align__(16) unsigned char tk_dyn[];
  double(*top_val)[TK_MAXK] = reinterpret_cast<double(*)[TK_MAXK]>(tk_dyn);
  double(*q_val)[TK_QCAP] = reinterpret_cast<double(*)[TK_QCAP]>(tk_dyn + sizeof(double) * TK_BM * TK_MAXK);
  int32_t(*top_idx)[TK_MAXK] =
      reinterpret_cast<int32_t(*)[TK_MAXK]>(tk_dyn + sizeof(double) * TK_BM * (TK_MAXK + TK_QCAP));
  int32_t(*q_idx)[TK_QCAP] = reinterpret_cast<int32_t(*)[TK_QCAP]>(
      tk_dyn + sizeof(double) * TK_BM * (TK_MAXK + TK_QCAP) + sizeof(int32_t) * TK_BM * TK_MAXK);

  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;  // thread tile: rows ty*4..+3, cols tx*4..+3
  const int64_t m0 = (int64_t)blockIdx.x * TK_BM;
  const double NEG_INF = -INFINITY;

  for (int i = t; i < TK_BM * TK_MAXK; i += TK_THREADS) {
    top_val[i / TK_MAXK][i % TK_MAXK] = NEG_INF;
    top_idx[i / TK_MAXK][i % TK_MAXK] = -1;
  }
  if (t < TK_BM) {
    thr[t] = NEG_INF;
    q_cnt[t] = 0;
  }
  __syncthreads();

  // loader mapping: 64 rows x 16 k per tile = 1024 doubles, 4 per thread (2 x double2)
  const int lr = t >> 2;         // 0..63 row inside the tile
  const int lk = (t & 3) * 4;    // 0,4,8,12

  for (int64_t n0 = 0; n0 < n; n0 += TK_BN) {
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;

    for (int64_t k0 = 0; k0 < d; k0 += TK_BK) {
      double a[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
      const int64_t ar = m0 + lr, br = n0 + lr;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t kk = k0 + lk + q;
        if (kk < d) {
          if (ar < n) a[q] = __ldg(x + ar * d + kk);
          if (br < n) b[q] = __ldg(x + br * d + kk);
        }
      }
      __syncthreads();  // previous chunk fully consumed
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        As[lk + q][lr] = a[q];
        Bs[lk + q][lr] = b[q];
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < TK_BK; ++kk) {
        const double2 a01 = *reinterpret_cast<const double2 *>(&As[kk][ty * 4]);
        const double2 a23 = *reinterpret_cast<const double2 *>(&As[kk][ty * 4 + 2]);
        const double2 b01 = *reinterpret_cast<const double2 *>(&Bs[kk][tx * 4]);
        const double2 b23 = *reinterpret_cast<const double2 *>(&Bs[kk][tx * 4 + 2]);
        const double av[4] = {a01.x, a01.y, a23.x, a23.y};
        const double bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
      }
    }

    // candidates above the row threshold go to the row's queue
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = ty * 4 + i;
      if (m0 + r >= n) continue;
      const double th = thr[r];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t c = n0 + tx * 4 + j;
        if (c < n && acc[i][j] > th) {
          const int slot = atomicAdd(&q_cnt[r], 1);  // <= 64 candidates per row per tile by construction
          q_val[r][slot] = acc[i][j];
          q_idx[r][slot] = (int32_t)c;
        }
      }
    }
    __syncthreads();
    if (t < TK_BM) {
      const int cnt = q_cnt[t];
      for (int s = 0; s < cnt; ++s) {
        const double v = q_val[t][s];
        const int32_t c = q_idx[t][s];
        if (!(v > top_val[t][k - 1])) continue;
        int pos = k - 1;  // insertion into the descending list
        while (pos > 0 && top_val[t][pos - 1] < v) {
          top_val[t][pos] = top_val[t][pos - 1];
          top_idx[t][pos] = top_idx[t][pos - 1];
          --pos;
        }
        top_val[t][pos] = v;
        top_idx[t][pos] = c;
      }
      q_cnt[t] = 0;
      thr[t] = top_val[t][k - 1];
    }
    __syncthreads();
  }

  for (int i = t; i < TK_BM * k; i += TK_THREADS) {
    const int r = i / k, j = i % k;
    if (m0 + r < n) {
      idx_out[(m0 + r) * k + j] = top_idx[r][j];
      val_out[(m0 + r) * k + j] = top_val[r][j];
    }
  }
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int gdr_gram_topk_f64(const double *x, int64_t n, int64_t d, int64_t k, int32_t *idx_out, double *val_out,
                      gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && d > 0 && x && idx_out && val_out, GDR_E_BADARG, "gram_topk: bad args");
  GDR_CHECK_ARG(k >= 1 && k <= TK_MAXK, GDR_E_SHAPE, "gram_topk: k=%lld outside 1..%d", (long long)k, TK_MAXK);
  GDR_CHECK_ARG(k <= n, GDR_E_SHAPE, "gram_topk: k=%lld exceeds n=%lld (np.argpartition raises too)", (long long)k,
                (long long)n);
  GDR_CHECK_ARG(n < (int64_t)INT32_MAX, GDR_E_SHAPE, "gram_topk: int32 index range exceeded");
  const int dyn = (int)((sizeof(double) + sizeof(int32_t)) * TK_BM * (TK_MAXK + TK_QCAP));
  GDR_CUDA(cudaFuncSetAttribute(gram_topk_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  gram_topk_f64_kernel<<<(unsigned)ceil_div<int64_t>(n, TK_BM), TK_THREADS, dyn, as_stream(stream)>>>(
      x, n, d, (int)k, idx_out, val_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int64_t gdr_gram_percentile_workspace_bytes(int64_t n, int64_t d) {
  if (n <= 0 || d <= 0) return GDR_E_BADARG;
  return align_up(GHIST_BINS * 8, 256) + 1024;
}

int64_t gdr_gram_percentile_tc_workspace_bytes(int64_t n, int64_t d) {
  if (n <= 0 || d <= 0) return GDR_E_BADARG;
  return align_up(GHIST_BINS * 8, 256) + 1024 + 2 * align_up(n * d * 4, 256);
}

static int gram_percentile_impl(const float *h, int64_t n, int64_t d, double q_percent, float *beta_out_host,
                                void *workspace, int64_t workspace_bytes, gdr_stream_t stream, bool use_tc) {
  GDR_CHECK_ARG(n > 0 && d > 0 && h && beta_out_host && workspace, GDR_E_BADARG, "gram_percentile: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "gram_percentile: d must be a multiple of 4");
  GDR_CHECK_ARG(q_percent >= 0.0 && q_percent <= 100.0, GDR_E_BADARG, "gram_percentile: q outside [0,100]");
  GDR_CHECK_ARG(workspace_bytes >= (use_tc ? gdr_gram_percentile_tc_workspace_bytes(n, d)
                                           : gdr_gram_percentile_workspace_bytes(n, d)),
                GDR_E_WORKSPACE, "gram_percentile: workspace too small");
  GDR_CHECK_ARG((double)n * (double)n < 9.0e15, GDR_E_SHAPE, "gram_percentile: n*n exceeds the exact fp64 index range");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  unsigned long long *hist = cv.take<unsigned long long>(GHIST_BINS);
  SelectState *sel = cv.take<SelectState>(1);
  float *hhi = nullptr, *hlo = nullptr;
  if (use_tc) {
    hhi = cv.take<float>(n * d);
    hlo = cv.take<float>(n * d);
    const int rc = tc_split(h, d, d, nullptr, 0, 0, n, hhi, hlo, st);
    if (rc) return rc;
  }

  // numpy's virtual index for method='linear' (alpha = beta = 1), same expression and rounding:
  //   n*q + (alpha + q*(1 - alpha - beta)) - 1          numpy/lib/_function_base_impl.py
  const double total = (double)n * (double)n;
  const double q = q_percent / 100.0;
  const double virt = total * q + (1.0 + q * (1.0 - 1.0 - 1.0)) - 1.0;
  double prev = floor(virt);
  if (prev < 0.0) prev = 0.0;
  if (prev > total - 1.0) prev = total - 1.0;
  const double gamma = (virt - prev > 0.0 && prev < total - 1.0) ? virt - prev : 0.0;
  const unsigned long long rank_lo = (unsigned long long)prev;

  select_init_kernel<<<1, 256, 0, st>>>(sel, rank_lo, hist);
  GDR_LAUNCHED();
  GemmParams g{};
  g.a[0] = h; g.lda[0] = d; g.b[0] = h; g.ldb[0] = d; g.kseg[0] = d; g.nseg = 1;
  g.m = n; g.n = n;
  g.sel = sel; g.hist = hist; g.sel_out = sel;
  const unsigned tiles = (unsigned)ceil_div<int64_t>(n, GBM);
  dim3 grid(tiles, tiles, 1);
  const int shifts[3] = {21, 10, 0};
  const int bins[3] = {2048, 2048, 1024};
  for (int pass = 0; pass < 3; ++pass) {
    g.shift = shifts[pass];
    g.bin_mask = (unsigned int)(bins[pass] - 1);
    if (use_tc) {
      const int rc = tc_gemm(EPI_HIST, hhi, hlo, hhi, hlo, n, n, d, d, d, d, 1, g, st);
      if (rc) return rc;
    } else {
      gemm_f32_kernel<true, true, EPI_HIST><<<grid, GTHREADS, 0, st>>>(g);
      GDR_LAUNCHED();
    }
    select_descend_kernel<<<1, 256, 0, st>>>(sel, hist, shifts[pass], bins[pass]);
    GDR_LAUNCHED();
  }
  SelectState hs;
  GDR_CUDA(cudaMemcpyAsync(&hs, sel, sizeof(SelectState), cudaMemcpyDeviceToHost, st));
  GDR_CUDA(cudaStreamSynchronize(st));
  const float v_lo = f32_from_key(hs.prefix);
  float v_hi = v_lo;
  if (gamma > 0.0 && hs.rank + 1 >= hs.count) {  // the next order statistic is a larger value
    if (use_tc) {
      const int rc = tc_gemm(EPI_MINGT, hhi, hlo, hhi, hlo, n, n, d, d, d, d, 1, g, st);
      if (rc) return rc;
    } else {
      gemm_f32_kernel<true, true, EPI_MINGT><<<grid, GTHREADS, 0, st>>>(g);
      GDR_LAUNCHED();
    }
    GDR_CUDA(cudaMemcpyAsync(&hs, sel, sizeof(SelectState), cudaMemcpyDeviceToHost, st));
    GDR_CUDA(cudaStreamSynchronize(st));
    if (hs.min_gt != 0xffffffffu) v_hi = f32_from_key(hs.min_gt);
  }
  // numpy _lerp: a + (b-a)*t, and b - (b-a)*(1-t) where t >= 0.5; (b-a) in fp32, the rest in fp64,
  // result cast back to the array dtype (fp32).
  const float diff = v_hi - v_lo;
  double r = (double)v_lo + (double)diff * gamma;
  if (gamma >= 0.5) r = (double)v_hi - (double)diff * (1.0 - gamma);
  *beta_out_host = (float)r;
  return GDR_OK;
}

int gdr_gram_percentile_f32(const float *h, int64_t n, int64_t d, double q_percent, float *beta_out_host,
                            void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  return gram_percentile_impl(h, n, d, q_percent, beta_out_host, workspace, workspace_bytes, stream, false);
}

// the same order statistic with the recomputed Gram tiles on the tensor cores (3xTF32, gemm_tc.cu)
int gdr_gram_percentile_tc_f32(const float *h, int64_t n, int64_t d, double q_percent, float *beta_out_host,
                               void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  return gram_percentile_impl(h, n, d, q_percent, beta_out_host, workspace, workspace_bytes, stream, true);
}

}  // extern "C"
