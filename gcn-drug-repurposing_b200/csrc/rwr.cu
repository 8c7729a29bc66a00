// rwr.cu — MSI diffusion profiles: batched personalised-PageRank power iteration, one column per
// seed (drug / indication), fp64.  Replaces, for ALL seeds at once,
//   convert_M_to_make_all_drugs_indications_sinks_except_selected  diffusion_profiles.py:29-47
//   refine_M_S                                                     diffusion_profiles.py:49-56
//   power_iteration                                                diffusion_profiles.py:65-90
// which the reference runs seed by seed in forked processes (:136-155).
//
// Layout in HBM: the iterate X, its successor and the per-seed row scale S are [n, SP] matrices
// (node-major, seeds contiguous, SP = n_seeds rounded up to 32): a warp = 32 seeds walking the
// same sparse row, so every gather X[u, s..s+31] is one coalesced 256-byte line and the matrix
// entries (col, val, flag) are warp-uniform broadcasts.  The per-seed matrix is never built:
// entry e of M is dropped for seed s when
//     flag bit0 (out-edge of a drug/indication row to one of its proteins) and row(e) != seed_s
//     flag bit1 (edge protein -> drug/indication it belongs to)             and col(e) == seed_s
// and kept entries are scaled by S_s[row] = 1/rowsum_s(row) (0 for an all-zero row).
//
// Bit-exactness with the reference (fp64): row sums run over the stored order of M's rows
// (scipy `M.sum(axis=1)` -> np.add.reduceat: row[0] + numpy pairwise sum of the rest; dropped
// entries are explicit zeros there and contribute +0.0); x*M is scipy's csc_matvec on M.T, i.e.
// y[v] accumulates (S[u]*M[u,v])*x[u] over sources u in ascending order with separate multiply and
// add (no FMA) — the transposed CSR (stable, sources ascending) walked sequentially by one thread
// reproduces it; the dangling mass is Python's builtin sum over ascending dangling nodes.  Only the
// L1 error (np.absolute(x-xlast).sum(), a pairwise sum) is reduced in a different (tree) order; it
// feeds nothing but the `err < N*tol` stop test.
#include <cub/cub.cuh>

#include "common.cuh"

namespace gdr {

constexpr int RW_SEEDS = 128;  // seeds per CTA (threadIdx.x)
constexpr int RW_ROWS = 4;     // rows per CTA step (threadIdx.y)
constexpr int RW_VTILE = 64;   // rows per CTA (each threadIdx.y walks RW_VTILE / RW_ROWS rows)

__device__ __forceinline__ bool rw_keep(unsigned flag, int32_t src, int32_t dst, int32_t seed) {
  return !(((flag & 1u) && src != seed) || ((flag & 2u) && dst == seed));
}

// S[u, s] = 1 / rowsum_s(u)  (0 if the row sums to 0).  Row sums in numpy reduceat order:
// first + pairwise(rest); dropped entries count as 0.0.  Rows longer than 129 fall back to the
// generic explicit-stack walk below.
__device__ double rw_pairwise(const double *__restrict__ val, const int32_t *__restrict__ col,
                              const uint8_t *__restrict__ flag, int32_t base, int n, int32_t u, int32_t seed) {
  // f(n) = n<8: sequential from 0 ; n<=128: 8 strided accumulators ; else split (n2 = n/2 - (n/2)%8)
  int off[28], len[28];
  unsigned char st[28];
  double stack[28];
  int sp = 0, vp = 0;
  off[0] = 0, len[0] = n, st[0] = 0;
  auto at = [&](int i) -> double {
    const int32_t e = base + i;
    return (flag == nullptr || rw_keep(flag[e], u, col[e], seed)) ? val[e] : 0.0;
  };
  while (sp >= 0) {
    const int o = off[sp], l = len[sp];
    if (l <= 128) {
      double res;
      if (l < 8) {
        res = 0.0;
        for (int i = 0; i < l; ++i) res = __dadd_rn(res, at(o + i));
      } else {
        double r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = at(o + j);
        int i = 8;
        for (; i < l - (l % 8); i += 8) {
#pragma unroll
          for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], at(o + i + j));
        }
        res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                        __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
        for (; i < l; ++i) res = __dadd_rn(res, at(o + i));
      }
      stack[vp++] = res;
      --sp;
      continue;
    }
    int n2 = l / 2;
    n2 -= n2 % 8;
    if (st[sp] == 0) {
      st[sp] = 1;
      ++sp;
      off[sp] = o, len[sp] = n2, st[sp] = 0;
    } else if (st[sp] == 1) {
      st[sp] = 2;
      ++sp;
      off[sp] = o + n2, len[sp] = l - n2, st[sp] = 0;
    } else {
      const double rgt = stack[--vp], lft = stack[--vp];
      stack[vp++] = __dadd_rn(lft, rgt);
      --sp;
    }
  }
  return stack[0];
}

__global__ void __launch_bounds__(RW_SEEDS)
rw_rowscale_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                   const double *__restrict__ val, const uint8_t *__restrict__ flag, int64_t n,
                   const int32_t *__restrict__ seed_node, int64_t n_seeds, int64_t sp,
                   double *__restrict__ S, uint8_t *__restrict__ dangling_any) {
  const int64_t s = (int64_t)blockIdx.y * RW_SEEDS + threadIdx.x;
  if (s >= sp) return;
  const int32_t seed = s < n_seeds ? seed_node[s] : -1;
  const int64_t u0 = (int64_t)blockIdx.x * RW_VTILE;
  const int64_t u1 = min(n, u0 + RW_VTILE);
  for (int64_t u = u0; u < u1; ++u) {
    const int32_t b = rowptr[u], e = rowptr[u + 1];
    double sum = 0.0;
    if (e > b) {
      const double first = rw_keep(flag[b], (int32_t)u, col[b], seed) ? val[b] : 0.0;
      sum = (e - b > 1) ? __dadd_rn(first, rw_pairwise(val, col, flag, b + 1, e - b - 1, (int32_t)u, seed)) : first;
    }
    const double sc = sum != 0.0 ? 1.0 / sum : 0.0;
    S[u * sp + s] = sc;
    if (sc == 0.0 && s < n_seeds) dangling_any[u] = 1;  // benign race: every writer stores 1
  }
}

__global__ void rw_init_kernel(double *__restrict__ x, int64_t count, double v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) x[i] = v;
}

// dangling mass per seed: Python's sum(x[is_dangling]) = sequential over ascending dangling nodes.
// cand[] lists (ascending) the nodes that are dangling for at least one seed.
__global__ void __launch_bounds__(128)
rw_dangling_kernel(const double *__restrict__ x, const double *__restrict__ S, const int32_t *__restrict__ cand,
                   const int32_t *__restrict__ n_cand, int64_t sp, const int32_t *__restrict__ done,
                   double *__restrict__ dsum) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= sp) return;
  if (done[s]) return;
  const int nc = *n_cand;
  double acc = 0.0;
  for (int i = 0; i < nc; ++i) {
    const int64_t u = cand[i];
    if (S[u * sp + s] == 0.0) acc = __dadd_rn(acc, x[u * sp + s]);
  }
  dsum[s] = acc;
}

// one power-iteration step for a [RW_VTILE rows] x [RW_SEEDS seeds] tile; Mt = transposed CSR
// (row v lists its sources u ascending).  Writes x_new and the per-tile L1 error partials.
__global__ void __launch_bounds__(RW_SEEDS *RW_ROWS)
rw_step_kernel(const int32_t *__restrict__ t_rowptr, const int32_t *__restrict__ t_col,
               const double *__restrict__ t_val, const uint8_t *__restrict__ t_flag, int64_t n,
               const int32_t *__restrict__ seed_node, int64_t n_seeds, int64_t sp, const double *__restrict__ S,
               const double *__restrict__ x, const double *__restrict__ dsum, const int32_t *__restrict__ done,
               double alpha, double one_minus_alpha, double *__restrict__ x_new, double *__restrict__ err_part) {
  __shared__ double red[RW_ROWS][RW_SEEDS];
  const int64_t s = (int64_t)blockIdx.y * RW_SEEDS + threadIdx.x;
  const bool live_col = s < sp;
  const int32_t seed = (live_col && s < n_seeds) ? seed_node[s] : -1;
  const bool is_done = live_col ? (done[s] != 0) : true;
  const double ds = (live_col && !is_done) ? dsum[s] : 0.0;
  const int64_t v0 = (int64_t)blockIdx.x * RW_VTILE;
  double err = 0.0;
  if (live_col) {
    for (int64_t v = v0 + threadIdx.y; v < min(n, v0 + (int64_t)RW_VTILE); v += RW_ROWS) {
      if (is_done) {  // frozen column: keep both ping-pong buffers identical
        x_new[v * sp + s] = x[v * sp + s];
        continue;
      }
      const int32_t b = t_rowptr[v], e = t_rowptr[v + 1];
      double acc = 0.0;
      for (int32_t j = b; j < e; ++j) {
        const int32_t u = t_col[j];
        if (rw_keep(t_flag[j], u, (int32_t)v, seed)) {
          const double m = __dmul_rn(S[(int64_t)u * sp + s], t_val[j]);   // refine_M_S: Q * M
          acc = __dadd_rn(acc, __dmul_rn(m, x[(int64_t)u * sp + s]));     // csc_matvec: y += A*x
        }
      }
      const double p = (v == seed) ? 1.0 : 0.0;
      const double t = __dadd_rn(acc, __dmul_rn(ds, p));
      const double xn = __dadd_rn(__dmul_rn(alpha, t), __dmul_rn(one_minus_alpha, p));
      x_new[v * sp + s] = xn;
      err += fabs(xn - x[v * sp + s]);
    }
  }
  red[threadIdx.y][threadIdx.x] = err;
  __syncthreads();
  if (threadIdx.y == 0 && live_col) {
    double tot = 0.0;
#pragma unroll
    for (int r = 0; r < RW_ROWS; ++r) tot += red[r][threadIdx.x];
    err_part[(int64_t)blockIdx.x * sp + s] = tot;
  }
}

// per seed: total L1 error -> converged?  done[s] = iteration number (1-based) at convergence.
__global__ void __launch_bounds__(128)
rw_check_kernel(const double *__restrict__ err_part, int64_t n_tiles, int64_t n_seeds, int64_t sp, double thresh,
                int32_t iter, int32_t *__restrict__ done, int32_t *__restrict__ n_done) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_seeds || done[s]) return;
  double tot = 0.0;
  for (int64_t t = 0; t < n_tiles; ++t) tot += err_part[t * sp + s];
  if (tot < thresh) {
    done[s] = iter;
    atomicAdd(n_done, 1);
  }
}

// x_out[s, v] = x[v, s]  (32x32 smem transpose)
__global__ void __launch_bounds__(256)
rw_transpose_out_kernel(const double *__restrict__ x, int64_t n, int64_t n_seeds, int64_t sp,
                        double *__restrict__ out) {
  __shared__ double tile[32][33];
  const int64_t v0 = (int64_t)blockIdx.x * 32, s0 = (int64_t)blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int64_t v = v0 + r, s = s0 + tx;
    tile[r][tx] = (v < n && s < sp) ? x[v * sp + s] : 0.0;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t s = s0 + r, v = v0 + tx;
    if (s < n_seeds && v < n) out[s * n + v] = tile[tx][r];
  }
}

__global__ void rw_iota_kernel(int32_t *p, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = (int32_t)i;
}

// =============================================================================================
// Per-seed API of the reference (DiffusionProfiles.refine_M_S / power_iteration called with a
// caller-supplied matrix): one vector, same accumulation orders as above.
// =============================================================================================

// refine_M_S (:49-56): S[u] = 1/rowsum(u) (numpy reduceat order; 0 stays 0) and val_out = S[row]*val
// (scipy's Q*M: one product per entry).
__global__ void __launch_bounds__(128)
rw1_refine_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                  const double *__restrict__ val, int64_t n, double *__restrict__ S, double *__restrict__ val_out) {
  const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n) return;
  const int32_t b = rowptr[u], e = rowptr[u + 1];
  double sum = 0.0;
  if (e > b) {
    const double first = val[b];
    sum = (e - b > 1) ? __dadd_rn(first, rw_pairwise(val, col, nullptr, b + 1, e - b - 1, (int32_t)u, -1)) : first;
  }
  const double sc = sum != 0.0 ? 1.0 / sum : 0.0;
  S[u] = sc;
  for (int32_t j = b; j < e; ++j) val_out[j] = __dmul_rn(sc, val[j]);
}

// p <- p / p.sum()  (numpy pairwise sum, one thread) ; x <- 1/N ; dangling list = ascending u with S[u] == 0
__global__ void rw1_prepare_kernel(double *__restrict__ p, const double *__restrict__ S, int64_t n,
                                   double *__restrict__ x, int32_t *__restrict__ dang, int32_t *__restrict__ n_dang) {
  __shared__ double total;
  if (threadIdx.x == 0) {
    total = rw_pairwise(p, nullptr, nullptr, 0, (int)n, 0, -1);
    int c = 0;
    for (int64_t u = 0; u < n; ++u)
      if (S[u] == 0.0) dang[c++] = (int32_t)u;
    *n_dang = c;
  }
  __syncthreads();
  const double t = total, x0 = 1.0 / (double)n;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
    p[i] = p[i] / t;
    x[i] = x0;
  }
}

// Python's sum(x[is_dangling]): sequential over ascending dangling nodes
__global__ void rw1_dangling_kernel(const double *__restrict__ x, const int32_t *__restrict__ dang,
                                    const int32_t *__restrict__ n_dang, double *__restrict__ dsum) {
  double acc = 0.0;
  const int c = *n_dang;
  for (int i = 0; i < c; ++i) acc = __dadd_rn(acc, x[dang[i]]);
  *dsum = acc;
}

// x_new = alpha*(x*M + dsum*p) + (1-alpha)*p with x*M as scipy's csc_matvec on M.T: sources ascending,
// separate multiply and add.  Mt = CSR of the transpose of the (already row-scaled) M.
__global__ void __launch_bounds__(128)
rw1_step_kernel(const int32_t *__restrict__ t_rowptr, const int32_t *__restrict__ t_col,
                const double *__restrict__ t_val, int64_t n, const double *__restrict__ x,
                const double *__restrict__ p, const double *__restrict__ dsum, double alpha, double oma,
                double *__restrict__ x_new, double *__restrict__ err_part) {
  __shared__ double red[4];
  const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double err = 0.0;
  if (v < n) {
    double acc = 0.0;
    for (int32_t j = t_rowptr[v]; j < t_rowptr[v + 1]; ++j) acc = __dadd_rn(acc, __dmul_rn(t_val[j], x[t_col[j]]));
    const double pv = p[v];
    const double t = __dadd_rn(acc, __dmul_rn(*dsum, pv));
    const double xn = __dadd_rn(__dmul_rn(alpha, t), __dmul_rn(oma, pv));
    x_new[v] = xn;
    err = fabs(xn - x[v]);
  }
  err = warp_sum(err);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = err;
  __syncthreads();
  if (threadIdx.x == 0) err_part[blockIdx.x] = (red[0] + red[1]) + (red[2] + red[3]);
}

__global__ void rw1_err_kernel(const double *__restrict__ err_part, int n_blocks, double *__restrict__ err_out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < n_blocks; i += blockDim.x) s += err_part[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    *err_out = t;
  }
}

struct RwPlan {
  int64_t sp, n_tiles;
  size_t select_tmp;
};
static RwPlan rw_plan(int64_t n, int64_t n_seeds) {
  RwPlan p;
  p.sp = align_up(n_seeds > 0 ? n_seeds : 1, 32);
  p.n_tiles = ceil_div<int64_t>(n > 0 ? n : 1, RW_VTILE);
  p.select_tmp = 0;
  cub::DeviceSelect::Flagged(nullptr, p.select_tmp, (const int32_t *)nullptr, (const uint8_t *)nullptr,
                             (int32_t *)nullptr, (int32_t *)nullptr, (int)n);
  return p;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int64_t gdr_rwr_workspace_bytes(int64_t n, int64_t nnz, int64_t n_seeds) {
  if (n < 0 || nnz < 0 || n_seeds < 0) return GDR_E_BADARG;
  const RwPlan p = rw_plan(n, n_seeds);
  const int64_t mat = align_up(n * p.sp * 8, 256);
  return 3 * mat + align_up(p.n_tiles * p.sp * 8, 256) + align_up(p.sp * 8, 256) + align_up(p.sp * 4, 256) +
         align_up(n + 1, 256) + 2 * align_up((n + 1) * 4, 256) + align_up((int64_t)p.select_tmp, 256) + 4096;
}

int gdr_rwr_solve_f64(const int32_t *m_rowptr, const int32_t *m_col, const double *m_val, const uint8_t *m_flag,
                      const int32_t *mt_rowptr, const int32_t *mt_col, const double *mt_val,
                      const uint8_t *mt_flag, int64_t n, int64_t nnz, const int32_t *seed_node, int64_t n_seeds,
                      double alpha, double tol, int64_t max_iter, double *x_out, int32_t *iters_out,
                      void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && nnz >= 0 && n_seeds >= 0 && max_iter >= 0, GDR_E_BADARG, "rwr_solve: bad sizes");
  GDR_CHECK_ARG(n < (int64_t)INT32_MAX, GDR_E_SHAPE, "rwr_solve: int32 node range exceeded");
  if (n_seeds == 0) return GDR_OK;
  GDR_CHECK_ARG(m_rowptr && mt_rowptr && seed_node && x_out && iters_out && workspace, GDR_E_BADARG,
                "rwr_solve: null pointer");
  GDR_CHECK_ARG(nnz == 0 || (m_col && m_val && m_flag && mt_col && mt_val && mt_flag), GDR_E_BADARG,
                "rwr_solve: null matrix arrays");
  GDR_CHECK_ARG(workspace_bytes >= gdr_rwr_workspace_bytes(n, nnz, n_seeds), GDR_E_WORKSPACE,
                "rwr_solve: workspace too small");
  cudaStream_t st = as_stream(stream);
  const RwPlan p = rw_plan(n, n_seeds);
  const int64_t sp = p.sp;
  Carver cv(workspace);
  double *S = cv.take<double>(n * sp);
  double *xa = cv.take<double>(n * sp);
  double *xb = cv.take<double>(n * sp);
  double *err_part = cv.take<double>(p.n_tiles * sp);
  double *dsum = cv.take<double>(sp);
  int32_t *done = cv.take<int32_t>(sp);
  uint8_t *dang = cv.take<uint8_t>(n + 1);
  int32_t *iota = cv.take<int32_t>(n + 1);
  int32_t *cand = cv.take<int32_t>(n + 1);
  int32_t *counters = cv.take<int32_t>(8);  // [0] = n_cand, [1] = n_done
  void *sel_tmp = cv.take<char>((int64_t)p.select_tmp);
  size_t sel_bytes = p.select_tmp;

  GDR_CUDA(cudaMemsetAsync(done, 0, sp * 4, st));
  GDR_CUDA(cudaMemsetAsync(dang, 0, n + 1, st));
  GDR_CUDA(cudaMemsetAsync(counters, 0, 32, st));
  GDR_CUDA(cudaMemsetAsync(dsum, 0, sp * 8, st));

  dim3 tiles((unsigned)p.n_tiles, (unsigned)ceil_div<int64_t>(sp, RW_SEEDS), 1);
  rw_rowscale_kernel<<<tiles, RW_SEEDS, 0, st>>>(m_rowptr, m_col, m_val, m_flag, n, seed_node, n_seeds, sp, S, dang);
  GDR_LAUNCHED();
  rw_iota_kernel<<<(unsigned)ceil_div<int64_t>(n, 256), 256, 0, st>>>(iota, n);
  GDR_LAUNCHED();
  GDR_CUDA(cub::DeviceSelect::Flagged(sel_tmp, sel_bytes, (const int32_t *)iota, (const uint8_t *)dang, cand,
                                      counters, (int)n, st));
  count_launch(2);
  rw_init_kernel<<<(unsigned)ceil_div<int64_t>(n * sp, 256), 256, 0, st>>>(xa, n * sp, 1.0 / (double)n);
  GDR_LAUNCHED();

  const double thresh = (double)n * tol;
  const double oma = 1.0 - alpha;
  double *cur = xa, *nxt = xb;
  int32_t n_done = 0;
  int64_t it = 0;
  while (n_done < n_seeds && it < max_iter) {
    ++it;
    rw_dangling_kernel<<<(unsigned)ceil_div<int64_t>(sp, 128), 128, 0, st>>>(cur, S, cand, counters, sp, done, dsum);
    GDR_LAUNCHED();
    rw_step_kernel<<<tiles, dim3(RW_SEEDS, RW_ROWS, 1), 0, st>>>(mt_rowptr, mt_col, mt_val, mt_flag, n, seed_node,
                                                                n_seeds, sp, S, cur, dsum, done, alpha, oma, nxt,
                                                                err_part);
    GDR_LAUNCHED();
    rw_check_kernel<<<(unsigned)ceil_div<int64_t>(n_seeds, 128), 128, 0, st>>>(err_part, p.n_tiles, n_seeds, sp,
                                                                              thresh, (int32_t)it, done,
                                                                              counters + 1);
    GDR_LAUNCHED();
    GDR_CUDA(cudaMemcpyAsync(&n_done, counters + 1, 4, cudaMemcpyDeviceToHost, st));
    GDR_CUDA(cudaStreamSynchronize(st));
    double *t = cur;
    cur = nxt;
    nxt = t;
  }
  dim3 tg((unsigned)ceil_div<int64_t>(n, 32), (unsigned)ceil_div<int64_t>(sp, 32), 1);
  rw_transpose_out_kernel<<<tg, 256, 0, st>>>(cur, n, n_seeds, sp, x_out);
  GDR_LAUNCHED();
  GDR_CUDA(cudaMemcpyAsync(iters_out, done, n_seeds * 4, cudaMemcpyDeviceToDevice, st));
  GDR_CUDA(cudaStreamSynchronize(st));
  if (n_done < n_seeds) {
    set_error("rwr_solve: power iteration failed to converge in %lld iterations for %lld of %lld seeds",
              (long long)max_iter, (long long)(n_seeds - n_done), (long long)n_seeds);
    return GDR_E_NOCONV;
  }
  return GDR_OK;
}

// refine_M_S (diffusion_profiles.py:49-56) for a caller-supplied matrix: s_out[n] = 1/rowsum (0 stays 0),
// val_out[nnz] = s_out[row] * val.  The host mirror rebuilds scipy's Q*M structure from it.
int gdr_rwr_refine_f64(const int32_t *rowptr, const int32_t *col, const double *val, int64_t n, int64_t nnz,
                       double *s_out, double *val_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && nnz >= 0 && rowptr && s_out, GDR_E_BADARG, "rwr_refine: bad args");
  GDR_CHECK_ARG(nnz == 0 || (col && val && val_out), GDR_E_BADARG, "rwr_refine: null matrix arrays");
  rw1_refine_kernel<<<(unsigned)ceil_div<int64_t>(n, 128), 128, 0, as_stream(stream)>>>(rowptr, col, val, n, s_out,
                                                                                      val_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int64_t gdr_rwr_power_iteration_workspace_bytes(int64_t n) {
  if (n < 0) return GDR_E_BADARG;
  const int64_t blocks = ceil_div<int64_t>(n > 0 ? n : 1, 128);
  return 3 * align_up(n * 8, 256) + align_up(blocks * 8, 256) + align_up((n + 1) * 4, 256) + 1024;
}

// power_iteration(M, S, nodelist, per_dict) (diffusion_profiles.py:65-90) for ONE personalisation vector and a
// caller-supplied, already row-scaled M (given as the CSR of its transpose, sources ascending).  s[n]: the S of
// refine_M_S (only S == 0, the dangling rows, matters); p[n]: per_dict values in nodelist order (normalised
// here, p / p.sum()).  x_out[n]; *iters_out_host = iterations used.  Synchronises the stream every iteration
// (the L1 stop test).  GDR_E_NOCONV after max_iter iterations (the reference raises).
int gdr_rwr_power_iteration_f64(const int32_t *mt_rowptr, const int32_t *mt_col, const double *mt_val, int64_t n,
                                const double *s, const double *p, double alpha, double tol, int64_t max_iter,
                                double *x_out, int32_t *iters_out_host, void *workspace, int64_t workspace_bytes,
                                gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && max_iter >= 0 && mt_rowptr && s && p && x_out && workspace, GDR_E_BADARG,
                "rwr_power_iteration: bad args");
  GDR_CHECK_ARG(n < (int64_t)INT32_MAX, GDR_E_SHAPE, "rwr_power_iteration: int32 node range exceeded");
  GDR_CHECK_ARG(workspace_bytes >= gdr_rwr_power_iteration_workspace_bytes(n), GDR_E_WORKSPACE,
                "rwr_power_iteration: workspace too small");
  cudaStream_t st = as_stream(stream);
  const int blocks = (int)ceil_div<int64_t>(n, 128);
  Carver cv(workspace);
  double *pn = cv.take<double>(n), *xa = cv.take<double>(n), *xb = cv.take<double>(n);
  double *err_part = cv.take<double>(blocks);
  int32_t *dang = cv.take<int32_t>(n + 1);
  double *scal = cv.take<double>(2);        // [0] dangling mass, [1] L1 error
  int32_t *n_dang = cv.take<int32_t>(1);
  GDR_CUDA(cudaMemcpyAsync(pn, p, n * 8, cudaMemcpyDeviceToDevice, st));
  rw1_prepare_kernel<<<1, 1024, 0, st>>>(pn, s, n, xa, dang, n_dang);
  GDR_LAUNCHED();
  const double thresh = (double)n * tol, oma = 1.0 - alpha;
  double *cur = xa, *nxt = xb;
  int64_t it = 0;
  bool converged = false;
  while (!converged && it < max_iter) {
    ++it;
    rw1_dangling_kernel<<<1, 1, 0, st>>>(cur, dang, n_dang, scal);
    GDR_LAUNCHED();
    rw1_step_kernel<<<blocks, 128, 0, st>>>(mt_rowptr, mt_col, mt_val, n, cur, pn, scal, alpha, oma, nxt, err_part);
    GDR_LAUNCHED();
    rw1_err_kernel<<<1, 1024, 0, st>>>(err_part, blocks, scal + 1);
    GDR_LAUNCHED();
    double err = 0.0;
    GDR_CUDA(cudaMemcpyAsync(&err, scal + 1, 8, cudaMemcpyDeviceToHost, st));
    GDR_CUDA(cudaStreamSynchronize(st));
    converged = err < thresh;
    double *t = cur;
    cur = nxt;
    nxt = t;
  }
  GDR_CUDA(cudaMemcpyAsync(x_out, cur, n * 8, cudaMemcpyDeviceToDevice, st));
  GDR_CUDA(cudaStreamSynchronize(st));
  if (iters_out_host) *iters_out_host = (int32_t)it;
  if (!converged) {
    set_error("pagerank_scipy: power iteration failed to converge in %lld iterations.", (long long)max_iter);
    return GDR_E_NOCONV;
  }
  return GDR_OK;
}

}  // extern "C"
