// csr_build.cu — CSR construction on the device: COO->CSR, transpose, (A+I) symmetric
// normalisation.  One-off structure work (not the timed hot loop): the stable sorts and
// scans use CUB (header templates from the CUDA toolkit); everything else is hand-written.
//
// Reference behaviour restated here:
//   helpers/helper.py:92-96  convert_sparse_matrix_to_sparse_tensor -> COO int64, row-major
//   helpers/helper.py:82-89  preprocess_graph: D^-1/2 (A+I) D^-1/2, value (a*d_j)*d_i, fp64
//   scipy A.T.tocsr()        transposed rows ordered by source row (stable counting sort)
#include <cub/cub.cuh>

#include "common.cuh"

namespace gdr {

// ---- small kernels --------------------------------------------------------------------
__global__ void make_keys_kernel(const int64_t *__restrict__ row, const int64_t *__restrict__ col,
                                 int64_t nnz, int64_t n_cols, uint64_t *__restrict__ keys,
                                 int64_t *__restrict__ iota) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nnz) {
    keys[i] = (uint64_t)row[i] * (uint64_t)n_cols + (uint64_t)col[i];
    iota[i] = i;
  }
}

// sorted keys -> col ids and rowptr.  Thread i owns the gap (row(i-1), row(i)].
__global__ void keys_to_csr_kernel(const uint64_t *__restrict__ keys, int64_t nnz, int64_t n_rows,
                                   int64_t n_cols, int32_t *__restrict__ rowptr,
                                   int32_t *__restrict__ col_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > nnz) return;
  if (i == nnz) {  // tail: rows after the last stored entry (also the nnz == 0 case)
    int64_t r_last = nnz ? (int64_t)(keys[nnz - 1] / (uint64_t)n_cols) : -1;
    for (int64_t r = r_last + 1; r <= n_rows; ++r) rowptr[r] = (int32_t)nnz;
    return;
  }
  uint64_t k = keys[i];
  int64_t r = (int64_t)(k / (uint64_t)n_cols);
  col_out[i] = (int32_t)(k - (uint64_t)r * (uint64_t)n_cols);
  int64_t r_prev = i ? (int64_t)(keys[i - 1] / (uint64_t)n_cols) : -1;
  for (int64_t q = r_prev + 1; q <= r; ++q) rowptr[q] = (int32_t)i;
}

template <typename T>
__global__ void permute_kernel(const T *__restrict__ in, const int64_t *__restrict__ perm, int64_t n,
                               T *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[perm[i]];
}

// per-entry source row (warp per row), col as 32-bit sort key, iota
__global__ void expand_rows_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                                   int64_t n_rows, int32_t *__restrict__ row_of, uint32_t *__restrict__ keys,
                                   int64_t *__restrict__ iota) {
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (warp >= n_rows) return;
  int32_t s = rowptr[warp], e = rowptr[warp + 1];
  for (int32_t j = s + lane; j < e; j += 32) {
    row_of[j] = (int32_t)warp;
    keys[j] = (uint32_t)col[j];
    iota[j] = j;
  }
}

__global__ void tkeys_to_csr_kernel(const uint32_t *__restrict__ keys, const int64_t *__restrict__ perm,
                                    const int32_t *__restrict__ row_of, int64_t nnz, int64_t n_trows,
                                    int32_t *__restrict__ t_rowptr, int32_t *__restrict__ t_col) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > nnz) return;
  if (i == nnz) {
    int64_t r_last = nnz ? (int64_t)keys[nnz - 1] : -1;
    for (int64_t r = r_last + 1; r <= n_trows; ++r) t_rowptr[r] = (int32_t)nnz;
    return;
  }
  int64_t r = keys[i];
  t_col[i] = row_of[perm[i]];
  int64_t r_prev = i ? (int64_t)keys[i - 1] : -1;
  for (int64_t q = r_prev + 1; q <= r; ++q) t_rowptr[q] = (int32_t)i;
}

// rows of A+I: length + (no stored diagonal)
__global__ void add_identity_count_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                                          int64_t n, int32_t *__restrict__ counts) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n) return;
  if (r == n) {
    counts[n] = 0;
    return;
  }
  int32_t s = rowptr[r], e = rowptr[r + 1];
  int has = 0;
  for (int32_t j = s; j < e; ++j) has |= (col[j] == (int32_t)r);
  counts[r] = (e - s) + (has ? 0 : 1);
}

// x^-1/2, correctly rounded in all but vanishingly rare cases (one FMA-residual Newton
// step on 1/sqrt(x)), so that it agrees with glibc's pow(x,-0.5) that numpy calls.
__device__ __forceinline__ double inv_sqrt_cr(double x) {
  double y0 = 1.0 / sqrt(x);
  if (!(x > 0.0) || isinf(x)) return pow(x, -0.5);  // inf / nan exactly as the reference gives
  double t = x * y0;
  double t_err = fma(x, y0, -t);
  double e = fma(-t, y0, 1.0) - t_err * y0;  // 1 - x*y0^2
  return fma(0.5 * e, y0, y0);
}

// numpy's pairwise summation (numpy/_core/src/umath/loops_utils.h.src, PW_BLOCKSIZE 128), restated:
// scipy's csr.sum(axis=1) is np.add.reduceat(data, indptr[:-1]) (scipy/sparse/_compressed.py, sum ->
// _minor_reduce), which evaluates every row as data[0] + pairwise_sum(data[1:]).  Reproducing that
// order makes the row sums (and through them D^-1/2) bit-identical to the reference's.
__device__ double np_pairwise_block(const double *a, int n) {  // n <= 128
  if (n < 8) {
    double r = 0.0;
    for (int i = 0; i < n; ++i) r = __dadd_rn(r, a[i]);
    return r;
  }
  double r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r[j] = a[j];
  int i = 8;
  for (; i < n - (n % 8); i += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
  }
  double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                         __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
  for (; i < n; ++i) res = __dadd_rn(res, a[i]);
  return res;
}

// explicit-stack version of the recursion  f(a,n) = n<=128 ? block : f(a,n2) + f(a+n2,n-n2),
// n2 = n/2 rounded down to a multiple of 8.  Post-order evaluation with a value stack.
__device__ double np_pairwise_sum(const double *a, int n) {
  if (n <= 128) return np_pairwise_block(a, n);
  // frames: (offset, length, state) ; depth <= log2(2^31/128)+1 < 28
  int off[28], len[28];
  unsigned char st[28];
  double val[28];
  int sp = 0, vp = 0;
  off[0] = 0, len[0] = n, st[0] = 0;
  while (sp >= 0) {
    const int o = off[sp], l = len[sp];
    if (l <= 128) {
      val[vp++] = np_pairwise_block(a + o, l);
      --sp;
      continue;
    }
    int n2 = l / 2;
    n2 -= n2 % 8;
    if (st[sp] == 0) {          // descend left
      st[sp] = 1;
      ++sp;
      off[sp] = o, len[sp] = n2, st[sp] = 0;
    } else if (st[sp] == 1) {   // descend right
      st[sp] = 2;
      ++sp;
      off[sp] = o + n2, len[sp] = l - n2, st[sp] = 0;
    } else {                    // combine
      const double rgt = val[--vp], lft = val[--vp];
      val[vp++] = __dadd_rn(lft, rgt);
      --sp;
    }
  }
  return val[0];
}

// merge the unit diagonal into every (sorted) row (scipy's A + eye), then the row sum in the
// reference's order: row[0] + pairwise_sum(row[1:]); d = rowsum^-1/2.
__global__ void add_identity_rowsum_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                                           const double *__restrict__ val, int64_t n,
                                           const int32_t *__restrict__ out_rowptr, int32_t *__restrict__ out_col,
                                           double *__restrict__ out_val, double *__restrict__ dis) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  int32_t s = rowptr[r], e = rowptr[r + 1];
  const int32_t o0 = out_rowptr[r];
  int32_t o = o0;
  bool placed = false;
  for (int32_t j = s; j < e; ++j) {
    int32_t c = col[j];
    double v = val[j];
    if (!placed && c >= (int32_t)r) {
      placed = true;
      if (c == (int32_t)r) {
        v += 1.0;  // stored diagonal: scipy's A + eye adds in place
      } else {
        out_col[o] = (int32_t)r;
        out_val[o] = 1.0;
        ++o;
      }
    }
    out_col[o] = c;
    out_val[o] = v;
    ++o;
  }
  if (!placed) {
    out_col[o] = (int32_t)r;
    out_val[o] = 1.0;
    ++o;
  }
  const int cnt = o - o0;  // >= 1
  const double sum = cnt > 1 ? __dadd_rn(out_val[o0], np_pairwise_sum(out_val + o0 + 1, cnt - 1)) : out_val[o0];
  dis[r] = inv_sqrt_cr(sum);
}

__global__ void sym_scale_kernel(const int32_t *__restrict__ out_rowptr, const int32_t *__restrict__ out_col,
                                 double *__restrict__ out_val, const double *__restrict__ dis, int64_t n) {
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (warp >= n) return;
  int32_t s = out_rowptr[warp], e = out_rowptr[warp + 1];
  double di = dis[warp];
  for (int32_t j = s + lane; j < e; j += 32) out_val[j] = (out_val[j] * dis[out_col[j]]) * di;
}

static int end_bit_for(uint64_t max_key) {
  int b = 1;
  while (b < 64 && (max_key >> b)) ++b;
  return b;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int64_t gdr_csr_from_coo_workspace_bytes(int64_t nnz, int64_t n_rows, int64_t n_cols) {
  if (nnz < 0 || n_rows < 0 || n_cols < 0) return GDR_E_BADARG;
  size_t tmp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                  (const int64_t *)nullptr, (int64_t *)nullptr, nnz, 0, 64);
  int64_t n1 = nnz + 1;
  return align_up(n1 * 8, 256) * 4 + align_up((int64_t)tmp, 256) + 1024;
}

int gdr_csr_from_coo(const int64_t *coo_row, const int64_t *coo_col, int64_t nnz, int64_t n_rows,
                     int64_t n_cols, int32_t *rowptr_out, int32_t *col_out, int64_t *perm_out,
                     void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(nnz >= 0 && n_rows >= 0 && n_cols > 0, GDR_E_BADARG, "csr_from_coo: bad sizes");
  GDR_CHECK_ARG(nnz < (int64_t)INT32_MAX && n_rows < (int64_t)INT32_MAX && n_cols < (int64_t)INT32_MAX,
                GDR_E_SHAPE, "csr_from_coo: int32 index range exceeded");
  GDR_CHECK_ARG(rowptr_out && workspace && (nnz == 0 || (coo_row && coo_col && col_out)), GDR_E_BADARG,
                "csr_from_coo: null pointer");
  GDR_CHECK_ARG(workspace_bytes >= gdr_csr_from_coo_workspace_bytes(nnz, n_rows, n_cols), GDR_E_WORKSPACE,
                "csr_from_coo: workspace too small");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  uint64_t *keys_in = cv.take<uint64_t>(nnz + 1);
  uint64_t *keys_out = cv.take<uint64_t>(nnz + 1);
  int64_t *iota = cv.take<int64_t>(nnz + 1);
  int64_t *perm = perm_out ? perm_out : cv.take<int64_t>(nnz + 1);
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                  (const int64_t *)nullptr, (int64_t *)nullptr, nnz, 0, 64);
  void *tmp = cv.take<char>((int64_t)tmp_bytes);
  const int T = 256;
  if (nnz > 0) {
    make_keys_kernel<<<(unsigned)ceil_div<int64_t>(nnz, T), T, 0, st>>>(coo_row, coo_col, nnz, n_cols, keys_in,
                                                                       iota);
    GDR_LAUNCHED();
    int eb = end_bit_for((uint64_t)n_rows * (uint64_t)n_cols);
    GDR_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, (const uint64_t *)keys_in, keys_out,
                                             (const int64_t *)iota, perm, nnz, 0, eb, st));
    count_launch(4);
  }
  keys_to_csr_kernel<<<(unsigned)ceil_div<int64_t>(nnz + 1, T), T, 0, st>>>(keys_out, nnz, n_rows, n_cols,
                                                                           rowptr_out, col_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_permute_values(const void *values_in, const int64_t *perm, int64_t nnz, int elem_bytes,
                       void *values_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(nnz >= 0 && (elem_bytes == 1 || elem_bytes == 4 || elem_bytes == 8), GDR_E_BADARG,
                "permute_values: bad args");
  if (nnz == 0) return GDR_OK;
  GDR_CHECK_ARG(values_in && perm && values_out, GDR_E_BADARG, "permute_values: null pointer");
  const int T = 256;
  unsigned g = (unsigned)ceil_div<int64_t>(nnz, T);
  if (elem_bytes == 1)
    permute_kernel<uint8_t><<<g, T, 0, as_stream(stream)>>>((const uint8_t *)values_in, perm, nnz,
                                                           (uint8_t *)values_out);
  else if (elem_bytes == 4)
    permute_kernel<uint32_t><<<g, T, 0, as_stream(stream)>>>((const uint32_t *)values_in, perm, nnz,
                                                            (uint32_t *)values_out);
  else
    permute_kernel<uint64_t><<<g, T, 0, as_stream(stream)>>>((const uint64_t *)values_in, perm, nnz,
                                                            (uint64_t *)values_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int64_t gdr_csr_transpose_workspace_bytes(int64_t nnz, int64_t n_rows, int64_t n_cols) {
  if (nnz < 0 || n_rows < 0 || n_cols < 0) return GDR_E_BADARG;
  size_t tmp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                  (const int64_t *)nullptr, (int64_t *)nullptr, nnz, 0, 32);
  int64_t n1 = nnz + 1;
  return align_up(n1 * 4, 256) * 3 + align_up(n1 * 8, 256) * 2 + align_up((int64_t)tmp, 256) + 1024;
}

int gdr_csr_transpose(const int32_t *rowptr, const int32_t *col, int64_t nnz, int64_t n_rows, int64_t n_cols,
                      int32_t *t_rowptr_out, int32_t *t_col_out, int64_t *perm_out, void *workspace,
                      int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(nnz >= 0 && n_rows >= 0 && n_cols > 0, GDR_E_BADARG, "csr_transpose: bad sizes");
  GDR_CHECK_ARG(rowptr && t_rowptr_out && workspace && (nnz == 0 || (col && t_col_out)), GDR_E_BADARG,
                "csr_transpose: null pointer");
  GDR_CHECK_ARG(workspace_bytes >= gdr_csr_transpose_workspace_bytes(nnz, n_rows, n_cols), GDR_E_WORKSPACE,
                "csr_transpose: workspace too small");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  int32_t *row_of = cv.take<int32_t>(nnz + 1);
  uint32_t *keys_in = cv.take<uint32_t>(nnz + 1);
  uint32_t *keys_out = cv.take<uint32_t>(nnz + 1);
  int64_t *iota = cv.take<int64_t>(nnz + 1);
  int64_t *perm = perm_out ? perm_out : cv.take<int64_t>(nnz + 1);
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                  (const int64_t *)nullptr, (int64_t *)nullptr, nnz, 0, 32);
  void *tmp = cv.take<char>((int64_t)tmp_bytes);
  const int T = 256;
  if (nnz > 0) {
    expand_rows_kernel<<<(unsigned)ceil_div<int64_t>(n_rows * 32, T), T, 0, st>>>(rowptr, col, n_rows, row_of,
                                                                                 keys_in, iota);
    GDR_LAUNCHED();
    int eb = end_bit_for((uint64_t)n_cols);
    GDR_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, (const uint32_t *)keys_in, keys_out,
                                             (const int64_t *)iota, perm, nnz, 0, eb, st));
    count_launch(4);
  }
  tkeys_to_csr_kernel<<<(unsigned)ceil_div<int64_t>(nnz + 1, T), T, 0, st>>>(keys_out, perm, row_of, nnz, n_cols,
                                                                            t_rowptr_out, t_col_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int64_t gdr_csr_add_identity_workspace_bytes(int64_t n) {
  if (n < 0) return GDR_E_BADARG;
  size_t tmp = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp, (const int32_t *)nullptr, (int32_t *)nullptr, n + 1);
  return align_up((n + 1) * 4, 256) + align_up((int64_t)tmp, 256) + 1024;
}

int gdr_csr_add_identity_count(const int32_t *rowptr, const int32_t *col, int64_t n, int32_t *out_rowptr,
                               int64_t *nnz_out_host, void *workspace, int64_t workspace_bytes,
                               gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && rowptr && out_rowptr && nnz_out_host && workspace, GDR_E_BADARG,
                "csr_add_identity_count: bad args");
  GDR_CHECK_ARG(workspace_bytes >= gdr_csr_add_identity_workspace_bytes(n), GDR_E_WORKSPACE,
                "csr_add_identity_count: workspace too small");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  int32_t *counts = cv.take<int32_t>(n + 1);
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (const int32_t *)nullptr, (int32_t *)nullptr, n + 1);
  void *tmp = cv.take<char>((int64_t)tmp_bytes);
  const int T = 256;
  add_identity_count_kernel<<<(unsigned)ceil_div<int64_t>(n + 1, T), T, 0, st>>>(rowptr, col, n, counts);
  GDR_LAUNCHED();
  GDR_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, (const int32_t *)counts, out_rowptr, n + 1, st));
  count_launch(2);
  int32_t last = 0;
  GDR_CUDA(cudaMemcpyAsync(&last, out_rowptr + n, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  GDR_CUDA(cudaStreamSynchronize(st));
  *nnz_out_host = last;
  return GDR_OK;
}

int gdr_csr_normalize_f64(const int32_t *rowptr, const int32_t *col, const double *val, int64_t n,
                          const int32_t *out_rowptr, int32_t *out_col, double *out_val,
                          double *deg_inv_sqrt_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && rowptr && out_rowptr && out_col && out_val && deg_inv_sqrt_out, GDR_E_BADARG,
                "csr_normalize: null pointer");
  if (n == 0) return GDR_OK;
  cudaStream_t st = as_stream(stream);
  const int T = 128;
  add_identity_rowsum_kernel<<<(unsigned)ceil_div<int64_t>(n, T), T, 0, st>>>(rowptr, col, val, n, out_rowptr,
                                                                             out_col, out_val, deg_inv_sqrt_out);
  GDR_LAUNCHED();
  sym_scale_kernel<<<(unsigned)ceil_div<int64_t>(n * 32, 256), 256, 0, st>>>(out_rowptr, out_col, out_val,
                                                                            deg_inv_sqrt_out, n);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
