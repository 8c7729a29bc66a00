// rank.cu — the consumer step right after the path: rank nodes by the proximity of their
// (L2-normalised) embedding to one query embedding, and rank drugs by a diffusion profile.
//
//   predict_drug.py:49-74   node_embs = normalize(np.loadtxt(gcn_emb_file), axis=1)      (sklearn, fp64)
//                           proximities = np.matmul(drug_embs, covid_emb); np.argsort(proximities)[::-1]
//   run_covid.py:294-306    protein_proximities = np.matmul(protein_embs, covid_emb)
//   predict_drug.py:111-124 np.argsort(profile[drug nodes])[::-1]
//
// fp64 like the reference (np.loadtxt yields fp64).  One warp scores one selected row: the row and the query are
// normalised element by element (x / max-free sklearn rule: a zero norm divides by 1) and multiplied, the d
// products are summed by a lane-strided loop + warp tree.  The order is a descending radix sort of the scores
// carrying the row positions (stable: equal scores keep ascending position; numpy's quicksort leaves ties in an
// unspecified order, so ties are excepted from parity).
#include <cub/cub.cuh>

#include "common.cuh"

namespace gdr {

__device__ __forceinline__ double row_norm_f64(const double *__restrict__ r, int64_t d, int lane) {
  double ss = 0.0;
  for (int64_t c = lane; c < d; c += 32) ss = fma(r[c], r[c], ss);
  ss = warp_sum(ss);
  const double nr = sqrt(ss);
  return nr == 0.0 ? 1.0 : nr;  // sklearn.preprocessing.normalize: norms[norms == 0] = 1
}

__global__ void __launch_bounds__(256)
rank_score_kernel(const double *__restrict__ embs, int64_t n, int64_t d, int normalize, int64_t query_row,
                  const int32_t *__restrict__ sel, int64_t m, double *__restrict__ prox) {
  const int lane = threadIdx.x & 31;
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= m) return;
  const int64_t r = sel ? (int64_t)sel[i] : i;
  const double *x = embs + r * d, *q = embs + query_row * d;
  double nx = 1.0, nq = 1.0;
  if (normalize) {
    nx = row_norm_f64(x, d, lane);
    nq = row_norm_f64(q, d, lane);
  }
  double acc = 0.0;
  for (int64_t c = lane; c < d; c += 32) acc = fma(x[c] / nx, q[c] / nq, acc);
  acc = warp_sum(acc);
  if (lane == 0) prox[i] = acc;
}

__global__ void rank_iota_kernel(int32_t *p, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = (int32_t)i;
}

__global__ void rank_gather_kernel(const double *__restrict__ v, const int32_t *__restrict__ sel, int64_t m,
                                   double *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) out[i] = v[sel ? sel[i] : i];
}

static size_t rank_sort_tmp_bytes(int64_t m) {
  size_t b = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, b, (const double *)nullptr, (double *)nullptr,
                                            (const int32_t *)nullptr, (int32_t *)nullptr, (int)m);
  return b;
}

static int rank_sort(const double *keys, int64_t m, double *sorted_out, int32_t *order_out, Carver &cv,
                     cudaStream_t st) {
  int32_t *iota = cv.take<int32_t>(m);
  size_t tmp_bytes = rank_sort_tmp_bytes(m);
  void *tmp = cv.take<char>((int64_t)tmp_bytes);
  rank_iota_kernel<<<(unsigned)ceil_div<int64_t>(m, 256), 256, 0, st>>>(iota, m);
  GDR_LAUNCHED();
  GDR_CUDA(cub::DeviceRadixSort::SortPairsDescending(tmp, tmp_bytes, keys, sorted_out, iota, order_out, (int)m, 0, 64,
                                                     st));
  count_launch(4);
  return GDR_OK;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int64_t gdr_rank_workspace_bytes(int64_t m) {
  if (m < 0) return GDR_E_BADARG;
  const int64_t mm = m > 0 ? m : 1;
  return 2 * align_up(mm * 8, 256) + align_up(mm * 4, 256) + align_up((int64_t)rank_sort_tmp_bytes(mm), 256) + 1024;
}

// prox_out[i] = <normalize(embs[sel[i]]), normalize(embs[query_row])> (normalize != 0) or the raw dot product;
// order_out[j] = position i (into sel) of the j-th largest score; sorted_out[j] = that score (optional, may be NULL).
// sel == NULL: rows 0..m-1.  embs: fp64 [n, d] row-major on the device.
int gdr_rank_embedding_f64(const double *embs, int64_t n, int64_t d, int normalize, int64_t query_row,
                           const int32_t *sel, int64_t m, double *prox_out, int32_t *order_out, double *sorted_out,
                           void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && d > 0 && m >= 0, GDR_E_BADARG, "rank_embedding: bad sizes");
  if (m == 0) return GDR_OK;
  GDR_CHECK_ARG(embs && prox_out && order_out && workspace, GDR_E_BADARG, "rank_embedding: null pointer");
  GDR_CHECK_ARG(query_row >= 0 && query_row < n, GDR_E_BADARG, "rank_embedding: query row %lld outside [0, %lld)",
                (long long)query_row, (long long)n);
  GDR_CHECK_ARG(m < (int64_t)INT32_MAX, GDR_E_SHAPE, "rank_embedding: int32 index range exceeded");
  GDR_CHECK_ARG(workspace_bytes >= gdr_rank_workspace_bytes(m), GDR_E_WORKSPACE, "rank_embedding: workspace too small");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  double *sorted = cv.take<double>(m);
  rank_score_kernel<<<(unsigned)ceil_div<int64_t>(m * 32, 256), 256, 0, st>>>(embs, n, d, normalize, query_row, sel, m,
                                                                             prox_out);
  GDR_LAUNCHED();
  return rank_sort(prox_out, m, sorted_out ? sorted_out : sorted, order_out, cv, st);
}

// values_out[i] = values[sel[i]] and its descending order (np.argsort(values[sel])[::-1]): the diffusion-profile
// consumer (predict_drug.py:111-124: rank the drug entries of the NodeCovid profile).
int gdr_rank_values_f64(const double *values, const int32_t *sel, int64_t m, double *values_out, int32_t *order_out,
                        double *sorted_out, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(m >= 0, GDR_E_BADARG, "rank_values: bad size");
  if (m == 0) return GDR_OK;
  GDR_CHECK_ARG(values && values_out && order_out && workspace, GDR_E_BADARG, "rank_values: null pointer");
  GDR_CHECK_ARG(m < (int64_t)INT32_MAX, GDR_E_SHAPE, "rank_values: int32 index range exceeded");
  GDR_CHECK_ARG(workspace_bytes >= gdr_rank_workspace_bytes(m), GDR_E_WORKSPACE, "rank_values: workspace too small");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  double *sorted = cv.take<double>(m);
  rank_gather_kernel<<<(unsigned)ceil_div<int64_t>(m, 256), 256, 0, st>>>(values, sel, m, values_out);
  GDR_LAUNCHED();
  return rank_sort(values_out, m, sorted_out ? sorted_out : sorted, order_out, cv, st);
}

}  // extern "C"
