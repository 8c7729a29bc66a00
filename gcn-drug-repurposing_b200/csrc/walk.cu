// walk.cu — node2vec / DeepWalk walk generation (multiscale/openne/walker.py:11-161), the stage that feeds the
// embeddings train.py reads.
//
// The reference precomputes one alias table per NODE (first step) and one per EDGE (second-order steps,
// get_alias_edge :106-125: weight/p back to the previous node, weight for neighbours the previous node is linked from,
// weight/q otherwise) — O(sum deg^2) memory — and walks in Python.  Here a walk is one thread and nothing per edge
// is stored: a candidate is drawn from the FIRST-order distribution of the current node (inverse CDF over the row's
// running weight sums, binary search) and accepted with probability factor / max(1/p, 1, 1/q) — rejection sampling
// whose accepted draws have exactly the reference's normalised second-order probabilities.  has_edge(x, prev) is a
// binary search in x's sorted adjacency row.  The random numbers come from a counter-based generator keyed by (seed,
// walk, draw), so a walk does not depend on scheduling; they are NOT numpy's stream: parity with the reference is
// distributional (tests compare transition frequencies with the probabilities of the reference's alias tables).
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace gdr {

__device__ __forceinline__ uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
// uniform in [0, 1) with 24 random bits, a pure function of (seed, walk, counter)
__device__ __forceinline__ float walk_uniform(uint64_t seed, uint64_t walk, uint32_t ctr) {
  const uint64_t h = mix64(mix64(seed + 0x9e3779b97f4a7c15ull * (walk + 1)) + ctr);
  return (float)(h >> 40) * (1.0f / 16777216.0f);
}

// cum[e] = running sum of the row's weights up to and including entry e (fp32, one warp per row would be overkill:
// MSI rows have a median of 5 entries)
__global__ void __launch_bounds__(256)
row_cumsum_kernel(const int32_t *__restrict__ rowptr, const float *__restrict__ w, int64_t n, float *__restrict__ cum) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  float s = 0.f;
  for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
    s += w ? w[e] : 1.f;
    cum[e] = s;
  }
}

__device__ __forceinline__ bool has_edge(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col, int32_t u,
                                         int32_t v) {
  int32_t lo = rowptr[u], hi = rowptr[u + 1];
  while (lo < hi) {
    const int32_t mid = (lo + hi) >> 1;
    const int32_t c = col[mid];
    if (c == v) return true;
    if (c < v) lo = mid + 1; else hi = mid;
  }
  return false;
}

// first-order draw from row `u`: smallest entry with cum > x * total
__device__ __forceinline__ int32_t draw_neighbour(const int32_t *__restrict__ rowptr, const float *__restrict__ cum, int32_t u,
                                                  float x) {
  const int32_t s = rowptr[u], e = rowptr[u + 1];
  const float t = x * cum[e - 1];
  int32_t lo = s, hi = e - 1;
  while (lo < hi) {
    const int32_t mid = (lo + hi) >> 1;
    if (cum[mid] > t) hi = mid; else lo = mid + 1;
  }
  return lo;
}

__global__ void __launch_bounds__(128)
node2vec_walk_kernel(const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col, const float *__restrict__ cum,
                     const int32_t *__restrict__ start, int64_t n_walks, int walk_length, float inv_p, float inv_q,
                     int second_order, uint64_t seed, int64_t walk_id0, int32_t *__restrict__ walks) {
  const int64_t wid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (wid >= n_walks) return;
  int32_t *out = walks + wid * walk_length;
  int32_t cur = start[wid], prev = -1;
  out[0] = cur;
  uint32_t ctr = 0;
  const float fmax = fmaxf(1.f, fmaxf(inv_p, inv_q));
  int len = 1;
  for (; len < walk_length; ++len) {
    if (rowptr[cur + 1] == rowptr[cur]) break;  // no out-neighbour: the walk ends here (walker.py:93-94)
    int32_t nxt;
    if (!second_order || prev < 0) {
      nxt = col[draw_neighbour(rowptr, cum, cur, walk_uniform(seed, (uint64_t)(walk_id0 + wid), ctr++))];
    } else {
      for (;;) {
        nxt = col[draw_neighbour(rowptr, cum, cur, walk_uniform(seed, (uint64_t)(walk_id0 + wid), ctr++))];
        const float f = nxt == prev ? inv_p : (has_edge(rowptr, col, nxt, prev) ? 1.f : inv_q);
        if (f >= fmax) break;                   // no second number needed for the largest factor
        if (walk_uniform(seed, (uint64_t)(walk_id0 + wid), ctr++) * fmax < f) break;
      }
    }
    out[len] = nxt;
    prev = cur;
    cur = nxt;
  }
  for (; len < walk_length; ++len) out[len] = -1;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int gdr_walk_row_cumsum_f32(const int32_t *rowptr, const float *weight, int64_t n, float *cum_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && rowptr && cum_out, GDR_E_BADARG, "walk_row_cumsum: bad args");
  if (n == 0) return GDR_OK;
  row_cumsum_kernel<<<(unsigned)ceil_div<int64_t>(n, 256), 256, 0, as_stream(stream)>>>(rowptr, weight, n, cum_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_node2vec_walks(const int32_t *rowptr, const int32_t *col, const float *cum, int64_t n, const int32_t *start,
                       int64_t n_walks, int64_t walk_length, double p, double q, int second_order, uint64_t seed,
                       int64_t walk_id0, int32_t *walks_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && n_walks >= 0 && walk_length >= 1 && rowptr && col && cum && start && walks_out, GDR_E_BADARG,
                "node2vec_walks: bad args");
  GDR_CHECK_ARG(p > 0.0 && q > 0.0, GDR_E_BADARG, "node2vec_walks: p and q must be positive");
  GDR_CHECK_ARG(walk_length <= INT32_MAX, GDR_E_SHAPE, "node2vec_walks: walk_length too large");
  if (n_walks == 0) return GDR_OK;
  node2vec_walk_kernel<<<(unsigned)ceil_div<int64_t>(n_walks, 128), 128, 0, as_stream(stream)>>>(
      rowptr, col, cum, start, n_walks, (int)walk_length, (float)(1.0 / p), (float)(1.0 / q), second_order, seed, walk_id0,
      walks_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
