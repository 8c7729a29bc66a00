// p2p.cu — the exchange step of the partitioned SpMM as direct stores into peer memory over
// NVLink 5 / NVSwitch (no staging buffer, no pack pass): rows of the local block are read by index
// and written straight into the destination GPU's receive region.  The peer pointers come from a
// symmetric-memory rendezvous done by the host side (torch.distributed._symmetric_memory); ordering
// across GPUs is the host side's barrier on the same stream (kernel completion orders the stores).
#include "common.cuh"

namespace gdr {

constexpr int PUSH_MAX_DST = 16;
struct PushTable {
  float *dst[PUSH_MAX_DST];           // peer (or local) destination base, row-major [count, d]
  const int64_t *idx[PUSH_MAX_DST];   // source row ids (device), or NULL: rows src_off .. src_off+count-1
  int64_t src_off[PUSH_MAX_DST];
  int64_t end[PUSH_MAX_DST];          // exclusive prefix ends over the concatenated row space
  int n;
};

// one warp per row, 4 rows in flight per warp; consecutive warps write consecutive remote rows
template <int ROWS_PER_ITER>
__global__ void __launch_bounds__(512)
push_rows_kernel(const float *__restrict__ src, int64_t d4, const PushTable t, int64_t total) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const float4 *s4 = reinterpret_cast<const float4 *>(src);
  for (int64_t base = warp * ROWS_PER_ITER; base < total; base += nwarps * ROWS_PER_ITER) {
    const float4 *sp[ROWS_PER_ITER];
    float4 *dp[ROWS_PER_ITER];
#pragma unroll
    for (int r = 0; r < ROWS_PER_ITER; ++r) {
      const int64_t i = base + r;
      sp[r] = nullptr;
      dp[r] = nullptr;
      if (i < total) {
        int k = 0;
        while (i >= t.end[k]) ++k;
        const int64_t j = i - (k ? t.end[k - 1] : 0);
        const int64_t row = t.idx[k] ? __ldg(t.idx[k] + j) : t.src_off[k] + j;
        sp[r] = s4 + row * d4;
        dp[r] = reinterpret_cast<float4 *>(t.dst[k]) + j * d4;
      }
    }
    for (int64_t c = lane; c < d4; c += 32) {
      float4 v[ROWS_PER_ITER];
#pragma unroll
      for (int r = 0; r < ROWS_PER_ITER; ++r)
        if (sp[r]) v[r] = __ldg(sp[r] + c);
#pragma unroll
      for (int r = 0; r < ROWS_PER_ITER; ++r)
        if (dp[r]) dp[r][c] = v[r];
    }
  }
}

}  // namespace gdr

using namespace gdr;

extern "C" {

// dst[k][j, :] = src[idx[k][j], :]  (idx[k] == NULL: src[src_off[k] + j, :]) for k < n_dst, j < counts[k].
// dst pointers may be peer-GPU addresses (P2P / symmetric memory).  The tables are HOST arrays.
int gdr_push_rows_f32(const float *src, int64_t d, int n_dst, float *const *dst, const int64_t *const *idx,
                      const int64_t *src_off, const int64_t *counts, int n_ctas, gdr_stream_t stream) {
  GDR_CHECK_ARG(n_dst >= 0 && n_dst <= PUSH_MAX_DST, GDR_E_SHAPE, "push_rows: at most %d destinations", PUSH_MAX_DST);
  GDR_CHECK_ARG(d > 0 && d % 4 == 0, GDR_E_SHAPE, "push_rows: d must be a multiple of 4");
  if (n_dst == 0) return GDR_OK;
  GDR_CHECK_ARG(src && dst && idx && src_off && counts, GDR_E_BADARG, "push_rows: null table");
  PushTable t{};
  int64_t total = 0;
  int n = 0;
  for (int k = 0; k < n_dst; ++k) {
    GDR_CHECK_ARG(counts[k] >= 0, GDR_E_BADARG, "push_rows: negative count");
    if (counts[k] == 0) continue;
    GDR_CHECK_ARG(dst[k] != nullptr, GDR_E_BADARG, "push_rows: null destination %d", k);
    t.dst[n] = dst[k]; t.idx[n] = idx[k]; t.src_off[n] = src_off[k];
    total += counts[k];
    t.end[n] = total;
    ++n;
  }
  t.n = n;
  if (total == 0) return GDR_OK;
  if (n_ctas <= 0) n_ctas = 64;
  const int64_t need = ceil_div<int64_t>(total, 4 * 16);  // 16 warps per CTA, 4 rows per warp and iteration
  if (n_ctas > need) n_ctas = (int)need;
  push_rows_kernel<4><<<(unsigned)n_ctas, 512, 0, as_stream(stream)>>>(src, d / 4, t, total);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
