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
  int cta_begin[PUSH_MAX_DST];        // first CTA serving destination k (CTAs dealt in proportion to the counts)
  int n;
};

// one warp per row, 4 rows in flight per warp.  CTAs are dealt to the destinations in proportion to their row
// counts (cta_begin), so every destination is written during the whole launch: with destination-major order
// all senders would hit the same receiver at the same time (incast on one NVSwitch port while the others idle).
template <int ROWS_PER_ITER>
__global__ void __launch_bounds__(512)
push_rows_kernel(const float *__restrict__ src, int64_t d4, const PushTable t) {
  int k = 0;
  while (k + 1 < t.n && (int)blockIdx.x >= t.cta_begin[k + 1]) ++k;
  const int ctas_k = (k + 1 < t.n ? t.cta_begin[k + 1] : (int)gridDim.x) - t.cta_begin[k];
  const int cta_in_k = (int)blockIdx.x - t.cta_begin[k];
  const int64_t count = t.end[k] - (k ? t.end[k - 1] : 0);
  const int lane = threadIdx.x & 31;
  const int64_t warp = (((int64_t)cta_in_k * blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)ctas_k * blockDim.x) >> 5;
  const float4 *s4 = reinterpret_cast<const float4 *>(src);
  float4 *d4p = reinterpret_cast<float4 *>(t.dst[k]);
  const int64_t *idx = t.idx[k];
  const int64_t soff = t.src_off[k];
  for (int64_t base = warp * ROWS_PER_ITER; base < count; base += nwarps * ROWS_PER_ITER) {
    const float4 *sp[ROWS_PER_ITER];
    float4 *dp[ROWS_PER_ITER];
#pragma unroll
    for (int r = 0; r < ROWS_PER_ITER; ++r) {
      const int64_t j = base + r;
      sp[r] = nullptr;
      dp[r] = nullptr;
      if (j < count) {
        const int64_t row = idx ? __ldg(idx + j) : soff + j;
        sp[r] = s4 + row * d4;
        dp[r] = d4p + j * d4;
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

// ---- NVSwitch multicast (NVLS) --------------------------------------------------------------------------
// mc_* pointers are addresses in a multicast mapping of a symmetric allocation: a multimem.st lands at the same
// offset on every GPU of the group, a multimem.ld_reduce returns the sum of that offset over all of them
// (the switch replicates / reduces; the issuing GPU moves each element over its own links once).
__global__ void __launch_bounds__(512)
mc_broadcast_kernel(const float4 *__restrict__ src, float4 *mc_dst, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(src + i);
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_dst + i), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
  }
}

__global__ void __launch_bounds__(512)
mc_reduce_add_kernel(float4 *__restrict__ y, const float4 *mc_src, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc_src + i)
                 : "memory");
    float4 a = y[i];
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    y[i] = a;
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
  if (n_ctas <= 0) n_ctas = 96;
  if (n_ctas < n) n_ctas = n;
  // deal the CTAs: at least one per destination, the rest in proportion to the row counts
  int given = 0;
  for (int k = 0; k < n; ++k) {
    const int64_t cnt = t.end[k] - (k ? t.end[k - 1] : 0);
    int share = 1 + (int)(((int64_t)(n_ctas - n) * cnt) / total);
    const int64_t need = ceil_div<int64_t>(cnt, 4 * 16);  // 16 warps per CTA, 4 rows per warp and iteration
    if (share > need) share = (int)need;
    t.cta_begin[k] = given;
    given += share;
  }
  push_rows_kernel<4><<<(unsigned)given, 512, 0, as_stream(stream)>>>(src, d / 4, t);
  GDR_LAUNCHED();
  return GDR_OK;
}

// Replicate src[0..count) to every GPU of the multicast group: mc_dst is the multicast address of the (symmetric)
// destination region.  Replaces an all-gather of a hub block of X: one copy leaves this GPU, the switch fans out.
int gdr_mc_broadcast_f32(const float *src, float *mc_dst, int64_t count, int n_ctas, gdr_stream_t stream) {
  GDR_CHECK_ARG(count >= 0 && count % 4 == 0, GDR_E_SHAPE, "mc_broadcast: count must be a multiple of 4");
  if (count == 0) return GDR_OK;
  GDR_CHECK_ARG(src && mc_dst && ((((uintptr_t)src | (uintptr_t)mc_dst) & 15) == 0), GDR_E_BADARG,
                "mc_broadcast: null or misaligned pointer");
  if (n_ctas <= 0) n_ctas = 96;
  mc_broadcast_kernel<<<(unsigned)n_ctas, 512, 0, as_stream(stream)>>>(reinterpret_cast<const float4 *>(src),
                                                                      reinterpret_cast<float4 *>(mc_dst), count / 4);
  GDR_LAUNCHED();
  return GDR_OK;
}

// y[0..count) += sum over the GPUs of the group of the symmetric region at mc_src (in-switch fp32 reduction).
// Replaces receiving one partial-row buffer per peer and adding them locally.
int gdr_mc_reduce_add_f32(float *y, const float *mc_src, int64_t count, int n_ctas, gdr_stream_t stream) {
  GDR_CHECK_ARG(count >= 0 && count % 4 == 0, GDR_E_SHAPE, "mc_reduce_add: count must be a multiple of 4");
  if (count == 0) return GDR_OK;
  GDR_CHECK_ARG(y && mc_src && ((((uintptr_t)y | (uintptr_t)mc_src) & 15) == 0), GDR_E_BADARG,
                "mc_reduce_add: null or misaligned pointer");
  if (n_ctas <= 0) n_ctas = 148;
  mc_reduce_add_kernel<<<(unsigned)n_ctas, 512, 0, as_stream(stream)>>>(reinterpret_cast<float4 *>(y),
                                                                       reinterpret_cast<const float4 *>(mc_src),
                                                                       count / 4);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
