// kipf.cu — the row-local pieces of the Kipf & Welling graph convolution as the reference's vendored OpenNE runs it
// (multiscale/openne/gcn/layers.py:133-190, metrics.py:4-21); the contractions themselves are the library's SpMM
// (spmm.cu: support·(X·W), sparse X·W) and tcgen05 GEMM (gemm_tc.cu: dense X·W, dW, dX).
//
//   bias_act:         Y = act(Z + b)                      layers.py:186-190  (act = relu | identity)
//   relu_bwd_colsum:  dZ = dY ⊙ [Y > 0] ; db = colsum(dZ) autograd of the above
//   dropout_apply:    y = x * floor(keep + u) / keep      layers.py:22-29 (sparse_dropout) and tf.nn.dropout; the uniform
//                     numbers u come from the caller's generator
//   masked_softmax_ce: loss = mean_i(CE_i * m_i / mean(m)), dLogits, masked accuracy     metrics.py:4-21
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace gdr {

__global__ void __launch_bounds__(256)
bias_act_kernel(const float *__restrict__ z, const float *__restrict__ bias, int64_t n, int64_t d, int relu,
                float *__restrict__ y) {
  const int64_t total = n * d;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float v = z[i];
    if (bias != nullptr) v += __ldg(bias + i % d);
    y[i] = relu ? fmaxf(v, 0.f) : v;
  }
}

// one CTA per 32 columns: threads (32 x 8) walk the rows, column sums reduced through shared memory in a fixed order
__global__ void __launch_bounds__(256)
act_bwd_colsum_kernel(const float *__restrict__ dy, const float *__restrict__ y, int64_t n, int64_t d, int relu,
                      float *__restrict__ dz, float *__restrict__ db) {
  __shared__ double red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t c = (int64_t)blockIdx.x * 32 + tx;
  double s = 0.0;
  if (c < d) {
    for (int64_t r = ty; r < n; r += 8) {
      float g = dy[r * d + c];
      if (relu && !(y[r * d + c] > 0.f)) g = 0.f;
      if (dz != nullptr) dz[r * d + c] = g;
      s += (double)g;
    }
  }
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < d && db != nullptr) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += red[k][tx];
    db[c] = (float)t;
  }
}

__global__ void __launch_bounds__(256)
dropout_apply_kernel(const float *__restrict__ x, const float *__restrict__ u, int64_t count, float keep,
                     float *__restrict__ y, float *__restrict__ mask_scale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const float m = floorf(keep + u[i]) * (1.f / keep);  // tf: floor(keep_prob + uniform) in {0, 1}, scaled by 1/keep_prob
    y[i] = x[i] * m;
    if (mask_scale != nullptr) mask_scale[i] = m;
  }
}

// one warp per row; C classes strided over the lanes.  partial[row] = CE_row * mask_row; dlogits scaled by
// mask_row / (mask_sum) — the reference's mask /= mean(mask) followed by reduce_mean over N rows: 1/(N*mean) = 1/sum.
__global__ void __launch_bounds__(256)
masked_softmax_ce_kernel(const float *__restrict__ logits, const float *__restrict__ labels,
                         const float *__restrict__ mask, int64_t n, int64_t c, int64_t ld, const double *__restrict__ mask_sum,
                         double *__restrict__ row_loss, float *__restrict__ correct, float *__restrict__ dlogits) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= n) return;
  const float *lg = logits + row * ld, *lb = labels + row * c;
  float mx = -INFINITY, lmx = -INFINITY;
  int amax = INT32_MAX, lamax = INT32_MAX;
  for (int64_t j = lane; j < c; j += 32) {
    const float v = lg[j], l = lb[j];
    if (v > mx) { mx = v; amax = (int)j; }
    if (l > lmx) { lmx = l; lamax = (int)j; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {  // first maximum wins, as tf.argmax
    const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oi = __shfl_xor_sync(0xffffffffu, amax, o);
    if (ov > mx || (ov == mx && oi < amax)) { mx = ov; amax = oi; }
    const float olv = __shfl_xor_sync(0xffffffffu, lmx, o);
    const int oli = __shfl_xor_sync(0xffffffffu, lamax, o);
    if (olv > lmx || (olv == lmx && oli < lamax)) { lmx = olv; lamax = oli; }
  }
  float se = 0.f, sl = 0.f, dot = 0.f;
  for (int64_t j = lane; j < c; j += 32) {
    se += expf(lg[j] - mx);
    sl += lb[j];
    dot += lb[j] * (lg[j] - mx);
  }
  se = warp_sum(se);
  sl = warp_sum(sl);
  dot = warp_sum(dot);
  const float lse = logf(se);
  const float m = mask[row];
  const float w = (float)((double)m / *mask_sum);
  if (lane == 0) {
    row_loss[row] = (double)(sl * lse - dot) * (double)m;  // -sum_j labels_j * log_softmax_j
    if (correct != nullptr) correct[row] = (amax == lamax ? 1.f : 0.f) * m;
  }
  if (dlogits != nullptr) {
    for (int64_t j = lane; j < ld; j += 32)
      dlogits[row * ld + j] = j < c ? (expf(lg[j] - mx - lse) * sl - lb[j]) * w : 0.f;
  }
}

__global__ void masked_ce_finalize_kernel(const double *__restrict__ row_loss, const float *__restrict__ correct,
                                          int64_t n, const double *__restrict__ mask_sum, float *__restrict__ loss_out,
                                          float *__restrict__ acc_out) {
  __shared__ double red[2][32];
  double s = 0.0, a = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
    s += row_loss[i];
    if (correct != nullptr) a += (double)correct[i];
  }
  s = warp_sum(s);
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = s;
    red[1][threadIdx.x >> 5] = a;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, ta = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      ts += red[0][w];
      ta += red[1][w];
    }
    *loss_out = (float)(ts / *mask_sum);
    if (acc_out != nullptr) *acc_out = (float)(ta / *mask_sum);
  }
}

__global__ void mask_sum_kernel(const float *__restrict__ mask, int64_t n, double *__restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += (double)mask[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    *out = t;
  }
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int gdr_bias_act_f32(const float *z, const float *bias, int64_t n, int64_t d, int relu, float *y, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && z && y, GDR_E_BADARG, "bias_act: bad args");
  if (n == 0) return GDR_OK;
  const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div<int64_t>(n * d, 256), (int64_t)kNumSMs * 16);
  bias_act_kernel<<<blocks, 256, 0, as_stream(stream)>>>(z, bias, n, d, relu, y);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_act_bwd_colsum_f32(const float *dy, const float *y, int64_t n, int64_t d, int relu, float *dz, float *db,
                           gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && dy && (y || !relu), GDR_E_BADARG, "act_bwd_colsum: bad args");
  act_bwd_colsum_kernel<<<(unsigned)ceil_div<int64_t>(d, 32), 256, 0, as_stream(stream)>>>(dy, y, n, d, relu, dz, db);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_dropout_apply_f32(const float *x, const float *uniform, int64_t count, float keep_prob, float *y,
                          float *mask_scale_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(count >= 0 && x && uniform && y, GDR_E_BADARG, "dropout_apply: bad args");
  GDR_CHECK_ARG(keep_prob > 0.f && keep_prob <= 1.f, GDR_E_BADARG, "dropout_apply: keep_prob must be in (0, 1]");
  if (count == 0) return GDR_OK;
  const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div<int64_t>(count, 256), (int64_t)kNumSMs * 16);
  dropout_apply_kernel<<<blocks, 256, 0, as_stream(stream)>>>(x, uniform, count, keep_prob, y, mask_scale_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int64_t gdr_masked_softmax_ce_workspace_bytes(int64_t n) {
  if (n < 0) return GDR_E_BADARG;
  return align_up(n * 8, 256) + align_up(n * 4, 256) + 512;
}

int gdr_masked_softmax_ce_f32(const float *logits, int64_t ld, const float *labels, const float *mask, int64_t n, int64_t c,
                              float *loss_out, float *accuracy_out, float *dlogits_out, void *workspace,
                              int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && c > 0 && ld >= c && logits && labels && mask && loss_out && workspace, GDR_E_BADARG,
                "masked_softmax_ce: bad args");
  GDR_CHECK_ARG(workspace_bytes >= gdr_masked_softmax_ce_workspace_bytes(n), GDR_E_WORKSPACE,
                "masked_softmax_ce: workspace too small");
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  double *row_loss = cv.take<double>(n);
  float *correct = cv.take<float>(n);
  double *msum = cv.take<double>(1);
  mask_sum_kernel<<<1, 1024, 0, st>>>(mask, n, msum);
  GDR_LAUNCHED();
  masked_softmax_ce_kernel<<<(unsigned)ceil_div<int64_t>(n, 8), 256, 0, st>>>(logits, labels, mask, n, c, ld, msum, row_loss,
                                                                           correct, dlogits_out);
  GDR_LAUNCHED();
  masked_ce_finalize_kernel<<<1, 1024, 0, st>>>(row_loss, correct, n, msum, loss_out, accuracy_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
