// loss.cu — GSS similarity-separation loss (modules/model.py:214-221), its gradient, the
// batch gather/scatter around it, and the Adam step (train.py:139,184).
//
//   E = embs[index] ; G = E·Eᵀ ; loss = mean(-0.5*alpha*(relu(G)-beta)^2)
//   dLoss/dG = -alpha*(relu(G)-beta)*[G>0] / B^2 ;  dLoss/dE = (dG + dGᵀ)·E = 2·dG·E  (G symmetric)
//
// Forward and gradient come out of one call: the Gram GEMM's epilogue turns each G tile
// into (loss partial, 2·dG tile) in registers, a second GEMM contracts 2·dG with E.
#include <stdlib.h>

#include <algorithm>
#include <atomic>

#include "gemm_tc.cuh"

namespace gdr {

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float *__restrict__ src, const int64_t *__restrict__ index, int64_t b, int64_t n,
                   int64_t d, float *__restrict__ dst) {
  const int lane = threadIdx.x & 31;
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= b) return;
  int64_t r = index ? index[i] : i;
  if (r < 0) r += n;  // torch-style negative indices
  const float *s = src + r * d;
  float *o = dst + i * d;
  for (int64_t c = lane * 4; c < d; c += 128) *reinterpret_cast<float4 *>(o + c) = ldg_f4(s + c);
}

__global__ void __launch_bounds__(256)
scatter_rows_kernel(const float *__restrict__ ge, const int64_t *__restrict__ index, int64_t b, int64_t n,
                    int64_t d, const float *__restrict__ scale_dev, float *__restrict__ gh) {
  const int lane = threadIdx.x & 31;
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= b) return;
  const float sc = scale_dev ? __ldg(scale_dev) : 1.f;
  int64_t r = index ? index[i] : i;
  if (r < 0) r += n;
  const float *s = ge + i * d;
  float *o = gh + r * d;
  // atomics: duplicate indices in a user-supplied batch must accumulate (index_put_ accumulate=True)
  for (int64_t c = lane; c < d; c += 32) atomicAdd(o + c, sc * s[c]);
}

__global__ void loss_finalize_kernel(const double *__restrict__ partials, int count, double inv_count,
                                     float *__restrict__ loss_out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) s += partials[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += red[w];
    *loss_out = (float)(tot * inv_count);
  }
}

// torch.optim.Adam (amsgrad=False, weight_decay=0, maximize=False), single-tensor formulas:
//   m = lerp(m, g, 1-b1) ; v = b2*v + (1-b2)*g*g ; denom = sqrt(v)/sqrt(1-b2^t) + eps
//   p -= (lr/(1-b1^t)) * m/denom
__global__ void __launch_bounds__(256)
adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
            int64_t count, float lr, float b1, float b2, float eps, const int32_t *__restrict__ step_dev) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const int step = *step_dev + 1;
  const double bc1 = 1.0 - pow((double)b1, (double)step);
  const double bc2 = 1.0 - pow((double)b2, (double)step);
  const float step_size = (float)((double)lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  const float gi = g[i];
  float mi = m[i], vi = v[i];
  mi = mi + (gi - mi) * (1.f - b1);
  vi = vi * b2 + (1.f - b2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  const float denom = sqrtf(vi) / bc2_sqrt + eps;
  p[i] = p[i] - step_size * (mi / denom);
}

__global__ void bump_step_kernel(int32_t *step_dev) { *step_dev += 1; }

// all parameter tensors of the model in ONE launch (train.py:139 builds Adam over 4 tensors)
constexpr int ADAM_MAX_TENSORS = 16;
struct AdamMulti {
  float *p[ADAM_MAX_TENSORS];
  const float *g[ADAM_MAX_TENSORS];
  float *m[ADAM_MAX_TENSORS];
  float *v[ADAM_MAX_TENSORS];
  int64_t end[ADAM_MAX_TENSORS];  // exclusive prefix ends of the concatenated index space
  int n;
};
__global__ void __launch_bounds__(256)
adam_multi_kernel(const AdamMulti a, int64_t total, float step_size, float bc2_sqrt, float omb1, float b2, float omb2,
                  float eps) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int t = 0;
  while (i >= a.end[t]) ++t;
  const int64_t j = i - (t ? a.end[t - 1] : 0);
  const float gi = a.g[t][j];
  float mi = a.m[t][j], vi = a.v[t][j];
  mi = mi + (gi - mi) * omb1;                              // exp_avg.lerp_(grad, 1 - beta1)
  vi = __fadd_rn(__fmul_rn(vi, b2), omb2 * gi * gi);       // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  a.m[t][j] = mi;
  a.v[t][j] = vi;
  const float denom = sqrtf(vi) / bc2_sqrt + eps;
  a.p[t][j] = a.p[t][j] - step_size * (mi / denom);
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int64_t gdr_gss_loss_workspace_bytes(int64_t b, int64_t d) {
  if (b < 0 || d <= 0) return GDR_E_BADARG;
  const int64_t ldw = align_up(b, 4);
  const int64_t tiles = ceil_div<int64_t>(b > 0 ? b : 1, GBM);
  return align_up(b * d * 4, 256) + align_up(b * ldw * 4, 256) + align_up(tiles * tiles * 8, 256) + 1024;
}

int gdr_gss_loss_fwd_bwd_f32(const float *embs, const int64_t *index, int64_t b, int64_t n, int64_t d, float alpha,
                             float beta, float *loss_out, float *grad_e_out, void *workspace,
                             int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(b > 0 && n > 0 && d > 0, GDR_E_BADARG, "gss_loss: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "gss_loss: d must be a multiple of 4");
  GDR_CHECK_ARG(embs && loss_out && grad_e_out && workspace, GDR_E_BADARG, "gss_loss: null pointer");
  GDR_CHECK_ARG(workspace_bytes >= gdr_gss_loss_workspace_bytes(b, d), GDR_E_WORKSPACE,
                "gss_loss: workspace too small (B=%lld needs a BxB tile buffer)", (long long)b);
  cudaStream_t st = as_stream(stream);
  const int64_t ldw = align_up(b, 4);
  const int64_t tiles = ceil_div<int64_t>(b, GBM);
  Carver cv(workspace);
  float *e = cv.take<float>(b * d);
  float *wm = cv.take<float>(b * ldw);
  double *partials = cv.take<double>(tiles * tiles);

  gather_rows_kernel<<<(unsigned)ceil_div<int64_t>(b * 32, 256), 256, 0, st>>>(embs, index, b, n, d, e);
  GDR_LAUNCHED();

  GemmParams g{};
  g.a[0] = e; g.lda[0] = d; g.b[0] = e; g.ldb[0] = d; g.kseg[0] = d; g.nseg = 1;
  g.m = b; g.n = b; g.c = wm; g.ldc = ldw;
  g.alpha = alpha; g.beta = beta;
  g.wscale = (float)(2.0 / ((double)b * (double)b));
  g.loss_partials = partials;
  dim3 grid((unsigned)tiles, (unsigned)tiles, 1);
  gemm_f32_kernel<true, true, EPI_GSS><<<grid, GTHREADS, 0, st>>>(g);
  GDR_LAUNCHED();

  loss_finalize_kernel<<<1, 256, 0, st>>>(partials, (int)(tiles * tiles), 1.0 / ((double)b * (double)b), loss_out);
  GDR_LAUNCHED();

  GemmParams q{};
  q.a[0] = wm; q.lda[0] = ldw;  // A(m=i, k=j) = Wm[i*ldw + j]
  q.b[0] = e; q.ldb[0] = d;     // B(n=c, k=j) = E[j*d + c]
  q.kseg[0] = b; q.nseg = 1;
  q.m = b; q.n = d; q.c = grad_e_out; q.ldc = d;
  dim3 grid2((unsigned)ceil_div<int64_t>(d, GBN), (unsigned)tiles, 1);
  gemm_f32_kernel<true, false, EPI_STORE><<<grid2, GTHREADS, 0, st>>>(q);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_gather_rows_f32(const float *src, const int64_t *index, int64_t b, int64_t n, int64_t d, float *dst,
                        gdr_stream_t stream) {
  GDR_CHECK_ARG(b >= 0 && n >= 0 && d > 0 && d % 4 == 0, GDR_E_BADARG, "gather_rows: bad sizes");
  if (b == 0) return GDR_OK;
  GDR_CHECK_ARG(src && dst, GDR_E_BADARG, "gather_rows: null pointer");
  gather_rows_kernel<<<(unsigned)ceil_div<int64_t>(b * 32, 256), 256, 0, as_stream(stream)>>>(src, index, b, n, d, dst);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_scatter_rows_f32(const float *ge, const int64_t *index, int64_t b, int64_t n, int64_t d,
                         const float *scale_dev, float *gh, gdr_stream_t stream) {
  GDR_CHECK_ARG(b >= 0 && d > 0 && ge && gh, GDR_E_BADARG, "scatter_rows: bad args");
  if (b == 0) return GDR_OK;
  scatter_rows_kernel<<<(unsigned)ceil_div<int64_t>(b * 32, 256), 256, 0, as_stream(stream)>>>(
      ge, index, b, n, d, scale_dev, gh);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_adam_step_f32(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, int64_t count, float lr,
                      float beta1, float beta2, float eps, int32_t *step_dev, gdr_stream_t stream) {
  GDR_CHECK_ARG(count >= 0 && param && grad && exp_avg && exp_avg_sq && step_dev, GDR_E_BADARG,
                "adam_step: bad args");
  if (count == 0) return GDR_OK;
  cudaStream_t st = as_stream(stream);
  adam_kernel<<<(unsigned)ceil_div<int64_t>(count, 256), 256, 0, st>>>(param, grad, exp_avg, exp_avg_sq, count, lr,
                                                                      beta1, beta2, eps, step_dev);
  GDR_LAUNCHED();
  bump_step_kernel<<<1, 1, 0, st>>>(step_dev);
  GDR_LAUNCHED();
  return GDR_OK;
}

// ---- tcgen05 variant: Gram and dG·E on the tensor cores (3xTF32, csrc/gemm_tc.cu) -------------------------
// The B x B matrices are never stored: the rows are processed in PANELS of R rows (R * B * 4 bytes <= 1 GiB).  Per
// panel one Gram GEMM (epilogue: loss partials + the 2·dG panel) and one product dE[panel] = (2·dG panel)·E; the loss
// partials of a panel are folded into one fp64 accumulator in a fixed order.  `index == NULL` with B = N = 2^20
// (train.py --batch-size 0, SURVEY 8 a9: "10^12 entries, must be tiled, never stored") needs 4 GB of panel instead of
// 4 TB; a B that fits one panel (every minibatch) takes exactly the two GEMMs it always did.
constexpr int64_t GSS_PANEL_BYTES = (int64_t)1 << 30;

static int64_t gss_panel_rows(int64_t b) {
  const int64_t ldw = align_up(b > 0 ? b : 1, 4);
  int64_t r = GSS_PANEL_BYTES / (ldw * 4);
  if (const char *env = getenv("GDR_GSS_PANEL_ROWS")) r = atoll(env);  // tests: force several panels at small B
  r = r / 128 * 128;
  if (r < 128) r = 128;
  return r < b ? r : b;
}

// dE panel = 2dG panel · E has few output tiles (rows/128 x d/128) and K = B: cut K so that every SM has a tile
static void gss_splitk_plan(int64_t rows, int64_t d, int64_t k, int &splits, int64_t &k_per_split) {
  const int64_t tiles = ceil_div<int64_t>(rows, 128) * ceil_div<int64_t>(d, 128);
  int64_t want = ceil_div<int64_t>(kNumSMs, tiles);
  const int64_t kblocks = ceil_div<int64_t>(k, 32);
  want = std::max<int64_t>(1, std::min<int64_t>(want, ceil_div<int64_t>(kblocks, 16)));  // >= 16 k-blocks per range
  k_per_split = ceil_div<int64_t>(kblocks, want) * 32;
  splits = (int)ceil_div<int64_t>(k, k_per_split);
}

// upper bound of splits * r * d for every panel height r <= rows: splits <= 148 / tiles + 1 and r * d <= tiles * 128^2
static int64_t gss_part_floats(int64_t rows, int64_t d) {
  const int64_t tiles = ceil_div<int64_t>(rows > 0 ? rows : 1, 128) * ceil_div<int64_t>(d, 128);
  return (kNumSMs + tiles) * 128 * 128;
}

// out[i] = sum_z part[z][i], z ascending (deterministic)
__global__ void __launch_bounds__(256)
gss_splitk_reduce_kernel(const float *__restrict__ part, int splits, int64_t count, float *__restrict__ out) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= count) return;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int z = 0; z < splits; ++z) {
    const float4 v = ldg_f4(part + (int64_t)z * count + i);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  *reinterpret_cast<float4 *>(out + i) = s;
}

__global__ void loss_accumulate_kernel(const double *__restrict__ partials, int64_t count, double *__restrict__ acc,
                                       int first) {
  __shared__ double red[32];
  double s = 0.0;
  for (int64_t i = threadIdx.x; i < count; i += blockDim.x) s += partials[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += red[w];
    *acc = first ? tot : *acc + tot;
  }
}

// workspace for a given panel height: the GDR_GSS_PANEL_ROWS override is read ONCE per entry-point call (by
// gss_panel_rows) and everything else derives from that value, so the size check of the run function and its carving
// cannot disagree even if the environment changes between the size query and the run.
static int64_t gss_tc_workspace_for(int64_t b, int64_t d, int64_t r) {
  const int64_t ldw = align_up(b, 4);
  const int64_t tiles_b = ceil_div<int64_t>(b > 0 ? b : 1, 128), tiles_r = ceil_div<int64_t>(r > 0 ? r : 1, 128);
  return 3 * align_up(b * d * 4, 256) + align_up(r * ldw * 4, 256) + 2 * align_up(d * ldw * 4, 256) +
         align_up(tiles_r * tiles_b * 8, 256) + align_up(gss_part_floats(r, d) * 4, 256) + 2048;
}

int64_t gdr_gss_loss_tc_workspace_bytes(int64_t b, int64_t d) {
  if (b < 0 || d <= 0) return GDR_E_BADARG;
  return gss_tc_workspace_for(b, d, gss_panel_rows(b));
}

int gdr_gss_loss_fwd_bwd_tc_f32(const float *embs, const int64_t *index, int64_t b, int64_t n, int64_t d, float alpha,
                                float beta, float *loss_out, float *grad_e_out, void *workspace,
                                int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(b > 0 && n > 0 && d > 0, GDR_E_BADARG, "gss_loss_tc: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "gss_loss_tc: d must be a multiple of 4");
  GDR_CHECK_ARG(embs && loss_out && grad_e_out && workspace, GDR_E_BADARG, "gss_loss_tc: null pointer");
  const int64_t rows = gss_panel_rows(b);
  GDR_CHECK_ARG(workspace_bytes >= gss_tc_workspace_for(b, d, rows), GDR_E_WORKSPACE,
                "gss_loss_tc: workspace too small");
  cudaStream_t st = as_stream(stream);
  const int64_t ldw = align_up(b, 4);
  const int64_t tiles_b = ceil_div<int64_t>(b, 128);
  Carver cv(workspace);
  float *e = cv.take<float>(b * d), *ehi = cv.take<float>(b * d), *elo = cv.take<float>(b * d);
  float *wm = cv.take<float>(rows * ldw);
  float *ethi = cv.take<float>(d * ldw), *etlo = cv.take<float>(d * ldw);
  double *partials = cv.take<double>(ceil_div<int64_t>(rows, 128) * tiles_b);
  double *acc = cv.take<double>(1);
  float *part = cv.take<float>(gss_part_floats(rows, d));
  const float *esrc = embs;  // index == NULL: E is embs itself, no copy
  if (index != nullptr) {
    gather_rows_kernel<<<(unsigned)ceil_div<int64_t>(b * 32, 256), 256, 0, st>>>(embs, index, b, n, d, e);
    GDR_LAUNCHED();
    esrc = e;
  }
  int rc;
  if ((rc = tc_split(esrc, d, d, nullptr, 0, 0, b, ehi, elo, st))) return rc;
  if ((rc = tc_split_transpose(esrc, d, b, d, ldw, ethi, etlo, st))) return rc;  // Eᵀ [d, ldw], zero padded
  for (int64_t r0 = 0; r0 < b; r0 += rows) {
    const int64_t r = std::min<int64_t>(rows, b - r0);
    const int64_t tiles_r = ceil_div<int64_t>(r, 128);
    GemmParams g{};
    g.c = wm; g.ldc = ldw;
    g.alpha = alpha; g.beta = beta;
    g.wscale = (float)(2.0 / ((double)b * (double)b));
    g.loss_partials = partials;
    // G panel = E[r0:r0+r]·Eᵀ -> loss partials, 2·dG panel
    if ((rc = tc_gemm(EPI_GSS, ehi + r0 * d, elo + r0 * d, ehi, elo, r, b, d, d, d, d, 1, g, st))) return rc;
    loss_accumulate_kernel<<<1, 1024, 0, st>>>(partials, tiles_r * tiles_b, acc, r0 == 0 ? 1 : 0);
    GDR_LAUNCHED();
    // dE[panel] = 2·dG panel · E   (the panel is split into TF32 hi/lo inside the kernel; K = B is cut into ranges so
    // that the few output tiles of a panel still occupy every SM, partials reduced in range order)
    int splits;
    int64_t kps;
    gss_splitk_plan(r, d, ldw, splits, kps);
    if (splits == 1) {
      GemmParams q{};
      q.c = grad_e_out + r0 * d; q.ldc = d;
      if ((rc = tc_gemm_raw_a(EPI_STORE, wm, ldw, ldw, nullptr, 0, 0, ethi, etlo, ldw, r, d, q, st))) return rc;
    } else {
      if ((rc = tc_gemm_raw_a_splitk(wm, ldw, ldw, ethi, etlo, ldw, r, d, kps, splits, part, st))) return rc;
      gss_splitk_reduce_kernel<<<(unsigned)ceil_div<int64_t>(r * d / 4, 256), 256, 0, st>>>(part, splits, r * d,
                                                                                         grad_e_out + r0 * d);
      GDR_LAUNCHED();
    }
  }
  loss_finalize_kernel<<<1, 32, 0, st>>>(acc, 1, 1.0 / ((double)b * (double)b), loss_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_adam_step_multi_f32(int n_tensors, float *const *params, const float *const *grads, float *const *exp_avg,
                            float *const *exp_avg_sq, const int64_t *counts, int64_t step, double lr, double beta1,
                            double beta2, double eps, gdr_stream_t stream) {
  GDR_CHECK_ARG(n_tensors >= 0 && n_tensors <= ADAM_MAX_TENSORS, GDR_E_SHAPE, "adam_step_multi: at most %d tensors",
                ADAM_MAX_TENSORS);
  GDR_CHECK_ARG(step >= 1, GDR_E_BADARG, "adam_step_multi: step counts from 1");
  if (n_tensors == 0) return GDR_OK;
  GDR_CHECK_ARG(params && grads && exp_avg && exp_avg_sq && counts, GDR_E_BADARG, "adam_step_multi: null table");
  AdamMulti a{};
  int64_t total = 0;
  for (int t = 0; t < n_tensors; ++t) {
    GDR_CHECK_ARG(params[t] && grads[t] && exp_avg[t] && exp_avg_sq[t] && counts[t] >= 0, GDR_E_BADARG,
                  "adam_step_multi: tensor %d", t);
    a.p[t] = params[t]; a.g[t] = grads[t]; a.m[t] = exp_avg[t]; a.v[t] = exp_avg_sq[t];
    total += counts[t];
    a.end[t] = total;
  }
  a.n = n_tensors;
  if (total == 0) return GDR_OK;
  // torch/optim/adam.py (_single_tensor_adam): every scalar (1 - beta, bias corrections, step size) is a python
  // double, rounded to fp32 only when it meets the tensor
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  adam_multi_kernel<<<(unsigned)ceil_div<int64_t>(total, 256), 256, 0, as_stream(stream)>>>(
      a, total, (float)(lr / bc1), (float)sqrt(bc2), (float)(1.0 - beta1), (float)beta2, (float)(1.0 - beta2),
      (float)eps);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"

// ---- GSS loss without any allocation: scratch lives inside the library ---------------------------------------
// Who this is for: code that may not (or cannot) obtain device memory at call time.  Examples: a plug-in hosted by a
// process whose memory pool is owned by somebody else and is not exposed to the plug-in, a real-time / serving loop
// that forbids allocator calls after start-up because cudaMalloc takes a device-wide lock and may synchronise, and a
// bare C or Fortran host that links libgdr.so but carries no allocator of its own.  Such a caller can still hand over
// its inputs and outputs (it has those anyway); what it cannot supply is the B x d + B x B workspace that
// gdr_gss_loss_fwd_bwd_f32 asks for.  Here that workspace is a fixed array in the library's own device image.
//
// Cost and limit: the array below is part of the module, so EVERY GPU on which this library's kernels get loaded keeps
// GSS_STATIC_BYTES (about 18.1 MiB) reserved until the process exits, used or not.  It is sized for the training
// batch of the reference (train.py --batch-size 2048, hidden 256): B <= 2048 and B*d <= 2048*256.  Anything larger
// is refused with GDR_E_SHAPE; use the workspace form for it.
constexpr int64_t GSS_STATIC_MAX_B = 2048;
constexpr int64_t GSS_STATIC_MAX_BD = GSS_STATIC_MAX_B * 256;
constexpr int64_t GSS_STATIC_E_BYTES = GSS_STATIC_MAX_BD * 4;                           // gathered rows E [B, d]
constexpr int64_t GSS_STATIC_W_BYTES = GSS_STATIC_MAX_B * GSS_STATIC_MAX_B * 4;         // 2·dG [B, ldw]
constexpr int64_t GSS_STATIC_P_BYTES = (GSS_STATIC_MAX_B / GBM) * (GSS_STATIC_MAX_B / GBM) * 8;  // one fp64 per CTA
constexpr int64_t GSS_STATIC_BYTES = GSS_STATIC_E_BYTES + GSS_STATIC_W_BYTES + GSS_STATIC_P_BYTES;
static_assert(GSS_STATIC_E_BYTES % 256 == 0 && GSS_STATIC_W_BYTES % 256 == 0, "sub-buffers stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_gss_static_scratch[GSS_STATIC_BYTES];

// In host code `g_gss_static_scratch` is only a handle for the runtime, not an address a kernel can use: the device
// address differs per GPU and has to be asked for with that GPU current.  One slot per device ordinal.
constexpr int GSS_STATIC_MAX_DEVICES = 64;
static std::atomic<unsigned char *> g_gss_static_ptr[GSS_STATIC_MAX_DEVICES];

static int gss_static_scratch(unsigned char **out) {
  int dev = -1;
  GDR_CUDA(cudaGetDevice(&dev));
  GDR_CHECK_ARG(dev >= 0 && dev < GSS_STATIC_MAX_DEVICES, GDR_E_INTERNAL,
                "gss_loss_static: device ordinal %d outside the %d slots of the address table", dev,
                GSS_STATIC_MAX_DEVICES);
  unsigned char *p = g_gss_static_ptr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {  // first call on this device (two threads racing here both get the same answer)
    void *sym = nullptr;
    GDR_CUDA(cudaGetSymbolAddress(&sym, g_gss_static_scratch));
    GDR_CHECK_ARG(sym != nullptr, GDR_E_INTERNAL, "gss_loss_static: cudaGetSymbolAddress returned a null address");
    p = static_cast<unsigned char *>(sym);
    g_gss_static_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return GDR_OK;
}

extern "C" {

// Same result as gdr_gss_loss_fwd_bwd_f32 (same kernels, same order of operations), no workspace argument, and no
// cudaMalloc / cudaFree / allocating library call anywhere below.
//   - limits: 0 < B <= 2048, B*d <= 524288, d % 4 == 0; otherwise GDR_E_SHAPE and nothing is launched.
//   - the first call on each device looks up the device address of the scratch array (cudaGetSymbolAddress, which may
//     also make the runtime load this module on that device).  Make that first call during set-up, outside any
//     stream capture, with the device you will run on current; later calls on that device only launch kernels.
//   - all calls on one device share the one array.  Two calls that can overlap on the GPU (different streams without
//     an event between them, or two host threads) overwrite each other's intermediate data: serialise them.
//   - the array still holds whatever the previous call left there.  Nothing in it has to start at zero: every byte
//     this call reads (E rows, the padded 2·dG rows, one loss partial per CTA) is written earlier in the same call.
int gdr_gss_loss_fwd_bwd_static_f32(const float *embs, const int64_t *index, int64_t b, int64_t n, int64_t d,
                                    float alpha, float beta, float *loss_out, float *grad_e_out,
                                    gdr_stream_t stream) {
  GDR_CHECK_ARG(b > 0 && n > 0 && d > 0, GDR_E_BADARG, "gss_loss_static: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "gss_loss_static: d must be a multiple of 4");
  GDR_CHECK_ARG(b <= GSS_STATIC_MAX_B && b * d <= GSS_STATIC_MAX_BD, GDR_E_SHAPE,
                "gss_loss_static: B=%lld, d=%lld exceeds the built-in scratch (B <= %lld and B*d <= %lld); use "
                "gdr_gss_loss_fwd_bwd_f32 with a caller-owned workspace",
                (long long)b, (long long)d, (long long)GSS_STATIC_MAX_B, (long long)GSS_STATIC_MAX_BD);
  GDR_CHECK_ARG(embs && loss_out && grad_e_out, GDR_E_BADARG, "gss_loss_static: null pointer");
  unsigned char *scratch = nullptr;
  if (int rc = gss_static_scratch(&scratch)) return rc;
  cudaStream_t st = as_stream(stream);
  const int64_t ldw = align_up(b, 4);
  const int64_t tiles = ceil_div<int64_t>(b, GBM);
  float *e = reinterpret_cast<float *>(scratch);
  float *wm = reinterpret_cast<float *>(scratch + GSS_STATIC_E_BYTES);
  double *partials = reinterpret_cast<double *>(scratch + GSS_STATIC_E_BYTES + GSS_STATIC_W_BYTES);

  gather_rows_kernel<<<(unsigned)ceil_div<int64_t>(b * 32, 256), 256, 0, st>>>(embs, index, b, n, d, e);
  GDR_LAUNCHED();

  GemmParams g{};  // G = E·Eᵀ; the epilogue leaves one loss partial per CTA and the 2·dG tile
  g.a[0] = e; g.lda[0] = d; g.b[0] = e; g.ldb[0] = d; g.kseg[0] = d; g.nseg = 1;
  g.m = b; g.n = b; g.c = wm; g.ldc = ldw;
  g.alpha = alpha; g.beta = beta;
  g.wscale = (float)(2.0 / ((double)b * (double)b));
  g.loss_partials = partials;
  dim3 grid((unsigned)tiles, (unsigned)tiles, 1);
  gemm_f32_kernel<true, true, EPI_GSS><<<grid, GTHREADS, 0, st>>>(g);
  GDR_LAUNCHED();

  loss_finalize_kernel<<<1, 256, 0, st>>>(partials, (int)(tiles * tiles), 1.0 / ((double)b * (double)b), loss_out);
  GDR_LAUNCHED();

  GemmParams q{};  // dE = 2·dG · E
  q.a[0] = wm; q.lda[0] = ldw;
  q.b[0] = e; q.ldb[0] = d;
  q.kseg[0] = b; q.nseg = 1;
  q.m = b; q.n = d; q.c = grad_e_out; q.ldc = d;
  dim3 grid2((unsigned)ceil_div<int64_t>(d, GBN), (unsigned)tiles, 1);
  gemm_f32_kernel<true, false, EPI_STORE><<<grid2, GTHREADS, 0, st>>>(q);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
