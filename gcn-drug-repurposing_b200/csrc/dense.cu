// dense.cu — the dense half of GSS_GNNLayer / ResidualGraphConvolutionalNetwork:
// fused layer transform (two Linear + bias + ELU + residual), row normalisation and the
// backward pieces autograd derives for them.  Reference: modules/model.py:163-175,197-207.
#include "gemm_tc.cuh"

namespace gdr {

// ---- F.normalize(x, dim=1): one warp per row ----------------------------------------------
__global__ void __launch_bounds__(256)
row_normalize_fwd_kernel(const float *__restrict__ x, int64_t n, int64_t d, float *__restrict__ h,
                         float *__restrict__ norm) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n) return;
  const float *xr = x + row * d;
  float ss = 0.f;
  for (int64_t c = lane * 4; c < d; c += 128) {
    const float4 v = ldg_f4(xr + c);
    ss = fmaf(v.x, v.x, ss);
    ss = fmaf(v.y, v.y, ss);
    ss = fmaf(v.z, v.z, ss);
    ss = fmaf(v.w, v.w, ss);
  }
  ss = warp_sum(ss);
  const float nr = sqrtf(ss);
  const float den = fmaxf(nr, 1e-12f);
  if (lane == 0) norm[row] = nr;
  float *hr = h + row * d;
  for (int64_t c = lane * 4; c < d; c += 128) {
    const float4 v = ldg_f4(xr + c);
    *reinterpret_cast<float4 *>(hr + c) = make_float4(v.x / den, v.y / den, v.z / den, v.w / den);
  }
}

// gx = (gh - h <h,gh>) / max(norm, eps)     (norm > eps case of torch's normalize backward;
// for norm <= eps torch gives gh/eps, reproduced below)
__global__ void __launch_bounds__(256)
row_normalize_bwd_kernel(const float *__restrict__ gh, const float *__restrict__ h,
                         const float *__restrict__ norm, int64_t n, int64_t d, float *__restrict__ gx) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n) return;
  const float *gr = gh + row * d, *hr = h + row * d;
  float dot = 0.f;
  for (int64_t c = lane * 4; c < d; c += 128) {
    const float4 a = ldg_f4(gr + c), b = ldg_f4(hr + c);
    dot = fmaf(a.x, b.x, dot);
    dot = fmaf(a.y, b.y, dot);
    dot = fmaf(a.z, b.z, dot);
    dot = fmaf(a.w, b.w, dot);
  }
  dot = warp_sum(dot);
  const float nr = norm[row];
  const bool clamped = nr <= 1e-12f;
  const float inv = 1.f / fmaxf(nr, 1e-12f);
  if (clamped) dot = 0.f;
  float *o = gx + row * d;
  for (int64_t c = lane * 4; c < d; c += 128) {
    const float4 a = ldg_f4(gr + c), b = ldg_f4(hr + c);
    *reinterpret_cast<float4 *>(o + c) = make_float4((a.x - b.x * dot) * inv, (a.y - b.y * dot) * inv,
                                                     (a.z - b.z * dot) * inv, (a.w - b.w * dot) * inv);
  }
}

// gZ = scale * gXnext * ELU'(Z) (+ gres)
__global__ void __launch_bounds__(256)
elu_bwd_kernel(const float *__restrict__ gxn, const float *__restrict__ z, const float *__restrict__ gres,
               float scale, int64_t count4, float *__restrict__ gz) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count4) return;
  const float4 g = ldg_f4(gxn + i * 4), zz = ldg_f4(z + i * 4);
  float4 o;
  o.x = scale * g.x * (zz.x > 0.f ? 1.f : expf(zz.x));
  o.y = scale * g.y * (zz.y > 0.f ? 1.f : expf(zz.y));
  o.z = scale * g.z * (zz.z > 0.f ? 1.f : expf(zz.z));
  o.w = scale * g.w * (zz.w > 0.f ? 1.f : expf(zz.w));
  if (gres) {
    const float4 r = ldg_f4(gres + i * 4);
    o.x += r.x;
    o.y += r.y;
    o.z += r.z;
    o.w += r.w;
  }
  *reinterpret_cast<float4 *>(gz + i * 4) = o;
}

// split-K partial reduce: out[i] (+)= sum_z part[z][i].  A block owns 32 consecutive outputs; its 8 warps
// stride over z (8 partial rows in flight, 128-byte coalesced each) and are combined in a fixed order in
// shared memory, so the result is deterministic.
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float *__restrict__ part, int splits, int64_t count, int accumulate,
                     float *__restrict__ out) {
  __shared__ float red[8][32];
  const int o = threadIdx.x & 31, zl = threadIdx.x >> 5;
  const int64_t i = (int64_t)blockIdx.x * 32 + o;
  float s = 0.f;
  if (i < count)
    for (int z = zl; z < splits; z += 8) s += part[(int64_t)z * count + i];
  red[zl][o] = s;
  __syncthreads();
  if (zl == 0 && i < count) {
    const float t = ((red[0][o] + red[1][o]) + (red[2][o] + red[3][o])) + ((red[4][o] + red[5][o]) + (red[6][o] + red[7][o]));
    out[i] = accumulate ? out[i] + t : t;
  }
}

// the same for partials of a [d, 2d] product whose left / right halves are two separate [d, d] outputs (dW1 | dW2);
// out2 == nullptr: only the left half exists (width d).  `dup` != 0: out2 receives the SAME sum as out1 (db1 = db2).
__global__ void __launch_bounds__(256)
splitk_reduce2_kernel(const float *__restrict__ part, int splits, int64_t rows, int64_t d, int accumulate, int dup,
                      float *__restrict__ out1, float *__restrict__ out2) {
  __shared__ float red[8][32];
  const int o = threadIdx.x & 31, zl = threadIdx.x >> 5;
  const int64_t width = dup ? d : 2 * d;
  const int64_t count = rows * width;
  const int64_t i = (int64_t)blockIdx.x * 32 + o;
  float s = 0.f;
  if (i < count)
    for (int z = zl; z < splits; z += 8) s += part[(int64_t)z * count + i];
  red[zl][o] = s;
  __syncthreads();
  if (zl == 0 && i < count) {
    const float t = ((red[0][o] + red[1][o]) + (red[2][o] + red[3][o])) + ((red[4][o] + red[5][o]) + (red[6][o] + red[7][o]));
    if (dup) {
      out1[i] = accumulate ? out1[i] + t : t;
      out2[i] = accumulate ? out2[i] + t : t;
    } else {
      const int64_t m = i / width, j = i - m * width;
      float *dst = j < d ? out1 + m * d + j : out2 + m * d + (j - d);
      *dst = accumulate ? *dst + t : t;
    }
  }
}

// column sums of G[n,d] -> partial[blocks][d]; each block owns a slab of CS_ROWS rows.  256 threads =
// 64 float4 column groups x 4 row lanes (4 rows in flight per column group), fixed-order combine in
// shared memory: deterministic.
constexpr int CS_ROWS = 128;
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const float *__restrict__ g, int64_t n, int64_t d, float *__restrict__ part) {
  __shared__ float4 red[4][64];
  const int64_t r0 = (int64_t)blockIdx.x * CS_ROWS;
  const int64_t r1 = min(n, r0 + CS_ROWS);
  const int cg = threadIdx.x & 63, rl = threadIdx.x >> 6;
  for (int64_t c = (int64_t)cg * 4; c < d; c += 256) {  // d % 4 == 0
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t r = r0 + rl; r < r1; r += 4) {
      const float4 v = ldg_f4(g + r * d + c);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    red[rl][cg] = s;
    __syncthreads();
    if (rl == 0) {
      const float4 a = red[0][cg], b = red[1][cg], e = red[2][cg], f = red[3][cg];
      *reinterpret_cast<float4 *>(part + (int64_t)blockIdx.x * d + c) =
          make_float4((a.x + b.x) + (e.x + f.x), (a.y + b.y) + (e.y + f.y), (a.z + b.z) + (e.z + f.z),
                      (a.w + b.w) + (e.w + f.w));
    }
    __syncthreads();
  }
}

static int splits_for(int64_t n) {
  // dW is d x d (a handful of tiles): split the node dimension so that ~2 waves of CTAs exist
  int64_t s = ceil_div<int64_t>(n, 512);
  if (s < 1) s = 1;
  if (s > 2 * kNumSMs) s = 2 * kNumSMs;
  return (int)s;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int gdr_gcn_layer_fwd_f32(const float *p, const float *t, const float *w1, const float *b1, const float *w2,
                          const float *b2, const float *residual, float decay, int64_t n, int64_t d,
                          float *z_out, float *xnext_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0, GDR_E_BADARG, "gcn_layer_fwd: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "gcn_layer_fwd: d=%lld must be a multiple of 4", (long long)d);
  GDR_CHECK_ARG(p && t && w1 && b1 && w2 && b2 && z_out && xnext_out, GDR_E_BADARG, "gcn_layer_fwd: null pointer");
  if (n == 0) return GDR_OK;
  GemmParams g{};
  g.a[0] = p; g.b[0] = w1; g.lda[0] = d; g.ldb[0] = d; g.kseg[0] = d;
  g.a[1] = t; g.b[1] = w2; g.lda[1] = d; g.ldb[1] = d; g.kseg[1] = d;
  g.nseg = 2;
  g.m = n; g.n = d;
  g.c = z_out; g.ldc = d;
  g.bias1 = b1; g.bias2 = b2; g.residual = residual; g.decay = decay; g.xnext = xnext_out;
  dim3 grid((unsigned)ceil_div<int64_t>(d, GBN), (unsigned)ceil_div<int64_t>(n, GBM), 1);
  gemm_f32_kernel<true, true, EPI_LAYER><<<grid, GTHREADS, 0, as_stream(stream)>>>(g);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_row_normalize_fwd_f32(const float *x, int64_t n, int64_t d, float *h_out, float *norm_out,
                              gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && x && h_out && norm_out, GDR_E_BADARG, "row_normalize_fwd: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "row_normalize_fwd: d must be a multiple of 4");
  if (n == 0) return GDR_OK;
  row_normalize_fwd_kernel<<<(unsigned)ceil_div<int64_t>(n * 32, 256), 256, 0, as_stream(stream)>>>(x, n, d, h_out,
                                                                                                   norm_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_row_normalize_bwd_f32(const float *gh, const float *h, const float *norm, int64_t n, int64_t d,
                              float *gx_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && gh && h && norm && gx_out, GDR_E_BADARG, "row_normalize_bwd: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "row_normalize_bwd: d must be a multiple of 4");
  if (n == 0) return GDR_OK;
  row_normalize_bwd_kernel<<<(unsigned)ceil_div<int64_t>(n * 32, 256), 256, 0, as_stream(stream)>>>(gh, h, norm, n, d,
                                                                                                   gx_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_elu_bwd_f32(const float *gxnext, const float *z, const float *gres, float scale, int64_t count,
                    float *gz_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(count >= 0 && gxnext && z && gz_out, GDR_E_BADARG, "elu_bwd: bad args");
  GDR_CHECK_ARG(count % 4 == 0, GDR_E_SHAPE, "elu_bwd: count must be a multiple of 4");
  if (count == 0) return GDR_OK;
  elu_bwd_kernel<<<(unsigned)ceil_div<int64_t>(count / 4, 256), 256, 0, as_stream(stream)>>>(gxnext, z, gres, scale,
                                                                                            count / 4, gz_out);
  GDR_LAUNCHED();
  return GDR_OK;
}

int gdr_linear_bwd_input_f32(const float *g, const float *w, const float *e1, const float *e2, int64_t n,
                             int64_t d, float *c_out, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && g && w && c_out, GDR_E_BADARG, "linear_bwd_input: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "linear_bwd_input: d must be a multiple of 4");
  GDR_CHECK_ARG((e1 == nullptr) == (e2 == nullptr), GDR_E_BADARG, "linear_bwd_input: e1/e2 must come together");
  if (n == 0) return GDR_OK;
  GemmParams q{};
  q.a[0] = g; q.lda[0] = d;   // A(m=node, k=out) = G[m*d + k]        (K-major)
  q.b[0] = w; q.ldb[0] = d;   // B(n=in,  k=out) = W[k*d + n]         (not K-major)
  q.kseg[0] = d; q.nseg = 1;
  q.m = n; q.n = d; q.c = c_out; q.ldc = d; q.e1 = e1; q.e2 = e2;
  dim3 grid((unsigned)ceil_div<int64_t>(d, GBN), (unsigned)ceil_div<int64_t>(n, GBM), 1);
  gemm_f32_kernel<true, false, EPI_STORE><<<grid, GTHREADS, 0, as_stream(stream)>>>(q);
  GDR_LAUNCHED();
  return GDR_OK;
}

int64_t gdr_linear_bwd_weight_workspace_bytes(int64_t n, int64_t d) {
  if (n < 0 || d <= 0) return GDR_E_BADARG;
  const int64_t splits = splits_for(n);
  const int64_t cs_blocks = ceil_div<int64_t>(n > 0 ? n : 1, CS_ROWS);
  return align_up(splits * d * d * 4, 256) + align_up(cs_blocks * d * 4, 256) + 512;
}

int gdr_linear_bwd_weight_f32(const float *g, const float *a, int64_t n, int64_t d, int accumulate, float *dw,
                              float *db, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && g && a && dw && workspace, GDR_E_BADARG, "linear_bwd_weight: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "linear_bwd_weight: d must be a multiple of 4");
  GDR_CHECK_ARG(workspace_bytes >= gdr_linear_bwd_weight_workspace_bytes(n, d), GDR_E_WORKSPACE,
                "linear_bwd_weight: workspace too small");
  cudaStream_t st = as_stream(stream);
  const int splits = splits_for(n);
  Carver cv(workspace);
  float *part = cv.take<float>((int64_t)splits * d * d);
  const int64_t cs_blocks = ceil_div<int64_t>(n > 0 ? n : 1, CS_ROWS);
  float *cpart = cv.take<float>(cs_blocks * d);
  if (n == 0) {
    if (!accumulate) {
      GDR_CUDA(cudaMemsetAsync(dw, 0, d * d * 4, st));
      if (db) GDR_CUDA(cudaMemsetAsync(db, 0, d * 4, st));
    }
    return GDR_OK;
  }
  GemmParams q{};
  q.a[0] = g; q.lda[0] = d;  // A(m=out, k=node) = G[k*d + m]
  q.b[0] = a; q.ldb[0] = d;  // B(n=in,  k=node) = A[k*d + n]
  q.kseg[0] = n; q.nseg = 1;
  q.m = d; q.n = d; q.c = part; q.ldc = d;
  q.k_per_split = ceil_div<int64_t>(ceil_div<int64_t>(n, splits), GBK) * GBK;
  dim3 grid((unsigned)ceil_div<int64_t>(d, GBN), (unsigned)ceil_div<int64_t>(d, GBM), (unsigned)splits);
  gemm_f32_kernel<false, false, EPI_SPLITK><<<grid, GTHREADS, 0, st>>>(q);
  GDR_LAUNCHED();
  splitk_reduce_kernel<<<(unsigned)ceil_div<int64_t>(d * d, 32), 256, 0, st>>>(part, splits, d * d, accumulate, dw);
  GDR_LAUNCHED();
  if (db) {
    colsum_partial_kernel<<<(unsigned)cs_blocks, 256, 0, st>>>(g, n, d, cpart);
    GDR_LAUNCHED();
    splitk_reduce_kernel<<<(unsigned)ceil_div<int64_t>(d, 32), 256, 0, st>>>(cpart, (int)cs_blocks, d, accumulate,
                                                                             db);
    GDR_LAUNCHED();
  }
  return GDR_OK;
}

// ---- tcgen05 (3xTF32) variants of the three contractions: same arithmetic contract, operands split into
// TF32 hi/lo pairs in a caller-owned workspace first (gemm_tc.cu) ------------------------------------------

int64_t gdr_gcn_layer_fwd_tc_workspace_bytes(int64_t n, int64_t d) {
  if (n < 0 || d <= 0) return GDR_E_BADARG;
  return 2 * align_up(n * 2 * d * 4, 256) + 2 * align_up(d * 2 * d * 4, 256) + 1024;
}

int gdr_gcn_layer_fwd_tc_f32(const float *p, const float *t, const float *w1, const float *b1, const float *w2,
                             const float *b2, const float *residual, float decay, int64_t n, int64_t d, float *z_out,
                             float *xnext_out, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0, GDR_E_BADARG, "gcn_layer_fwd_tc: bad sizes");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "gcn_layer_fwd_tc: d=%lld must be a multiple of 4", (long long)d);
  GDR_CHECK_ARG(p && t && w1 && b1 && w2 && b2 && z_out && xnext_out && workspace, GDR_E_BADARG,
                "gcn_layer_fwd_tc: null pointer");
  GDR_CHECK_ARG(workspace_bytes >= gdr_gcn_layer_fwd_tc_workspace_bytes(n, d), GDR_E_WORKSPACE,
                "gcn_layer_fwd_tc: workspace too small");
  if (n == 0) return GDR_OK;
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  float *ahi = cv.take<float>(n * 2 * d), *alo = cv.take<float>(n * 2 * d);
  float *bhi = cv.take<float>(d * 2 * d), *blo = cv.take<float>(d * 2 * d);
  int rc;
  if ((rc = tc_split(w1, d, d, w2, d, d, d, bhi, blo, st))) return rc;    // B = [W1 | W2] (d x 2d)
  GemmParams g{};
  g.c = z_out; g.ldc = d;
  g.bias1 = b1; g.bias2 = b2; g.residual = residual; g.decay = decay; g.xnext = xnext_out;
  if (d % TC_K_ALIGN == 0)                                                 // A = [P | T], split inside the kernel
    return tc_gemm_raw_a(EPI_LAYER, p, d, d, t, d, d, bhi, blo, 2 * d, n, d, g, st);
  if ((rc = tc_split(p, d, d, t, d, d, n, ahi, alo, st))) return rc;      // odd widths: pre-split [P | T] (n x 2d)
  return tc_gemm(EPI_LAYER, ahi, alo, bhi, blo, n, d, 2 * d, 2 * d, 2 * d, 2 * d, 1, g, st);
}

int64_t gdr_linear_bwd_input_tc_workspace_bytes(int64_t n, int64_t d) {
  if (n < 0 || d <= 0) return GDR_E_BADARG;
  return 2 * align_up(n * d * 4, 256) + 2 * align_up(d * d * 4, 256) + 1024;
}

int gdr_linear_bwd_input_tc_f32(const float *g, const float *w, const float *e1, const float *e2, int64_t n, int64_t d,
                                float *c_out, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && g && w && c_out && workspace, GDR_E_BADARG, "linear_bwd_input_tc: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "linear_bwd_input_tc: d must be a multiple of 4");
  GDR_CHECK_ARG((e1 == nullptr) == (e2 == nullptr), GDR_E_BADARG, "linear_bwd_input_tc: e1/e2 must come together");
  GDR_CHECK_ARG(workspace_bytes >= gdr_linear_bwd_input_tc_workspace_bytes(n, d), GDR_E_WORKSPACE,
                "linear_bwd_input_tc: workspace too small");
  if (n == 0) return GDR_OK;
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  float *ahi = cv.take<float>(n * d), *alo = cv.take<float>(n * d);
  float *bhi = cv.take<float>(d * d), *blo = cv.take<float>(d * d);
  int rc;
  (void)ahi; (void)alo;                                                      // A(m=node, k=out) = G, split in-kernel
  if ((rc = tc_split_transpose(w, d, d, d, d, bhi, blo, st))) return rc;     // B(n=in, k=out)  = Wᵀ
  GemmParams q{};
  q.c = c_out; q.ldc = d; q.e1 = e1; q.e2 = e2;
  return tc_gemm_raw_a(EPI_STORE, g, d, d, nullptr, 0, 0, bhi, blo, d, n, d, q, st);
}

// TN kernel (gemm_tc_tn_kernel): one work item per SM — K ranges so that tiles * splits <= 148
static int tn_splits_for(int64_t n, int64_t tiles, int64_t *k_per_split) {
  int64_t s = kNumSMs / tiles;
  if (s < 1) s = 1;
  const int64_t per = align_up(ceil_div<int64_t>(n > 0 ? n : 1, s), TC_K_ALIGN);
  *k_per_split = per;
  return (int)ceil_div<int64_t>(n > 0 ? n : 1, per);
}

static int tc_splits_for(int64_t n, int64_t d, int64_t *k_per_split) {
  const int64_t tiles = ceil_div<int64_t>(d, 128) * ceil_div<int64_t>(d, 128);
  int64_t s = ceil_div<int64_t>(2 * kNumSMs, tiles);                  // ~2 CTAs per SM in total
  int64_t per = align_up(ceil_div<int64_t>(n > 0 ? n : 1, s), TC_K_ALIGN);
  if (per < 4 * TC_K_ALIGN) per = 4 * TC_K_ALIGN;                     // at least 4 k-blocks per CTA
  *k_per_split = per;
  return (int)ceil_div<int64_t>(n > 0 ? n : 1, per);
}

int64_t gdr_linear_bwd_weight_tc_workspace_bytes(int64_t n, int64_t d) {
  if (n < 0 || d <= 0) return GDR_E_BADARG;
  int64_t per;
  const int64_t splits = tc_splits_for(n, d, &per);
  const int64_t ldt = align_up(n > 0 ? n : 1, 4);
  const int64_t cs_blocks = ceil_div<int64_t>(n > 0 ? n : 1, CS_ROWS);
  int64_t tn_per;
  const int64_t tn = tn_splits_for(n, ceil_div<int64_t>(d, 128) * ceil_div<int64_t>(d, 128), &tn_per);
  return 4 * align_up(d * ldt * 4, 256) + align_up(std::max(splits, tn) * d * d * 4, 256) +
         align_up(std::max(cs_blocks, tn) * d * 4, 256) + 1024;
}

int gdr_linear_bwd_weight_tc_f32(const float *g, const float *a, int64_t n, int64_t d, int accumulate, float *dw,
                                 float *db, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && g && a && dw && workspace, GDR_E_BADARG, "linear_bwd_weight_tc: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "linear_bwd_weight_tc: d must be a multiple of 4");
  GDR_CHECK_ARG(workspace_bytes >= gdr_linear_bwd_weight_tc_workspace_bytes(n, d), GDR_E_WORKSPACE,
                "linear_bwd_weight_tc: workspace too small");
  cudaStream_t st = as_stream(stream);
  if (n == 0) {
    if (!accumulate) {
      GDR_CUDA(cudaMemsetAsync(dw, 0, d * d * 4, st));
      if (db) GDR_CUDA(cudaMemsetAsync(db, 0, d * 4, st));
    }
    return GDR_OK;
  }
  // raw operands straight into the TN kernel: no transposed copies, column sums from the A converter
  int64_t per;
  int rc;
  const int tn = tn_splits_for(n, ceil_div<int64_t>(d, 128) * ceil_div<int64_t>(d, 128), &per);
  Carver cv(workspace);
  float *part = cv.take<float>((int64_t)tn * d * d);
  float *cpart = cv.take<float>((int64_t)tn * d);
  if ((rc = tc_gemm_tn_splitk(g, d, d, a, d, d, nullptr, 0, 0, n, per, tn, part, db ? cpart : nullptr, st))) return rc;
  splitk_reduce_kernel<<<(unsigned)ceil_div<int64_t>(d * d, 32), 256, 0, st>>>(part, tn, d * d, accumulate, dw);
  GDR_LAUNCHED();
  if (db) {
    splitk_reduce_kernel<<<(unsigned)ceil_div<int64_t>(d, 32), 256, 0, st>>>(cpart, tn, d, accumulate, db);
    GDR_LAUNCHED();
  }
  return GDR_OK;
}

// Both weight gradients of a layer from ONE split-K product: [dW1 | dW2] = Gᵀ·[A1 | A2] (G is transposed and split
// once, N = 2d gives twice the tiles per K range), and db1 = db2 = colsum(G) from one pass (Z = Z1 + Z2, so both
// biases see the same gradient).  modules/model.py:165,170 under autograd.
int64_t gdr_linear_bwd_weight2_tc_workspace_bytes(int64_t n, int64_t d) {
  if (n < 0 || d <= 0) return GDR_E_BADARG;
  const int64_t tiles = ceil_div<int64_t>(d, 128) * ceil_div<int64_t>(2 * d, 128);
  int64_t s = ceil_div<int64_t>(2 * kNumSMs, tiles);
  int64_t per = align_up(ceil_div<int64_t>(n > 0 ? n : 1, s), TC_K_ALIGN);
  if (per < 4 * TC_K_ALIGN) per = 4 * TC_K_ALIGN;
  const int64_t splits = ceil_div<int64_t>(n > 0 ? n : 1, per);
  const int64_t ldt = align_up(n > 0 ? n : 1, 4);
  const int64_t cs_blocks = ceil_div<int64_t>(n > 0 ? n : 1, CS_ROWS);
  int64_t tn_per;
  const int64_t tn = tn_splits_for(n, tiles, &tn_per);
  return 2 * align_up(d * ldt * 4, 256) + 2 * align_up(2 * d * ldt * 4, 256) +
         align_up(std::max(splits, tn) * d * 2 * d * 4, 256) + align_up(std::max(cs_blocks, tn) * d * 4, 256) + 1024;
}

int gdr_linear_bwd_weight2_tc_f32(const float *g, const float *a1, const float *a2, int64_t n, int64_t d,
                                  int accumulate, float *dw1, float *db1, float *dw2, float *db2, void *workspace,
                                  int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n >= 0 && d > 0 && g && a1 && a2 && dw1 && dw2 && db1 && db2 && workspace, GDR_E_BADARG,
                "linear_bwd_weight2_tc: bad args");
  GDR_CHECK_ARG(d % 4 == 0, GDR_E_SHAPE, "linear_bwd_weight2_tc: d must be a multiple of 4");
  GDR_CHECK_ARG(workspace_bytes >= gdr_linear_bwd_weight2_tc_workspace_bytes(n, d), GDR_E_WORKSPACE,
                "linear_bwd_weight2_tc: workspace too small");
  cudaStream_t st = as_stream(stream);
  if (n == 0) {
    if (!accumulate) {
      GDR_CUDA(cudaMemsetAsync(dw1, 0, d * d * 4, st));
      GDR_CUDA(cudaMemsetAsync(dw2, 0, d * d * 4, st));
      GDR_CUDA(cudaMemsetAsync(db1, 0, d * 4, st));
      GDR_CUDA(cudaMemsetAsync(db2, 0, d * 4, st));
    }
    return GDR_OK;
  }
  const int64_t tiles = ceil_div<int64_t>(d, 128) * ceil_div<int64_t>(2 * d, 128);
  int rc;
  if (d % 128 == 0) {  // an N tile never straddles A1 | A2: raw operands straight into the TN kernel
    int64_t tn_per;
    const int tn = tn_splits_for(n, tiles, &tn_per);
    Carver cv(workspace);
    float *part = cv.take<float>((int64_t)tn * d * 2 * d);
    float *cpart = cv.take<float>((int64_t)tn * d);
    if ((rc = tc_gemm_tn_splitk(g, d, d, a1, d, d, a2, d, d, n, tn_per, tn, part, cpart, st))) return rc;
    splitk_reduce2_kernel<<<(unsigned)ceil_div<int64_t>(d * 2 * d, 32), 256, 0, st>>>(part, tn, d, d, accumulate, 0, dw1,
                                                                                   dw2);
    GDR_LAUNCHED();
    splitk_reduce2_kernel<<<(unsigned)ceil_div<int64_t>(d, 32), 256, 0, st>>>(cpart, tn, 1, d, accumulate, 1, db1, db2);
    GDR_LAUNCHED();
    return GDR_OK;
  }
  int64_t per = align_up(ceil_div<int64_t>(n, ceil_div<int64_t>(2 * kNumSMs, tiles)), TC_K_ALIGN);
  if (per < 4 * TC_K_ALIGN) per = 4 * TC_K_ALIGN;
  const int splits = (int)ceil_div<int64_t>(n, per);
  const int64_t ldt = align_up(n, 4);
  const int64_t cs_blocks = ceil_div<int64_t>(n, CS_ROWS);
  Carver cv(workspace);
  float *ghi = cv.take<float>(d * ldt), *glo = cv.take<float>(d * ldt);
  float *ahi = cv.take<float>(2 * d * ldt), *alo = cv.take<float>(2 * d * ldt);
  float *part = cv.take<float>((int64_t)splits * d * 2 * d);
  float *cpart = cv.take<float>(cs_blocks * d);
  if ((rc = tc_split_transpose(g, d, n, d, ldt, ghi, glo, st))) return rc;                         // A(m=out, k=node)
  if ((rc = tc_split_transpose(a1, d, n, d, ldt, ahi, alo, st))) return rc;                        // B rows 0..d-1
  if ((rc = tc_split_transpose(a2, d, n, d, ldt, ahi + d * ldt, alo + d * ldt, st))) return rc;    // B rows d..2d-1
  GemmParams q{};
  q.c = part; q.ldc = 2 * d;
  if ((rc = tc_gemm(EPI_SPLITK, ghi, glo, ahi, alo, d, 2 * d, n, ldt, ldt, per, splits, q, st))) return rc;
  splitk_reduce2_kernel<<<(unsigned)ceil_div<int64_t>(d * 2 * d, 32), 256, 0, st>>>(part, splits, d, d, accumulate, 0,
                                                                                 dw1, dw2);
  GDR_LAUNCHED();
  colsum_partial_kernel<<<(unsigned)cs_blocks, 256, 0, st>>>(g, n, d, cpart);
  GDR_LAUNCHED();
  splitk_reduce2_kernel<<<(unsigned)ceil_div<int64_t>(d, 32), 256, 0, st>>>(cpart, (int)cs_blocks, 1, d, accumulate, 1,
                                                                         db1, db2);
  GDR_LAUNCHED();
  return GDR_OK;
}

}  // extern "C"
