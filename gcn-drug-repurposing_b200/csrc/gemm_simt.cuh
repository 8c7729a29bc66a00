// gemm_simt.cuh — fp32 FFMA GEMM with fused epilogues (128x128x16 tiles, 8x8 per thread,
// register-prefetch double buffering).  Exact fp32 arithmetic: the reference tolerance
// (1e-5 relative) rules out single-pass TF32/BF16 tensor-core products for these
// contractions; the tcgen05 3xTF32 variant lives in gemm_tc.cu and replaces this kernel
// where it has been validated.
#pragma once
#include <string.h>

#include "common.cuh"

namespace gdr {

constexpr int GBM = 128, GBN = 128, GBK = 16, GTHREADS = 256, GPAD = 4;

enum GemmEpi { EPI_STORE = 0, EPI_LAYER = 1, EPI_SPLITK = 2, EPI_GSS = 3, EPI_HIST = 4, EPI_MINGT = 5 };

constexpr int GHIST_BINS = 2048;

// radix-select state of the Gram order-statistic search (gram.cu)
struct SelectState {
  unsigned int prefix;       // key bits fixed so far
  unsigned int mask;         // which bits those are
  unsigned long long rank;   // rank still to descend inside the prefix bucket
  unsigned long long count;  // population of the bucket chosen last
  unsigned int min_gt;       // EPI_MINGT: smallest key strictly above `prefix`
};

// order-preserving fp32 -> uint32 key
__device__ __forceinline__ unsigned int f32_key(float v) {
  const unsigned int b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float f32_from_key(unsigned int k) {
  const unsigned int b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

struct GemmParams {
  // up to two K-segments: C = sum_s A_s · B_sᵀ   (A_s(m,k), B_s(n,k))
  const float *a[2];
  const float *b[2];
  int64_t lda[2], ldb[2];
  int64_t kseg[2];
  int nseg;
  int64_t m, n;
  int64_t k_per_split;  // EPI_SPLITK: k range handled by blockIdx.z (segment 0 only)
  float *c;             // EPI_STORE / EPI_SPLITK (partials [z][m][n]) / EPI_LAYER: z_out / EPI_GSS: w_out
  int64_t ldc;
  // EPI_STORE addend: c += e1 ⊙ e2 (same layout as c)
  const float *e1, *e2;
  // EPI_LAYER
  const float *bias1, *bias2, *residual;
  float decay;
  float *xnext;
  // EPI_GSS
  float alpha, beta, wscale;
  double *loss_partials;  // one per CTA
  // EPI_HIST / EPI_MINGT (symmetric Gram: only tiles with blockIdx.x >= blockIdx.y run, off-diagonal
  // tiles count twice)
  const SelectState *sel;
  unsigned long long *hist;  // [GHIST_BINS]
  int shift;                 // digit position of this pass
  unsigned int bin_mask;     // (bins of this pass) - 1
  SelectState *sel_out;      // EPI_MINGT: atomicMin target
};

__device__ __forceinline__ float elu1(float z) { return z > 0.f ? z : expm1f(z); }

// load one 128 x 16 operand tile into registers (2 float4 per thread)
template <bool KMAJOR>
__device__ __forceinline__ void gemm_load_tile(const float *__restrict__ base, int64_t ld, int64_t rows,
                                               int64_t r0, int64_t k0, int64_t k_end, float4 (&reg)[2]) {
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int idx = t + i * GTHREADS;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (KMAJOR) {
      const int64_t r = r0 + (idx >> 2);
      const int64_t k = k0 + (idx & 3) * 4;
      if (r < rows && k < k_end) {
        const float *p = base + r * ld + k;
        if (k + 3 < k_end) {
          v = ldg_f4(p);
        } else {
          v.x = __ldg(p);
          if (k + 1 < k_end) v.y = __ldg(p + 1);
          if (k + 2 < k_end) v.z = __ldg(p + 2);
        }
      }
    } else {
      const int64_t k = k0 + (idx >> 5);
      const int64_t r = r0 + (idx & 31) * 4;
      if (k < k_end && r < rows) v = ldg_f4(base + k * ld + r);  // rows % 4 == 0 guaranteed by callers
    }
    reg[i] = v;
  }
}

template <bool KMAJOR>
__device__ __forceinline__ void gemm_store_tile(float (*sm)[GBM + GPAD], const float4 (&reg)[2]) {
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int idx = t + i * GTHREADS;
    if (KMAJOR) {
      const int r = idx >> 2, kq = (idx & 3) * 4;
      sm[kq + 0][r] = reg[i].x;
      sm[kq + 1][r] = reg[i].y;
      sm[kq + 2][r] = reg[i].z;
      sm[kq + 3][r] = reg[i].w;
    } else {
      const int k = idx >> 5, r = (idx & 31) * 4;
      *reinterpret_cast<float4 *>(&sm[k][r]) = reg[i];
    }
  }
}

template <bool A_KMAJOR, bool B_KMAJOR, int EPI>
__global__ void __launch_bounds__(GTHREADS) gemm_f32_kernel(const GemmParams p) {
  __shared__ __align__(16) float As[2][GBK][GBM + GPAD];
  __shared__ __align__(16) float Bs[2][GBK][GBN + GPAD];

  if (EPI == EPI_HIST || EPI == EPI_MINGT) {
    if (blockIdx.x < blockIdx.y) return;  // symmetric: the mirrored tile is counted by weight 2
  }
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int64_t m0 = (int64_t)blockIdx.y * GBM;
  const int64_t n0 = (int64_t)blockIdx.x * GBN;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  for (int s = 0; s < p.nseg; ++s) {
    const float *A = p.a[s];
    const float *B = p.b[s];
    const int64_t lda = p.lda[s], ldb = p.ldb[s];
    int64_t k_begin = 0, k_end = p.kseg[s];
    if (EPI == EPI_SPLITK) {
      k_begin = (int64_t)blockIdx.z * p.k_per_split;
      k_end = min(k_end, k_begin + p.k_per_split);
    }
    if (k_begin >= k_end) continue;

    float4 ra[2], rb[2];
    gemm_load_tile<A_KMAJOR>(A, lda, p.m, m0, k_begin, k_end, ra);
    gemm_load_tile<B_KMAJOR>(B, ldb, p.n, n0, k_begin, k_end, rb);
    __syncthreads();  // previous segment's readers are done with buffer 0
    gemm_store_tile<A_KMAJOR>(As[0], ra);
    gemm_store_tile<B_KMAJOR>(Bs[0], rb);
    __syncthreads();

    int buf = 0;
    for (int64_t k0 = k_begin; k0 < k_end; k0 += GBK) {
      const bool has_next = (k0 + GBK) < k_end;
      if (has_next) {
        gemm_load_tile<A_KMAJOR>(A, lda, p.m, m0, k0 + GBK, k_end, ra);
        gemm_load_tile<B_KMAJOR>(B, ldb, p.n, n0, k0 + GBK, k_end, rb);
      }
#pragma unroll
      for (int kk = 0; kk < GBK; ++kk) {
        const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][kk][ty * 4]);
        const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][kk][64 + ty * 4]);
        const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][tx * 4]);
        const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][64 + tx * 4]);
        const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
      if (has_next) {
        gemm_store_tile<A_KMAJOR>(As[buf ^ 1], ra);
        gemm_store_tile<B_KMAJOR>(Bs[buf ^ 1], rb);
        __syncthreads();
        buf ^= 1;
      }
    }
  }

  // ---- epilogue ----------------------------------------------------------------------
  if (EPI == EPI_HIST) {
    __shared__ unsigned int sh_hist[GHIST_BINS];
    for (int i = t; i < GHIST_BINS; i += GTHREADS) sh_hist[i] = 0u;
    __syncthreads();
    const unsigned int prefix = p.sel->prefix, mask = p.sel->mask;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
      if (m >= p.m) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int64_t n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
        if (n >= p.n) continue;
        const unsigned int key = f32_key(acc[i][j]);
        if ((key & mask) == prefix) atomicAdd(&sh_hist[(key >> p.shift) & p.bin_mask], 1u);
      }
    }
    __syncthreads();
    const unsigned long long w = (blockIdx.x == blockIdx.y) ? 1ull : 2ull;
    for (int i = t; i < GHIST_BINS; i += GTHREADS) {
      const unsigned int c = sh_hist[i];
      if (c) atomicAdd(p.hist + i, w * c);
    }
    return;
  }
  if (EPI == EPI_MINGT) {
    const unsigned int base = p.sel->prefix;
    unsigned int best = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
      if (m >= p.m) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int64_t n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
        if (n >= p.n) continue;
        const unsigned int key = f32_key(acc[i][j]);
        if (key > base) best = min(best, key);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((t & 31) == 0 && best != 0xffffffffu) atomicMin(&p.sel_out->min_gt, best);
    return;
  }
  double lsum = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= p.m) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int64_t n = n0 + jh * 64 + tx * 4;
      if (n >= p.n) continue;  // n % 4 == 0 for every caller except EPI_GSS (handled per element)
      float v[4] = {acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]};
      if (EPI == EPI_STORE) {
        float *cp = p.c + m * p.ldc + n;
        if (p.e1) {
          const float4 a = ldg_f4(p.e1 + m * p.ldc + n), b = ldg_f4(p.e2 + m * p.ldc + n);
          v[0] = fmaf(a.x, b.x, v[0]);
          v[1] = fmaf(a.y, b.y, v[1]);
          v[2] = fmaf(a.z, b.z, v[2]);
          v[3] = fmaf(a.w, b.w, v[3]);
        }
        *reinterpret_cast<float4 *>(cp) = make_float4(v[0], v[1], v[2], v[3]);
      } else if (EPI == EPI_SPLITK) {
        float *cp = p.c + ((int64_t)blockIdx.z * p.m + m) * p.ldc + n;
        *reinterpret_cast<float4 *>(cp) = make_float4(v[0], v[1], v[2], v[3]);
      } else if (EPI == EPI_LAYER) {
        const float4 b1 = ldg_f4(p.bias1 + n), b2 = ldg_f4(p.bias2 + n);
        // reference order: (P·W1ᵀ + b1) + (T·W2ᵀ + b2); the two products share one accumulator here
        v[0] += b1.x + b2.x;
        v[1] += b1.y + b2.y;
        v[2] += b1.z + b2.z;
        v[3] += b1.w + b2.w;
        *reinterpret_cast<float4 *>(p.c + m * p.ldc + n) = make_float4(v[0], v[1], v[2], v[3]);
        float y[4] = {elu1(v[0]), elu1(v[1]), elu1(v[2]), elu1(v[3])};
        if (p.residual) {
          const float4 r = ldg_f4(p.residual + m * p.ldc + n);
          y[0] = fmaf(p.decay, y[0], r.x);
          y[1] = fmaf(p.decay, y[1], r.y);
          y[2] = fmaf(p.decay, y[2], r.z);
          y[3] = fmaf(p.decay, y[3], r.w);
        }
        *reinterpret_cast<float4 *>(p.xnext + m * p.ldc + n) = make_float4(y[0], y[1], y[2], y[3]);
      } else {  // EPI_GSS: p.n need not be a multiple of 4; p.ldc is (padded to 4)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float w = 0.f;
          if (n + q < p.n) {
            const float g = v[q];
            const float r = fmaxf(g, 0.f) - p.beta;
            lsum += (double)(-0.5f * p.alpha * r * r);
            w = g > 0.f ? -p.alpha * r * p.wscale : 0.f;
          }
          v[q] = w;
        }
        *reinterpret_cast<float4 *>(p.c + m * p.ldc + n) = make_float4(v[0], v[1], v[2], v[3]);
      }
    }
  }
  if (EPI == EPI_GSS) {
    __shared__ double red[GTHREADS / 32];
    lsum = warp_sum(lsum);
    if ((t & 31) == 0) red[t >> 5] = lsum;
    __syncthreads();
    if (t == 0) {
      double s = 0.0;
      for (int w = 0; w < GTHREADS / 32; ++w) s += red[w];
      p.loss_partials[(int64_t)blockIdx.y * gridDim.x + blockIdx.x] = s;
    }
  }
}

}  // namespace gdr
