// gemm_tc.cu — fp32-accurate NT GEMM on the 5th-generation tensor cores (tcgen05, sm_100a).
//
//   C[M,N] = A[M,K] · B[N,K]ᵀ   A, B row-major with K contiguous ("K-major"), fp32 in / fp32 out.
//
// Why not plain TF32: the reference runs these contractions in fp32 (torch.mm / nn.Linear / np.dot,
// modules/model.py:163-173, 218-221) and parity is 1e-5 relative.  Each operand is therefore split as
// x = hi + lo with hi = rn_tf32(x), lo = rn_tf32(x - hi) (split_tf32_kernel, exact residual), and the
// product is evaluated as  hi·hi  +  (hi·lo + lo·hi)  — three TF32 MMAs per K step ("3xTF32").  The
// dropped lo·lo term and the rounding of lo are both <= 2^-22 relative per term.  Tensor-core
// accumulation may round toward zero, which biases long accumulation chains; two measures keep that
// below the fp32 rounding noise of the reference itself:
//   * the small correction products accumulate in their OWN TMEM accumulator (their rounding errors
//     are 2^-11 smaller), and
//   * the hi·hi chain is cut into up to TC_KCHUNKS accumulators over K; all of them are summed in
//     round-to-nearest fp32 in the epilogue.
//
// Structure (one 128 x 128 output tile per CTA, 192 threads):
//   warp 0   TMA producer: four tiles per stage (A_hi, A_lo, B_hi, B_lo; 128 rows x 32 fp32 = 128-byte
//            rows, SWIZZLE_128B) into a 3-stage ring, completion on `full` mbarriers
//   warp 1   allocates 512 TMEM columns; one elected lane issues tcgen05.mma (kind::tf32, M=128, N=128,
//            K=8) 12 times per stage and releases the stage with tcgen05.commit -> `empty` mbarrier
//   warps 2-5 epilogue: tcgen05.ld (32 lanes x 32 columns per warp and step), sum of the accumulators,
//            fused epilogue (bias / ELU / residual, GSS loss + dG, split-K partials, radix histogram)
#include <algorithm>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace gdr {

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 32, TC_STAGES = 3, TC_THREADS = 192;
constexpr int TC_KCHUNKS = 3;                                 // hi·hi accumulators (+1 for the corrections)
constexpr int TC_TILE_BYTES = TC_BM * TC_BK * 4;              // 16 KB per operand tile
constexpr int TC_STAGE_BYTES = 4 * TC_TILE_BYTES;             // A_hi, A_lo, B_hi, B_lo
constexpr int TC_SMEM_BYTES = TC_STAGES * TC_STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;

struct TcParams {
  int64_t m, n, k;         // k = full reduction length (all of it unless split-K)
  int64_t k_per_split;     // EPI_SPLITK: K range of blockIdx.z (multiple of TC_BK); otherwise >= k
  int64_t k_a2;            // RAW_A: columns [k_a2, k) of A come from the second raw source (k_a2 % TC_BK == 0)
  GemmParams g;            // epilogue operands, same meaning as in gemm_simt.cuh (a/b/lda/ldb unused)
};

// ---- operand split ---------------------------------------------------------------------------------
__device__ __forceinline__ float rn_tf32(float x) {
  uint32_t u = __float_as_uint(x);
  u += 0x00000fffu + ((u >> 13) & 1u);  // round to nearest even on the 13 dropped bits
  return __uint_as_float(u & 0xffffe000u);
}

// hi/lo [rows, k0+k1] (ld = k0+k1) from up to two row-major sources placed side by side along K
__global__ void __launch_bounds__(256)
split_tf32_kernel(const float *__restrict__ s0, int64_t ld0, int64_t k0, const float *__restrict__ s1, int64_t ld1,
                  int64_t k1, int64_t rows, float *__restrict__ hi, float *__restrict__ lo) {
  const int64_t kq = (k0 + k1) >> 2;  // float4 per output row
  const int64_t total = rows * kq;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / kq, c = (i - r * kq) << 2;
    const float4 v = c < k0 ? ldg_f4(s0 + r * ld0 + c) : ldg_f4(s1 + r * ld1 + (c - k0));
    float4 h, l;
    h.x = rn_tf32(v.x); l.x = rn_tf32(v.x - h.x);
    h.y = rn_tf32(v.y); l.y = rn_tf32(v.y - h.y);
    h.z = rn_tf32(v.z); l.z = rn_tf32(v.z - h.z);
    h.w = rn_tf32(v.w); l.w = rn_tf32(v.w - h.w);
    *reinterpret_cast<float4 *>(hi + r * (k0 + k1) + c) = h;
    *reinterpret_cast<float4 *>(lo + r * (k0 + k1) + c) = l;
  }
}

// hi/lo of the TRANSPOSE of a row-major [rows, cols] matrix: out[c, r], ld_out = rows_padded
__global__ void __launch_bounds__(256)
split_tf32_transpose_kernel(const float *__restrict__ s, int64_t ld, int64_t rows, int64_t cols, int64_t ld_out,
                            float *__restrict__ hi, float *__restrict__ lo) {
  __shared__ float tile[32][33];
  const int64_t r0 = (int64_t)blockIdx.x * 32, c0 = (int64_t)blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = r0 + ty + i * 8, c = c0 + tx;
    tile[ty + i * 8][tx] = (r < rows && c < cols) ? __ldg(s + r * ld + c) : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t c = c0 + ty + i * 8, r = r0 + tx;
    if (c < cols && r < ld_out) {
      const float v = tile[tx][ty + i * 8];  // zero beyond `rows`: the padding columns of the output
      const float h = rn_tf32(v);
      hi[c * ld_out + r] = h;
      lo[c * ld_out + r] = rn_tf32(v - h);
    }
  }
}

// ---- fused epilogues: one thread owns 32 consecutive columns [n, n+32) of output row m -----------------
template <int EPI>
__device__ __forceinline__ void tc_epilogue_row(const TcParams &p, int64_t m, int64_t n, float (&v)[32], double &lsum,
                                                unsigned int *sh_hist, unsigned int &best) {
  const GemmParams &g = p.g;
  if (EPI == EPI_HIST) {
    const unsigned int prefix = g.sel->prefix, mask = g.sel->mask;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (n + j < p.n) {
        const unsigned int key = f32_key(v[j]);
        if ((key & mask) == prefix) atomicAdd(&sh_hist[(key >> g.shift) & g.bin_mask], 1u);
      }
    }
    return;
  }
  if (EPI == EPI_MINGT) {
    const unsigned int base = g.sel->prefix;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (n + j < p.n) {
        const unsigned int key = f32_key(v[j]);
        if (key > base) best = min(best, key);
      }
    }
    return;
  }
  if (EPI == EPI_GSS) {  // n need not be a multiple of 4; ldc is
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float w = 0.f;
      if (n + j < p.n) {
        const float gv = v[j];
        const float r = fmaxf(gv, 0.f) - g.beta;
        lsum += (double)(-0.5f * g.alpha * r * r);
        w = gv > 0.f ? -g.alpha * r * g.wscale : 0.f;
      }
      v[j] = w;
    }
  }
  float *crow = g.c + (EPI == EPI_SPLITK ? ((int64_t)blockIdx.z * p.m + m) : m) * g.ldc + n;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int64_t nn = n + q * 4;
    if (nn >= (EPI == EPI_GSS ? g.ldc : p.n)) break;  // every other caller has n % 4 == 0
    float a0 = v[q * 4 + 0], a1 = v[q * 4 + 1], a2 = v[q * 4 + 2], a3 = v[q * 4 + 3];
    if (EPI == EPI_STORE && g.e1) {
      const float4 x = ldg_f4(g.e1 + m * g.ldc + nn), y = ldg_f4(g.e2 + m * g.ldc + nn);
      a0 = fmaf(x.x, y.x, a0); a1 = fmaf(x.y, y.y, a1); a2 = fmaf(x.z, y.z, a2); a3 = fmaf(x.w, y.w, a3);
    }
    if (EPI == EPI_LAYER) {
      const float4 b1 = ldg_f4(g.bias1 + nn), b2 = ldg_f4(g.bias2 + nn);
      a0 += b1.x + b2.x; a1 += b1.y + b2.y; a2 += b1.z + b2.z; a3 += b1.w + b2.w;
    }
    *reinterpret_cast<float4 *>(crow + q * 4) = make_float4(a0, a1, a2, a3);
    if (EPI == EPI_LAYER) {
      float y0 = elu1(a0), y1 = elu1(a1), y2 = elu1(a2), y3 = elu1(a3);
      if (g.residual) {
        const float4 r = ldg_f4(g.residual + m * g.ldc + nn);
        y0 = fmaf(g.decay, y0, r.x); y1 = fmaf(g.decay, y1, r.y);
        y2 = fmaf(g.decay, y2, r.z); y3 = fmaf(g.decay, y3, r.w);
      }
      *reinterpret_cast<float4 *>(g.xnext + m * g.ldc + nn) = make_float4(y0, y1, y2, y3);
    }
  }
}

// RAW_A: operand A arrives as plain fp32 (map_ahi = first source, map_alo = second source placed after it along
// K) and is split into hi/lo in shared memory by the four epilogue warps while the pipeline runs: the TMA tile is
// rewritten in place as hi, lo goes to the neighbouring slot, fence.proxy.async makes both visible to the
// tensor core, and a `conv` mbarrier (128 arrivals) replaces `full` as the MMA issuer's go signal.
template <int EPI, bool RAW_A>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_ahi, const __grid_constant__ CUtensorMap map_alo,
               const __grid_constant__ CUtensorMap map_bhi, const __grid_constant__ CUtensorMap map_blo,
               const TcParams p) {
  constexpr bool SYMMETRIC = (EPI == EPI_HIST || EPI == EPI_MINGT);
  if (SYMMETRIC && blockIdx.x < blockIdx.y) return;  // mirrored tile counted with weight 2

  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + TC_STAGES * TC_STAGE_BYTES);
  uint64_t *empty = full + TC_STAGES;
  uint64_t *acc_ready = empty + TC_STAGES;
  uint64_t *conv = acc_ready + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(conv + TC_STAGES);
  __shared__ unsigned int sh_hist[SYMMETRIC ? GHIST_BINS : 1];
  __shared__ double red[4];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t m0 = (int64_t)blockIdx.y * TC_BM, n0 = (int64_t)blockIdx.x * TC_BN;
  int64_t k_begin = 0, k_end = p.k;
  if (EPI == EPI_SPLITK) {
    k_begin = (int64_t)blockIdx.z * p.k_per_split;
    k_end = min(p.k, k_begin + p.k_per_split);
  }
  const int nkb = (int)((k_end - k_begin + TC_BK - 1) / TC_BK);  // >= 1 by construction of the grid
  // hi·hi chunk boundaries (in k-blocks): chunk c covers [c*per, (c+1)*per)
  const int per = (nkb + TC_KCHUNKS - 1) / TC_KCHUNKS;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_STAGES; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
      tc::mbar_init(&conv[s], 128);
    }
    tc::mbar_init(acc_ready, 1);
    tc::mbar_fence_init();
    tc::tma_prefetch_desc(&map_ahi);
    tc::tma_prefetch_desc(&map_alo);
    tc::tma_prefetch_desc(&map_bhi);
    tc::tma_prefetch_desc(&map_blo);
  }
  if (SYMMETRIC)
    for (int i = threadIdx.x; i < GHIST_BINS; i += TC_THREADS) sh_hist[i] = 0u;
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % TC_STAGES;
        const uint32_t ph = (uint32_t)(kb / TC_STAGES) & 1u;
        tc::mbar_wait(&empty[s], ph ^ 1u);  // first lap passes immediately
        unsigned char *st = smem + s * TC_STAGE_BYTES;
        tc::mbar_expect_tx(&full[s], RAW_A ? 3 * TC_TILE_BYTES : TC_STAGE_BYTES);
        const int32_t kc = (int32_t)(k_begin + (int64_t)kb * TC_BK);
        if (RAW_A) {
          if (kc < p.k_a2)
            tc::tma_load_2d(st + 0 * TC_TILE_BYTES, &map_ahi, &full[s], kc, (int32_t)m0);
          else
            tc::tma_load_2d(st + 0 * TC_TILE_BYTES, &map_alo, &full[s], kc - (int32_t)p.k_a2, (int32_t)m0);
        } else {
          tc::tma_load_2d(st + 0 * TC_TILE_BYTES, &map_ahi, &full[s], kc, (int32_t)m0);
          tc::tma_load_2d(st + 1 * TC_TILE_BYTES, &map_alo, &full[s], kc, (int32_t)m0);
        }
        tc::tma_load_2d(st + 2 * TC_TILE_BYTES, &map_bhi, &full[s], kc, (int32_t)n0);
        tc::tma_load_2d(st + 3 * TC_TILE_BYTES, &map_blo, &full[s], kc, (int32_t)n0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_tf32(TC_BM, TC_BN);
      const uint32_t acc_corr = tmem + TC_KCHUNKS * TC_BN;  // columns [384, 512)
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % TC_STAGES;
        const uint32_t ph = (uint32_t)(kb / TC_STAGES) & 1u;
        tc::mbar_wait(RAW_A ? &conv[s] : &full[s], ph);
        tc::tc_fence_after();
        unsigned char *st = smem + s * TC_STAGE_BYTES;
        const uint64_t d_ahi = tc::smem_desc_k_sw128(st + 0 * TC_TILE_BYTES);
        const uint64_t d_alo = tc::smem_desc_k_sw128(st + 1 * TC_TILE_BYTES);
        const uint64_t d_bhi = tc::smem_desc_k_sw128(st + 2 * TC_TILE_BYTES);
        const uint64_t d_blo = tc::smem_desc_k_sw128(st + 3 * TC_TILE_BYTES);
        const int chunk = kb / per;
        const uint32_t acc_main = tmem + (uint32_t)chunk * TC_BN;
        const bool first_of_chunk = (kb % per) == 0;
#pragma unroll
        for (int k = 0; k < TC_BK / 8; ++k) {
          const uint32_t off = (uint32_t)k * 32u;  // 8 TF32 = 32 bytes inside the swizzled 128-byte row
          tc::mma_tf32(acc_main, tc::smem_desc_advance(d_ahi, off), tc::smem_desc_advance(d_bhi, off), idesc,
                       (first_of_chunk && k == 0) ? 0u : 1u);
          tc::mma_tf32(acc_corr, tc::smem_desc_advance(d_ahi, off), tc::smem_desc_advance(d_blo, off), idesc,
                       (kb == 0 && k == 0) ? 0u : 1u);
          tc::mma_tf32(acc_corr, tc::smem_desc_advance(d_alo, off), tc::smem_desc_advance(d_bhi, off), idesc, 1u);
        }
        tc::mma_commit(&empty[s]);  // frees the stage once these MMAs have read it
      }
      tc::mma_commit(acc_ready);
    }
  } else {
    // ===== epilogue (warps 2..5; TMEM lane quarter = warp % 4) =====
    const int q = warp & 3;
    if (RAW_A) {
      const int t = threadIdx.x - 64;  // 0..127
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % TC_STAGES;
        const uint32_t ph = (uint32_t)(kb / TC_STAGES) & 1u;
        tc::mbar_wait(&full[s], ph);
        float4 *hi4 = reinterpret_cast<float4 *>(smem + s * TC_STAGE_BYTES);
        float4 *lo4 = reinterpret_cast<float4 *>(smem + s * TC_STAGE_BYTES + TC_TILE_BYTES);
#pragma unroll
        for (int i = 0; i < TC_TILE_BYTES / 16 / 128; ++i) {  // the swizzle is the same in both tiles
          const float4 x = hi4[t + i * 128];
          float4 h, l;
          h.x = rn_tf32(x.x); l.x = rn_tf32(x.x - h.x);
          h.y = rn_tf32(x.y); l.y = rn_tf32(x.y - h.y);
          h.z = rn_tf32(x.z); l.z = rn_tf32(x.z - h.z);
          h.w = rn_tf32(x.w); l.w = rn_tf32(x.w - h.w);
          hi4[t + i * 128] = h;
          lo4[t + i * 128] = l;
        }
        tc::fence_proxy_async();
        tc::mbar_arrive(&conv[s]);
      }
    }
    tc::mbar_wait(acc_ready, 0u);
    tc::tc_fence_after();
    const int64_t m = m0 + q * 32 + lane;
    const int nchunks = (nkb + per - 1) / per;  // hi·hi accumulators actually written
    double lsum = 0.0;
    unsigned int best = 0xffffffffu;
#pragma unroll 1
    for (int c0 = 0; c0 < TC_BN; c0 += 32) {
      const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
      float v[32], t[32];
      tc::tmem_ld_32x32(lane_base + (uint32_t)(TC_KCHUNKS * TC_BN + c0), v);  // corrections first (smallest)
      for (int c = nchunks - 1; c >= 0; --c) {
        tc::tmem_ld_32x32(lane_base + (uint32_t)(c * TC_BN + c0), t);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] += t[j];
      }
      if (m < p.m && n0 + c0 < p.n) tc_epilogue_row<EPI>(p, m, n0 + c0, v, lsum, sh_hist, best);
    }
    if (EPI == EPI_GSS) {
      lsum = warp_sum(lsum);
      if (lane == 0) red[q] = lsum;
    }
    if (EPI == EPI_MINGT) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
      if (lane == 0 && best != 0xffffffffu) atomicMin(&p.g.sel_out->min_gt, best);
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem, 512);
  }
  if (EPI == EPI_GSS && threadIdx.x == 0)
    p.g.loss_partials[(int64_t)blockIdx.y * gridDim.x + blockIdx.x] = (red[0] + red[1]) + (red[2] + red[3]);
  if (EPI == EPI_HIST) {
    const unsigned long long w = (blockIdx.x == blockIdx.y) ? 1ull : 2ull;
    for (int i = threadIdx.x; i < GHIST_BINS; i += TC_THREADS) {
      const unsigned int c = sh_hist[i];
      if (c) atomicAdd(p.g.hist + i, w * c);
    }
  }
}

// ---- host side ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// [rows, k] fp32 row-major (leading dimension ld elements) -> boxes of 128 rows x 32 columns, 128B swizzle,
// out-of-range elements read as zero (ragged M, N and K tails need no masking in the kernel)
static int make_operand_map(CUtensorMap *map, const float *base, int64_t rows, int64_t k, int64_t ld) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) {
    set_error("gemm_tc: cuTensorMapEncodeTiled is not available from this driver");
    return GDR_E_INTERNAL;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)ld * 4u};
  const cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)TC_BM};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("gemm_tc: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return GDR_E_INTERNAL;
  }
  return GDR_OK;
}

// RAW_A: ahi = first raw source [m, k_a2] (lda), alo = second raw source [m, k - k_a2] (lda2) or nullptr
template <int EPI, bool RAW_A>
static int launch_gemm_tc(const float *ahi, const float *alo, const float *bhi, const float *blo, int64_t m, int64_t n,
                          int64_t k, int64_t lda, int64_t ldb, int64_t k_per_split, int splits, const GemmParams &g,
                          cudaStream_t st, int64_t k_a2 = 0, int64_t lda2 = 0) {
  int rc;
  if ((lda % 4) || (ldb % 4) || (((uintptr_t)ahi | (uintptr_t)alo | (uintptr_t)bhi | (uintptr_t)blo) & 15))
  {
    set_error("gemm_tc: operands must be 16-byte aligned with leading dimensions %% 4 == 0");
    return GDR_E_SHAPE;
  }
  CUtensorMap ma, ml, mb, mbl;
  if (RAW_A) {
    if (k_a2 <= 0 || k_a2 > k) k_a2 = k;
    if (k_a2 < k && (!alo || (k_a2 % TC_BK) || (lda2 % 4))) {
      set_error("gemm_tc: the second raw A source needs k_a2 %% %d == 0 and lda2 %% 4 == 0", TC_BK);
      return GDR_E_SHAPE;
    }
    if ((rc = make_operand_map(&ma, ahi, m, k_a2, lda))) return rc;
    if ((rc = make_operand_map(&ml, k_a2 < k ? alo : ahi, m, k_a2 < k ? k - k_a2 : k_a2, k_a2 < k ? lda2 : lda))) return rc;
  } else {
    if ((rc = make_operand_map(&ma, ahi, m, k, lda))) return rc;
    if ((rc = make_operand_map(&ml, alo, m, k, lda))) return rc;
  }
  if ((rc = make_operand_map(&mb, bhi, n, k, ldb))) return rc;
  if ((rc = make_operand_map(&mbl, blo, n, k, ldb))) return rc;
  TcParams p{};
  p.m = m; p.n = n; p.k = k; p.k_per_split = k_per_split; p.k_a2 = RAW_A ? k_a2 : k; p.g = g;
  static bool attr_set = false;  // per template instance
  if (!attr_set) {
    GDR_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<EPI, RAW_A>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  TC_SMEM_BYTES));
    attr_set = true;
  }
  dim3 grid((unsigned)ceil_div<int64_t>(n, TC_BN), (unsigned)ceil_div<int64_t>(m, TC_BM), (unsigned)splits);
  gemm_tc_kernel<EPI, RAW_A><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(ma, ml, mb, mbl, p);
  GDR_LAUNCHED();
  return GDR_OK;
}

int tc_split(const float *s0, int64_t ld0, int64_t k0, const float *s1, int64_t ld1, int64_t k1, int64_t rows,
             float *hi, float *lo, cudaStream_t st) {
  const int64_t total = rows * ((k0 + k1) >> 2);
  const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div<int64_t>(total, 256), 148 * 16);
  split_tf32_kernel<<<blocks, 256, 0, st>>>(s0, ld0, k0, s1, ld1, k1, rows, hi, lo);
  GDR_LAUNCHED();
  return GDR_OK;
}

int tc_split_transpose(const float *s, int64_t ld, int64_t rows, int64_t cols, int64_t ld_out, float *hi, float *lo,
                       cudaStream_t st) {
  dim3 grid((unsigned)ceil_div<int64_t>(ld_out, 32), (unsigned)ceil_div<int64_t>(cols, 32));
  split_tf32_transpose_kernel<<<grid, 256, 0, st>>>(s, ld, rows, cols, ld_out, hi, lo);
  GDR_LAUNCHED();
  return GDR_OK;
}

int tc_gemm(int epi, const float *ahi, const float *alo, const float *bhi, const float *blo, int64_t m, int64_t n,
            int64_t k, int64_t lda, int64_t ldb, int64_t k_per_split, int splits, const GemmParams &g,
            cudaStream_t st) {
  switch (epi) {
    case EPI_STORE: return launch_gemm_tc<EPI_STORE, false>(ahi, alo, bhi, blo, m, n, k, lda, ldb, k, 1, g, st);
    case EPI_LAYER: return launch_gemm_tc<EPI_LAYER, false>(ahi, alo, bhi, blo, m, n, k, lda, ldb, k, 1, g, st);
    case EPI_SPLITK: return launch_gemm_tc<EPI_SPLITK, false>(ahi, alo, bhi, blo, m, n, k, lda, ldb, k_per_split, splits, g, st);
    case EPI_GSS: return launch_gemm_tc<EPI_GSS, false>(ahi, alo, bhi, blo, m, n, k, lda, ldb, k, 1, g, st);
    case EPI_HIST: return launch_gemm_tc<EPI_HIST, false>(ahi, alo, bhi, blo, m, n, k, lda, ldb, k, 1, g, st);
    case EPI_MINGT: return launch_gemm_tc<EPI_MINGT, false>(ahi, alo, bhi, blo, m, n, k, lda, ldb, k, 1, g, st);
  }
  set_error("gemm_tc: unknown epilogue %d", epi);
  return GDR_E_BADARG;
}

int tc_gemm_raw_a(int epi, const float *a1, int64_t lda1, int64_t k1, const float *a2, int64_t lda2, int64_t k2,
                  const float *bhi, const float *blo, int64_t ldb, int64_t m, int64_t n, const GemmParams &g,
                  cudaStream_t st) {
  const int64_t k = k1 + k2;
  switch (epi) {
    case EPI_STORE:
      return launch_gemm_tc<EPI_STORE, true>(a1, a2, bhi, blo, m, n, k, lda1, ldb, k, 1, g, st, k1, lda2);
    case EPI_LAYER:
      return launch_gemm_tc<EPI_LAYER, true>(a1, a2, bhi, blo, m, n, k, lda1, ldb, k, 1, g, st, k1, lda2);
  }
  set_error("gemm_tc: raw-A form supports the store and layer epilogues only (got %d)", epi);
  return GDR_E_BADARG;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int64_t gdr_gemm_nt_tc_workspace_bytes(int64_t m, int64_t n, int64_t k) {
  if (m < 0 || n < 0 || k <= 0) return GDR_E_BADARG;
  return 2 * align_up(m * k * 4, 256) + 2 * align_up(n * k * 4, 256) + 1024;
}

// C[m,n] = A[m,k] · B[n,k]ᵀ  (both row-major, K contiguous), fp32 in/out, 3xTF32 on tcgen05.
int gdr_gemm_nt_tc_f32(const float *a, int64_t lda, const float *b, int64_t ldb, int64_t m, int64_t n, int64_t k,
                       float *c, int64_t ldc, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(m >= 0 && n >= 0 && k > 0 && a && b && c && workspace, GDR_E_BADARG, "gemm_nt_tc: bad args");
  GDR_CHECK_ARG(k % 4 == 0 && n % 4 == 0 && lda % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0, GDR_E_SHAPE,
                "gemm_nt_tc: k, n and the leading dimensions must be multiples of 4");
  GDR_CHECK_ARG(workspace_bytes >= gdr_gemm_nt_tc_workspace_bytes(m, n, k), GDR_E_WORKSPACE,
                "gemm_nt_tc: workspace too small");
  if (m == 0 || n == 0) return GDR_OK;
  cudaStream_t st = as_stream(stream);
  Carver cv(workspace);
  float *ahi = cv.take<float>(m * k), *alo = cv.take<float>(m * k);
  float *bhi = cv.take<float>(n * k), *blo = cv.take<float>(n * k);
  int rc;
  (void)ahi; (void)alo;  // A is split inside the kernel
  if ((rc = tc_split(b, ldb, k, nullptr, 0, 0, n, bhi, blo, st))) return rc;
  GemmParams g{};
  g.c = c; g.ldc = ldc;
  return tc_gemm_raw_a(EPI_STORE, a, lda, k, nullptr, 0, 0, bhi, blo, k, m, n, g, st);
}

}  // extern "C"
