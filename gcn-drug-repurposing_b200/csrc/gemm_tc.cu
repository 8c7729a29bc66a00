// gemm_tc.cu — fp32-accurate NT GEMM on the 5th-generation tensor cores (tcgen05, sm_100a).
//
//   C[M,N] = A[M,K] · B[N,K]ᵀ   A, B row-major with K contiguous ("K-major"), fp32 in / fp32 out.
//
// Why not plain TF32: the reference runs these contractions in fp32 (torch.mm / nn.Linear / np.dot,
// modules/model.py:163-173, 218-221) and parity is 1e-5 relative.  Each operand is therefore split as
// x = hi + lo with hi = rn_tf32(x), lo = rn_tf32(x - hi) (exact residual), and the product is evaluated as
// hi·hi  +  (hi·lo + lo·hi)  — "3xTF32".  The dropped lo·lo term and the rounding of lo are both <= 2^-22
// relative per term.  The small correction products accumulate in their OWN TMEM accumulator and are added to
// the hi·hi accumulator in round-to-nearest fp32 in the epilogue.
//
// Structure: persistent warp-specialised kernel, one CTA per SM walking 128 x 128 output tiles (gemm_tc_kernel):
// TMA producer warp, single-thread MMA issuer, eight epilogue warps, and — when A arrives as plain fp32 (RAW_A) —
// four converter warps that split A on the fly and feed it to the MMAs through TENSOR MEMORY, so that the only
// shared-memory MMA operand is B.  Measured at the MSI layer shape 29,959 x 256 x 512 (profiles/tc_pipeline_trace.py,
// CTA-0 clock stamps): 888 clk per k-block in the main loop (12 x 64 = 768 is the tensor-pipe floor), 68.6 us for
// the layer forward with the weight split, 105.5 us before the A-in-TMEM / 8-warp epilogue change.
// Accuracy (tests/test_gpu_tc.py): max error 3-8e-7 of sum|a||b| (torch-CPU fp32: 2-5e-7).
#include <algorithm>
#include <atomic>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace gdr {

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 32;
constexpr int TC_STAGES = 3, TC_STAGES_RAW = 4;               // pre-split: 4 tiles per stage; RAW_A: 3 tiles per stage
constexpr int TC_EPI_WARPS = 8;                               // two per TMEM lane quarter: 64 columns each
constexpr int TC_THREADS = 64 + 32 * TC_EPI_WARPS;            // producer, MMA issuer, epilogue warps
constexpr int TC_THREADS_RAW = TC_THREADS + 128;              // + 4 converter warps when A is split in the kernel
constexpr int TC_TILE_BYTES = TC_BM * TC_BK * 4;              // 16 KB per operand tile
constexpr int TC_RING_BYTES = TC_STAGES * 4 * TC_TILE_BYTES;   // = TC_STAGES_RAW * 3 * TC_TILE_BYTES = 192 KB
constexpr int TC_STG_BYTES = TC_EPI_WARPS * 32 * 32 * 4;      // per-warp 32 x 32 transposition buffers of the epilogue
constexpr int tc_smem_bytes(bool staging) {
  return TC_RING_BYTES + (staging ? TC_STG_BYTES : 0) + 1024 /*alignment slack*/ + 512 /*barriers*/;
}

struct TcParams {
  int64_t m, n, k;         // k = full reduction length (all of it unless split-K)
  int64_t k_per_split;     // EPI_SPLITK: K range of one split (multiple of TC_BK); otherwise >= k
  int64_t splits;          // EPI_SPLITK: number of K ranges (tiles = tiles_m * tiles_n * splits); otherwise 1
  int64_t k_a2;            // RAW_A: columns [k_a2, k) of A come from the second raw source (k_a2 % TC_BK == 0)
  GemmParams g;            // epilogue operands, same meaning as in gemm_simt.cuh (a/b/lda/ldb unused)
  float *colsum_part;      // TN form: column sums of the A source per K range, [splits][m] (nullptr: not wanted)
  long long *trace;        // diagnostic: CTA 0 records clock64() at the pipeline hand-offs (gdr_debug_tc_trace)
};

// trace[role][index < 128][event < 4]; roles: 0 TMA producer, 1 MMA issuer, 2 converter (warp 6), 3 epilogue (warp 2)
__device__ __forceinline__ void tc_trace(const TcParams &p, int role, uint32_t idx, int ev) {
  if (p.trace != nullptr && blockIdx.x == 0 && idx < 128u) p.trace[(role * 128 + (int)idx) * 4 + ev] = clock64();
}
static long long *g_tc_trace = nullptr;
constexpr int TC_MAX_DEVICES = 64;  // slots of the per-device "attribute already set" flags (beyond: set every call)
long long *tc_trace_buffer() { return g_tc_trace; }

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

// ---- fused epilogues ---------------------------------------------------------------------------------------
// After the TMEM read a thread owns 32 consecutive columns of ONE output row; stores in that layout would touch
// 32 different rows per instruction.  Each epilogue warp therefore transposes its 32x32 block through a
// shared-memory buffer and works on it as 8 steps of 4 rows x 32 columns: lane l owns columns 4*(l%8)..+3 of
// row 4*i + l/8, so every global load / store of the warp covers 4 row segments of 128 contiguous bytes; the
// addend loads (residual, e1/e2) of 4 steps are issued together before any of them is consumed.
// The staging buffer is 32 rows x 8 chunks of 16 bytes, chunk c of row r stored at position c ^ (r % 8): the
// row-per-thread writes and the 4-rows-per-instruction reads are both conflict-free 128-bit accesses.
__device__ __forceinline__ uint32_t tc_stg_off(int row, int chunk) { return (uint32_t)(row * 8 + (chunk ^ (row & 7))) << 4; }

template <int EPI>
__device__ __forceinline__ void tc_epilogue_block(const TcParams &p, int64_t z, int64_t m_base, int64_t n_base,
                                                  uint32_t stg, int lane, double &lsum) {
  const GemmParams &g = p.g;
  const int rq = lane >> 3, cq = (lane & 7) * 4;
  const int64_t nn = n_base + cq;
  const int64_t n_lim = (EPI == EPI_GSS) ? g.ldc : p.n;  // every other caller has n % 4 == 0
  if (nn >= n_lim) return;
  float4 bias = make_float4(0.f, 0.f, 0.f, 0.f);
  if (EPI == EPI_LAYER) {
    const float4 b1 = ldg_f4(g.bias1 + nn), b2 = ldg_f4(g.bias2 + nn);
    bias = make_float4(b1.x + b2.x, b1.y + b2.y, b1.z + b2.z, b1.w + b2.w);
  }
  const bool has_add = (EPI == EPI_STORE && g.e1 != nullptr) || (EPI == EPI_LAYER && g.residual != nullptr);
#pragma unroll
  for (int h = 0; h < 2; ++h) {  // two batches of 4 steps: the addend loads of a batch are in flight together
    float4 add1[4], add2[4];
    if (has_add) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t m = m_base + (h * 4 + i) * 4 + rq;
        if (m < p.m) {
          if (EPI == EPI_STORE) {
            add1[i] = ldg_f4(g.e1 + m * g.ldc + nn);
            add2[i] = ldg_f4(g.e2 + m * g.ldc + nn);
          } else {
            add1[i] = ldg_f4(g.residual + m * g.ldc + nn);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = (h * 4 + i) * 4 + rq;
      const int64_t m = m_base + r;
      if (m >= p.m) continue;
      const float4 sv = tc::lds_f4(stg + tc_stg_off(r, lane & 7));
      float a0 = sv.x, a1 = sv.y, a2 = sv.z, a3 = sv.w;
      if (EPI == EPI_GSS) {
        float v[4] = {a0, a1, a2, a3};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float w = 0.f;
          if (nn + j < p.n) {
            const float gv = v[j];
            const float rr = fmaxf(gv, 0.f) - g.beta;
            lsum += (double)(-0.5f * g.alpha * rr * rr);
            w = gv > 0.f ? -g.alpha * rr * g.wscale : 0.f;
          }
          v[j] = w;
        }
        a0 = v[0]; a1 = v[1]; a2 = v[2]; a3 = v[3];
      }
      if (EPI == EPI_STORE && has_add) {
        a0 = fmaf(add1[i].x, add2[i].x, a0); a1 = fmaf(add1[i].y, add2[i].y, a1);
        a2 = fmaf(add1[i].z, add2[i].z, a2); a3 = fmaf(add1[i].w, add2[i].w, a3);
      }
      if (EPI == EPI_LAYER) {
        a0 += bias.x; a1 += bias.y; a2 += bias.z; a3 += bias.w;
      }
      float *crow = g.c + (EPI == EPI_SPLITK ? (z * p.m + m) : m) * g.ldc + nn;
      *reinterpret_cast<float4 *>(crow) = make_float4(a0, a1, a2, a3);
      if (EPI == EPI_LAYER) {
        float y0 = elu1(a0), y1 = elu1(a1), y2 = elu1(a2), y3 = elu1(a3);
        if (has_add) {
          y0 = fmaf(g.decay, y0, add1[i].x); y1 = fmaf(g.decay, y1, add1[i].y);
          y2 = fmaf(g.decay, y2, add1[i].z); y3 = fmaf(g.decay, y3, add1[i].w);
        }
        *reinterpret_cast<float4 *>(g.xnext + m * g.ldc + nn) = make_float4(y0, y1, y2, y3);
      }
    }
  }
}

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Persistent kernel: one CTA per SM loops over output tiles (n fastest, so neighbouring CTAs share an A row
// block in L2).
//   warp 0        TMA producer (whole tile stream, one ring shared by all tiles)
//   warp 1        TMEM allocation + single-thread MMA issue
//   warps 2-9     epilogue: two warps per TMEM lane quarter, 64 accumulator columns each.  A warp first DRAINS its
//                 columns (main + corrections, summed in fp32 RN) into registers and hands the accumulator back
//                 (`acc_empty`), then does the global-memory part of the epilogue from registers while the MMA
//                 issuer is already inside the next tile.
//   warps 10-13   RAW_A only: operand A never becomes a shared-memory MMA operand.  TMA lands the raw fp32 tile
//                 (128 rows x 32 columns, SWIZZLE_128B); each converter thread owns one row: 8 conflict-free
//                 LDS.128, hi = rn_tf32(x), lo = rn_tf32(x - hi), tcgen05.st into a 4-deep TMEM ring (64 columns
//                 per stage: hi | lo), and the MMAs read A from tensor memory (tcgen05.mma [d], [a], b-desc).
// Shared-memory traffic per k-block: pre-split form 64 KB TMA + 96 KB operand reads; RAW_A form 48 KB TMA + 16 KB
// converter reads + 48 KB B reads — the SS form of kind::tf32 at M = N = 128 needs the full 128 B/clk for its
// operand reads alone.
// TMEM columns:  pre-split  [set0 main | set0 corr | set1 main | set1 corr]      (2 accumulator sets)
//                RAW_A      [main | corr | A ring: 4 x (hi 32 | lo 32)]          (1 set; the drain into registers
//                           costs the MMA pipe ~1k of the ~17k clocks of a K = 512 tile)
// The hi·hi products and the corrections keep separate accumulators: the tensor core truncates when it adds into
// the accumulator, so the error grows with the number of MMAs that hit the LARGE accumulator (measured: one shared
// accumulator, 192 adds at K = 512, 2.1e-6 of sum|a||b|; separate, 64 adds, 8e-7).
template <int EPI, bool RAW_A>
__global__ void __launch_bounds__(RAW_A ? TC_THREADS_RAW : TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_ahi, const __grid_constant__ CUtensorMap map_alo,
               const __grid_constant__ CUtensorMap map_bhi, const __grid_constant__ CUtensorMap map_blo,
               const TcParams p) {
  constexpr bool SYMMETRIC = (EPI == EPI_HIST || EPI == EPI_MINGT);
  constexpr int STAGES = RAW_A ? TC_STAGES_RAW : TC_STAGES;
  constexpr int STAGE_BYTES = RAW_A ? 3 * TC_TILE_BYTES : 4 * TC_TILE_BYTES;  // RAW_A: A raw, B_hi, B_lo
  constexpr uint32_t ACC_SETS = RAW_A ? 1 : 2;                                // [main | corr] pairs of 2 * TC_BN columns
  constexpr uint32_t A_RING_COL = 2 * TC_BN;                                  // RAW_A: first column of the A ring
  constexpr bool STAGING = !SYMMETRIC;                                        // histogram epilogues never store tiles
  static_assert(STAGES * STAGE_BYTES == TC_RING_BYTES, "both forms use the same ring size");

  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char *smem = tc_smem_raw + ((1024u - (tc::smem_u32(tc_smem_raw) & 1023u)) & 1023u);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + TC_RING_BYTES + (STAGING ? TC_STG_BYTES : 0));
  uint64_t *empty = full + 4;    // B (and pre-split A) stage consumed by the MMAs
  uint64_t *afull = empty + 4;   // RAW_A: raw A tile landed
  uint64_t *aempty = afull + 4;  // RAW_A: raw A tile read by the converter warps
  uint64_t *tfull = aempty + 4;  // RAW_A: hi/lo of the stage are in tensor memory
  uint64_t *tempty = tfull + 4;  // RAW_A: the MMAs have read that TMEM stage
  uint64_t *acc_full = tempty + 4;
  uint64_t *acc_empty = acc_full + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
  __shared__ unsigned int sh_hist[SYMMETRIC ? GHIST_BINS : 1];
  __shared__ double red[TC_EPI_WARPS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t tiles_n = ceil_div<int64_t>(p.n, TC_BN), tiles_m = ceil_div<int64_t>(p.m, TC_BM);
  const int64_t tiles_mn = tiles_m * tiles_n;
  const int64_t n_tiles = tiles_mn * p.splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
      tc::mbar_init(&afull[s], 1);
      tc::mbar_init(&aempty[s], 128);
      tc::mbar_init(&tfull[s], 128);
      tc::mbar_init(&tempty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(&acc_full[a], 1);
      tc::mbar_init(&acc_empty[a], TC_EPI_WARPS);
    }
    tc::mbar_fence_init();
    tc::tma_prefetch_desc(&map_ahi);
    tc::tma_prefetch_desc(&map_alo);
    tc::tma_prefetch_desc(&map_bhi);
    tc::tma_prefetch_desc(&map_blo);
  }
  if (SYMMETRIC)
    for (int i = threadIdx.x; i < GHIST_BINS; i += blockDim.x) sh_hist[i] = 0u;
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // every role walks the same tile sequence; `it` counts k-blocks, `tl` processed tiles of this CTA
  auto tile_coords = [&](int64_t tile, int64_t &z, int64_t &mb, int64_t &nb, int &nkb, int64_t &k_begin) -> bool {
    z = tile / tiles_mn;
    const int64_t r = tile - z * tiles_mn;
    mb = r / tiles_n;
    nb = r - mb * tiles_n;
    if (SYMMETRIC && nb < mb) return false;  // mirrored tile counted with weight 2
    k_begin = 0;
    int64_t k_end = p.k;
    if (EPI == EPI_SPLITK) {
      k_begin = z * p.k_per_split;
      k_end = min(p.k, k_begin + p.k_per_split);
    }
    nkb = (int)((k_end - k_begin + TC_BK - 1) / TC_BK);  // >= 1 by construction of the split
    return true;
  };

  if (warp == 0) {
    // ===== TMA producer: the whole warp walks the loop, one elected lane issues (tc::elect_one) =====
    {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        int64_t z, mb, nb, k_begin;
        int nkb;
        if (!tile_coords(tile, z, mb, nb, nkb, k_begin)) continue;
        const int32_t m0 = (int32_t)(mb * TC_BM), n0 = (int32_t)(nb * TC_BN);
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1u;
          unsigned char *st = smem + s * STAGE_BYTES;
          const int32_t kc = (int32_t)(k_begin + (int64_t)kb * TC_BK);
          if (RAW_A) {
            tc::mbar_wait(&aempty[s], ph ^ 1u);  // first lap passes immediately
            if (lane == 0) tc_trace(p, 0, it, 0);
            if (tc::elect_one()) {
              tc::mbar_expect_tx(&afull[s], TC_TILE_BYTES);
              if (kc < p.k_a2)
                tc::tma_load_2d(st, &map_ahi, &afull[s], kc, m0);
              else
                tc::tma_load_2d(st, &map_alo, &afull[s], kc - (int32_t)p.k_a2, m0);
            }
            __syncwarp();
            tc::mbar_wait(&empty[s], ph ^ 1u);
            if (lane == 0) tc_trace(p, 0, it, 1);
            if (tc::elect_one()) {
              tc::mbar_expect_tx(&full[s], 2 * TC_TILE_BYTES);
              tc::tma_load_2d(st + 1 * TC_TILE_BYTES, &map_bhi, &full[s], kc, n0);
              tc::tma_load_2d(st + 2 * TC_TILE_BYTES, &map_blo, &full[s], kc, n0);
            }
            __syncwarp();
            if (lane == 0) tc_trace(p, 0, it, 2);
          } else {
            tc::mbar_wait(&empty[s], ph ^ 1u);
            if (tc::elect_one()) {
              tc::mbar_expect_tx(&full[s], 4 * TC_TILE_BYTES);
              tc::tma_load_2d(st + 0 * TC_TILE_BYTES, &map_ahi, &full[s], kc, m0);
              tc::tma_load_2d(st + 1 * TC_TILE_BYTES, &map_alo, &full[s], kc, m0);
              tc::tma_load_2d(st + 2 * TC_TILE_BYTES, &map_bhi, &full[s], kc, n0);
              tc::tma_load_2d(st + 3 * TC_TILE_BYTES, &map_blo, &full[s], kc, n0);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: uniform control flow for the whole warp, the MMAs and commits under tc::elect_one() =====
    {
      constexpr uint32_t idesc = tc::idesc_tf32(TC_BM, TC_BN), idesc2 = tc::idesc_tf32(TC_BM, 2 * TC_BN);
      (void)idesc2;
      uint32_t it = 0, tl = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        int64_t z, mb, nb, k_begin;
        int nkb;
        if (!tile_coords(tile, z, mb, nb, nkb, k_begin)) continue;
        const uint32_t a = tl % ACC_SETS, aph = (tl / ACC_SETS) & 1u;
        tc::mbar_wait(&acc_empty[a], aph ^ 1u);  // the epilogue has drained this set (first use passes)
        tc::tc_fence_after();
        const uint32_t acc_main = tmem + a * (2 * TC_BN), acc_corr = acc_main + TC_BN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1u;
          unsigned char *st = smem + s * STAGE_BYTES;
          if (RAW_A) {
            tc::mbar_wait(&full[s], ph);
            if (lane == 0) tc_trace(p, 1, it, 0);
            tc::mbar_wait(&tfull[s], ph);
            if (lane == 0) tc_trace(p, 1, it, 1);
            tc::tc_fence_after();
            const uint32_t t_hi = tmem + A_RING_COL + (uint32_t)s * 64u, t_lo = t_hi + 32u;
            const uint64_t d_bhi = tc::smem_desc_k_sw128(st + 1 * TC_TILE_BYTES);  // B_lo follows at + 16 KB
            if (tc::elect_one()) {
#pragma unroll
              for (int k = 0; k < TC_BK / 8; ++k) {
                const uint32_t off = (uint32_t)k * 32u, col = (uint32_t)k * 8u;  // 8 TF32 = 32 bytes = 8 TMEM columns
                const uint32_t acc = (kb == 0 && k == 0) ? 0u : 1u;
                // B_hi and B_lo are adjacent 128-row tiles: one N = 256 MMA writes [hi·hi | hi·lo] into [main | corr]
                tc::mma_tf32_ts(acc_main, t_hi + col, tc::smem_desc_advance(d_bhi, off), idesc2, acc);
                tc::mma_tf32_ts(acc_corr, t_lo + col, tc::smem_desc_advance(d_bhi, off), idesc, 1u);
              }
              tc::mma_commit(&empty[s]);   // B stage free once these MMAs have read it
              tc::mma_commit(&tempty[s]);  // and so is the TMEM stage of A
            }
            __syncwarp();
            if (lane == 0) tc_trace(p, 1, it, 2);
          } else {
            tc::mbar_wait(&full[s], ph);
            tc::tc_fence_after();
            const uint64_t d_ahi = tc::smem_desc_k_sw128(st + 0 * TC_TILE_BYTES);
            const uint64_t d_alo = tc::smem_desc_k_sw128(st + 1 * TC_TILE_BYTES);
            const uint64_t d_bhi = tc::smem_desc_k_sw128(st + 2 * TC_TILE_BYTES);
            const uint64_t d_blo = tc::smem_desc_k_sw128(st + 3 * TC_TILE_BYTES);
            if (tc::elect_one()) {
#pragma unroll
              for (int k = 0; k < TC_BK / 8; ++k) {
                const uint32_t off = (uint32_t)k * 32u;  // 8 TF32 = 32 bytes inside the swizzled 128-byte row
                const uint32_t acc = (kb == 0 && k == 0) ? 0u : 1u;
                tc::mma_tf32(acc_main, tc::smem_desc_advance(d_ahi, off), tc::smem_desc_advance(d_bhi, off), idesc, acc);
                tc::mma_tf32(acc_corr, tc::smem_desc_advance(d_ahi, off), tc::smem_desc_advance(d_blo, off), idesc, acc);
                tc::mma_tf32(acc_corr, tc::smem_desc_advance(d_alo, off), tc::smem_desc_advance(d_bhi, off), idesc, 1u);
              }
              tc::mma_commit(&empty[s]);  // frees the stage once these MMAs have read it
            }
            __syncwarp();
          }
        }
        if (tc::elect_one()) tc::mma_commit(&acc_full[a]);
        __syncwarp();
        ++tl;
      }
    }
  } else if (warp < 2 + TC_EPI_WARPS) {
    // ===== epilogue (warps 2..9; TMEM lane quarter = warp % 4, column half = (warp - 2) / 4) =====
    const int q = warp & 3, half = (warp - 2) >> 2;
    const uint32_t stg = tc::smem_u32(smem + TC_RING_BYTES) + (uint32_t)(warp - 2) * 4096u;
    uint32_t tl = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      int64_t z, mb, nb, k_begin;
      int nkb;
      if (!tile_coords(tile, z, mb, nb, nkb, k_begin)) continue;
      const uint32_t a = tl % ACC_SETS, aph = (tl / ACC_SETS) & 1u;
      tc::mbar_wait(&acc_full[a], aph);
      if (threadIdx.x == 64) tc_trace(p, 3, tl, 0);
      tc::tc_fence_after();
      const int64_t m0 = mb * TC_BM, n0 = nb * TC_BN + half * 64;
      const int64_t m_row = m0 + q * 32 + lane;  // the row this thread owns in the TMEM layout
      const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + a * (2 * TC_BN) + (uint32_t)(half * 64);
      float v[2][32];
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        float t[32];
        tc::tmem_ld_32x32(lane_base + (uint32_t)(TC_BN + b * 32), v[b]);  // corrections first (smallest)
        tc::tmem_ld_32x32(lane_base + (uint32_t)(b * 32), t);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[b][j] += t[j];
      }
      tc::tc_fence_before();  // this warp's share of the set is in registers: hand it back to the MMA issuer
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[a]);
      if (threadIdx.x == 64) tc_trace(p, 3, tl, 1);
      double lsum = 0.0;
      unsigned int best = 0xffffffffu;
      const unsigned int wgt = (SYMMETRIC && nb != mb) ? 2u : 1u;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int64_t nc = n0 + b * 32;
        if (nc >= p.n) continue;
        if (EPI == EPI_HIST) {
          if (m_row < p.m) {
            const unsigned int prefix = p.g.sel->prefix, mask = p.g.sel->mask;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (nc + j < p.n) {
                const unsigned int key = f32_key(v[b][j]);
                if ((key & mask) == prefix) atomicAdd(&sh_hist[(key >> p.g.shift) & p.g.bin_mask], wgt);
              }
            }
          }
        } else if (EPI == EPI_MINGT) {
          if (m_row < p.m) {
            const unsigned int base = p.g.sel->prefix;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (nc + j < p.n) {
                const unsigned int key = f32_key(v[b][j]);
                if (key > base) best = min(best, key);
              }
            }
          }
        } else {
          __syncwarp();  // the previous block's staging reads are done
#pragma unroll
          for (int c = 0; c < 8; ++c)
            tc::sts_f4(stg + tc_stg_off(lane, c), make_float4(v[b][4 * c], v[b][4 * c + 1], v[b][4 * c + 2], v[b][4 * c + 3]));
          __syncwarp();
          tc_epilogue_block<EPI>(p, z, m0 + q * 32, nc, stg, lane, lsum);
        }
      }
      if (EPI == EPI_GSS) {
        lsum = warp_sum(lsum);
        named_bar_sync(1, 32 * TC_EPI_WARPS);  // the previous tile's partial has been consumed
        if (lane == 0) red[warp - 2] = lsum;
        named_bar_sync(1, 32 * TC_EPI_WARPS);
        if (warp == 2 && lane == 0)
          p.g.loss_partials[mb * tiles_n + nb] =
              ((red[0] + red[1]) + (red[2] + red[3])) + ((red[4] + red[5]) + (red[6] + red[7]));
      }
      if (EPI == EPI_MINGT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0 && best != 0xffffffffu) atomicMin(&p.g.sel_out->min_gt, best);
      }
      if (threadIdx.x == 64) tc_trace(p, 3, tl, 2);
      ++tl;
    }
  } else if (RAW_A) {
    // ===== converter warps (10..13): raw fp32 A row -> TF32 hi / lo -> tensor memory =====
    const int q = warp & 3;                      // the TMEM lane quarter this warp may touch
    const int r = q * 32 + lane;                 // tile row owned by this thread
    const uint32_t row_off = (uint32_t)r * 128u; // 128-byte rows; chunk j of row r sits at j ^ (r % 8) (SWIZZLE_128B)
    const uint32_t sw = (uint32_t)(r & 7);
    const uint32_t smem_base = tc::smem_u32(smem);
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + A_RING_COL;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      int64_t z, mb, nb, k_begin;
      int nkb;
      if (!tile_coords(tile, z, mb, nb, nkb, k_begin)) continue;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        tc::mbar_wait(&afull[s], ph);
        if (threadIdx.x == TC_THREADS) tc_trace(p, 2, it, 0);
        const uint32_t src = smem_base + (uint32_t)s * STAGE_BYTES + row_off;
        float4 x[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) x[j] = tc::lds_f4(src + ((((uint32_t)j) ^ sw) << 4));
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float e[4] = {x[j].x, x[j].y, x[j].z, x[j].w};
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const uint32_t h = tc::cvt_tf32(e[c]);
            hi[j * 4 + c] = h;
            lo[j * 4 + c] = tc::cvt_tf32(e[c] - __uint_as_float(h));
          }
        }
        tc::mbar_arrive(&aempty[s]);         // the raw tile is in registers: TMA may refill the slot
        if (threadIdx.x == TC_THREADS) tc_trace(p, 2, it, 1);
        tc::mbar_wait(&tempty[s], ph ^ 1u);  // the MMAs of the previous lap have read this TMEM stage
        if (threadIdx.x == TC_THREADS) tc_trace(p, 2, it, 2);
        tc::tc_fence_after();
        tc::tmem_st_32x32(t_lane + (uint32_t)s * 64u, hi);
        tc::tmem_st_32x32(t_lane + (uint32_t)s * 64u + 32u, lo);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        tc::mbar_arrive(&tfull[s]);
        if (threadIdx.x == TC_THREADS) tc_trace(p, 2, it, 3);
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem, 512);
  }
  if (EPI == EPI_HIST) {
    for (int i = threadIdx.x; i < GHIST_BINS; i += blockDim.x) {
      const unsigned int c = sh_hist[i];
      if (c) atomicAdd(p.g.hist + i, (unsigned long long)c);
    }
  }
}

// ---- TN form: C[M,N] = Aᵀ·B with A [K, M] and B [K, N] row-major (the weight gradient dW = Gᵀ·[A1 | A2],
// modules/model.py:165,170 under autograd: the reduction runs over the NODE dimension, so both operands are
// "MN-major").  Same pipeline as gemm_tc_kernel<., RAW_A>, both operands raw fp32:
//   A: TMA box [32 k x 128 m], unswizzled; converter thread m reads its COLUMN (32 conflict-free LDS.32), splits it
//      and stores it as its TMEM lane — the transposition costs nothing; its fp64 running sum is the column sum of
//      the A source (db = colsum(G)) for free.
//   B: four TMA boxes [32 k x 32 n], SWIZZLE_128B_ATOM_32B = the canonical MN-major TF32 UMMA layout; four more warps split the
//      tile in place (hi) and into the neighbouring slot (lo); the MMAs read it through an MN-major descriptor.
// Split-K: a work item is (m tile, n tile, K range); partial tiles go to part[z][M][N], reduced by
// splitk_reduce2_kernel in a fixed order (deterministic).  This replaces three transposing split kernels, the
// pre-split GEMM and the column-sum kernel of the round-1 path.
constexpr int TN_THREADS = TC_THREADS + 256;  // + 4 A-converter warps + 4 B-converter warps

__global__ void __launch_bounds__(TN_THREADS, 1)
gemm_tc_tn_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b1,
                  const __grid_constant__ CUtensorMap map_b2, const TcParams p) {
  constexpr int STAGES = TC_STAGES_RAW;
  constexpr int STAGE_BYTES = 3 * TC_TILE_BYTES;  // A raw | B raw -> B_hi | B_lo
  constexpr uint32_t A_RING_COL = 2 * TC_BN;

  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char *smem = tc_smem_raw + ((1024u - (tc::smem_u32(tc_smem_raw) & 1023u)) & 1023u);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + TC_RING_BYTES + TC_STG_BYTES);  // raw B tile landed
  uint64_t *empty = full + 4;    // B stage consumed by the MMAs
  uint64_t *afull = empty + 4;   // raw A tile landed
  uint64_t *aempty = afull + 4;  // raw A tile read by the A converter
  uint64_t *tfull = aempty + 4;  // A hi/lo in tensor memory
  uint64_t *tempty = tfull + 4;  // the MMAs have read that TMEM stage
  uint64_t *bconv = tempty + 4;  // B hi/lo written to shared memory
  uint64_t *acc_full = bconv + 4;
  uint64_t *acc_empty = acc_full + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t tiles_n = ceil_div<int64_t>(p.n, TC_BN), tiles_m = ceil_div<int64_t>(p.m, TC_BM);
  const int64_t tiles_mn = tiles_m * tiles_n;
  const int64_t n_tiles = tiles_mn * p.splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
      tc::mbar_init(&afull[s], 1);
      tc::mbar_init(&aempty[s], 128);
      tc::mbar_init(&tfull[s], 128);
      tc::mbar_init(&tempty[s], 1);
      tc::mbar_init(&bconv[s], 128);
    }
    tc::mbar_init(acc_full, 1);
    tc::mbar_init(acc_empty, TC_EPI_WARPS);
    tc::mbar_fence_init();
    tc::tma_prefetch_desc(&map_a);
    tc::tma_prefetch_desc(&map_b1);
    tc::tma_prefetch_desc(&map_b2);
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto tile_coords = [&](int64_t tile, int64_t &z, int64_t &mb, int64_t &nb, int &nkb, int64_t &k_begin) {
    z = tile / tiles_mn;
    const int64_t r = tile - z * tiles_mn;
    mb = r / tiles_n;
    nb = r - mb * tiles_n;
    k_begin = z * p.k_per_split;
    const int64_t k_end = min(p.k, k_begin + p.k_per_split);
    nkb = (int)((k_end - k_begin + TC_BK - 1) / TC_BK);
  };

  if (warp == 0) {
    // ===== TMA producer (whole warp, elected lane issues) =====
    {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        int64_t z, mb, nb, k_begin;
        int nkb;
        tile_coords(tile, z, mb, nb, nkb, k_begin);
        const int32_t m0 = (int32_t)(mb * TC_BM);
        int32_t n0 = (int32_t)(nb * TC_BN);
        const bool second = n0 >= p.k_a2;  // k_a2 = width of the first B source (a multiple of TC_BN)
        if (second) n0 -= (int32_t)p.k_a2;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1u;
          unsigned char *st = smem + s * STAGE_BYTES;
          const int32_t kc = (int32_t)(k_begin + (int64_t)kb * TC_BK);
          tc::mbar_wait(&aempty[s], ph ^ 1u);
          if (tc::elect_one()) {
            tc::mbar_expect_tx(&afull[s], TC_TILE_BYTES);
            tc::tma_load_2d(st, &map_a, &afull[s], m0, kc);
          }
          __syncwarp();
          tc::mbar_wait(&empty[s], ph ^ 1u);
          if (tc::elect_one()) {
            tc::mbar_expect_tx(&full[s], TC_TILE_BYTES);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (second)
                tc::tma_load_2d(st + TC_TILE_BYTES + j * 4096, &map_b2, &full[s], n0 + 32 * j, kc);
              else
                tc::tma_load_2d(st + TC_TILE_BYTES + j * 4096, &map_b1, &full[s], n0 + 32 * j, kc);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (whole warp, elected lane issues) =====
    {
      constexpr uint32_t idesc = tc::idesc_tf32_bmn(TC_BM, TC_BN), idesc2 = tc::idesc_tf32_bmn(TC_BM, 2 * TC_BN);
      uint32_t it = 0, tl = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        int64_t z, mb, nb, k_begin;
        int nkb;
        tile_coords(tile, z, mb, nb, nkb, k_begin);
        tc::mbar_wait(acc_empty, (tl & 1u) ^ 1u);
        tc::tc_fence_after();
        const uint32_t acc_main = tmem, acc_corr = tmem + TC_BN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1u;
          unsigned char *st = smem + s * STAGE_BYTES;
          tc::mbar_wait(&bconv[s], ph);
          tc::mbar_wait(&tfull[s], ph);
          tc::tc_fence_after();
          const uint32_t t_hi = tmem + A_RING_COL + (uint32_t)s * 64u, t_lo = t_hi + 32u;
          const uint64_t d_bhi = tc::smem_desc_mn_sw128_32b(st + TC_TILE_BYTES);  // B_lo = the next four 4 KB blocks
          if (tc::elect_one()) {
#pragma unroll
            for (int k = 0; k < TC_BK / 8; ++k) {
              const uint32_t off = (uint32_t)k * 1024u, col = (uint32_t)k * 8u;  // 8 K-rows of 128 bytes
              const uint32_t acc = (kb == 0 && k == 0) ? 0u : 1u;
              tc::mma_tf32_ts(acc_main, t_hi + col, tc::smem_desc_advance(d_bhi, off), idesc2, acc);
              tc::mma_tf32_ts(acc_corr, t_lo + col, tc::smem_desc_advance(d_bhi, off), idesc, 1u);
            }
            tc::mma_commit(&empty[s]);
            tc::mma_commit(&tempty[s]);
          }
          __syncwarp();
        }
        if (tc::elect_one()) tc::mma_commit(acc_full);
        __syncwarp();
        ++tl;
      }
    }
  } else if (warp < 2 + TC_EPI_WARPS) {
    // ===== epilogue: partial tile -> part[z][M][N] =====
    const int q = warp & 3, half = (warp - 2) >> 2;
    const uint32_t stg = tc::smem_u32(smem + TC_RING_BYTES) + (uint32_t)(warp - 2) * 4096u;
    uint32_t tl = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      int64_t z, mb, nb, k_begin;
      int nkb;
      tile_coords(tile, z, mb, nb, nkb, k_begin);
      tc::mbar_wait(acc_full, tl & 1u);
      tc::tc_fence_after();
      const int64_t m0 = mb * TC_BM, n0 = nb * TC_BN + half * 64;
      const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 64);
      float v[2][32];
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        float t[32];
        tc::tmem_ld_32x32(lane_base + (uint32_t)(TC_BN + b * 32), v[b]);
        tc::tmem_ld_32x32(lane_base + (uint32_t)(b * 32), t);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[b][j] += t[j];
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(acc_empty);
      double lsum = 0.0;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int64_t nc = n0 + b * 32;
        if (nc >= p.n) continue;
        __syncwarp();
#pragma unroll
        for (int c = 0; c < 8; ++c)
          tc::sts_f4(stg + tc_stg_off(lane, c), make_float4(v[b][4 * c], v[b][4 * c + 1], v[b][4 * c + 2], v[b][4 * c + 3]));
        __syncwarp();
        tc_epilogue_block<EPI_SPLITK>(p, z, m0 + q * 32, nc, stg, lane, lsum);
      }
      ++tl;
    }
  } else if (warp < 2 + TC_EPI_WARPS + 4) {
    // ===== A converter (warps 10..13): column m of the raw [32 k x 128 m] tile -> TMEM lane m =====
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const uint32_t smem_base = tc::smem_u32(smem) + (uint32_t)m * 4u;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + A_RING_COL;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      int64_t z, mb, nb, k_begin;
      int nkb;
      tile_coords(tile, z, mb, nb, nkb, k_begin);
      double csum = 0.0;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        tc::mbar_wait(&afull[s], ph);
        const uint32_t src = smem_base + (uint32_t)s * STAGE_BYTES;
        float x[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) x[k] = tc::lds_f32(src + (uint32_t)k * 512u);
#pragma unroll
        for (int k = 0; k < 32; ++k) csum += (double)x[k];
        tc::mbar_arrive(&aempty[s]);
        tc::mbar_wait(&tempty[s], ph ^ 1u);
        tc::tc_fence_after();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t hi[16], lo[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            hi[c] = tc::cvt_tf32(x[h * 16 + c]);
            lo[c] = tc::cvt_tf32(x[h * 16 + c] - __uint_as_float(hi[c]));
          }
          tc::tmem_st_32x16(t_lane + (uint32_t)s * 64u + (uint32_t)h * 16u, hi);
          tc::tmem_st_32x16(t_lane + (uint32_t)s * 64u + 32u + (uint32_t)h * 16u, lo);
        }
        tc::tmem_st_wait();
        tc::tc_fence_before();
        tc::mbar_arrive(&tfull[s]);
      }
      if (p.colsum_part != nullptr && nb == 0 && mb * TC_BM + m < p.m)
        p.colsum_part[z * p.m + mb * TC_BM + m] = (float)csum;
    }
  } else {
    // ===== B converter (warps 14..17): raw tile -> hi in place, lo in the neighbouring slot (same swizzle) =====
    const uint32_t t = threadIdx.x - (TC_THREADS + 128);
    const uint32_t smem_base = tc::smem_u32(smem) + TC_TILE_BYTES + t * 16u;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      int64_t z, mb, nb, k_begin;
      int nkb;
      tile_coords(tile, z, mb, nb, nkb, k_begin);
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        tc::mbar_wait(&full[s], ph);
        const uint32_t src = smem_base + (uint32_t)s * STAGE_BYTES;
        float4 x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = tc::lds_f4(src + (uint32_t)i * 2048u);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float4 h, l;
          h.x = __uint_as_float(tc::cvt_tf32(x[i].x)); l.x = __uint_as_float(tc::cvt_tf32(x[i].x - h.x));
          h.y = __uint_as_float(tc::cvt_tf32(x[i].y)); l.y = __uint_as_float(tc::cvt_tf32(x[i].y - h.y));
          h.z = __uint_as_float(tc::cvt_tf32(x[i].z)); l.z = __uint_as_float(tc::cvt_tf32(x[i].z - h.z));
          h.w = __uint_as_float(tc::cvt_tf32(x[i].w)); l.w = __uint_as_float(tc::cvt_tf32(x[i].w - h.w));
          tc::sts_f4(src + (uint32_t)i * 2048u, h);
          tc::sts_f4(src + (uint32_t)i * 2048u + TC_TILE_BYTES, l);
        }
        tc::fence_proxy_async();
        tc::mbar_arrive(&bconv[s]);
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem, 512);
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
int tc_make_operand_map(CUtensorMap *map, const float *base, int64_t rows, int64_t k, int64_t ld) {
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

// [k_rows, width] fp32 row-major -> boxes of box_w columns x 32 rows (the reduction index is the slow one)
static int make_mn_map(CUtensorMap *map, const float *base, int64_t k_rows, int64_t width, int64_t ld, int box_w,
                       bool swizzle) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) {
    set_error("gemm_tc: cuTensorMapEncodeTiled is not available from this driver");
    return GDR_E_INTERNAL;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)width, (cuuint64_t)k_rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)ld * 4u};
  const cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)TC_BK};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("gemm_tc: cuTensorMapEncodeTiled (MN-major operand) failed (%d)", (int)r);
    return GDR_E_INTERNAL;
  }
  return GDR_OK;
}

// part[z][m][n1 + n2] = A[k range z, 0:m]ᵀ · [B1 | B2][k range z, :]; colsum_part[z][m] = column sums of A (optional)
int tc_gemm_tn_splitk(const float *a, int64_t lda, int64_t m, const float *b1, int64_t ldb1, int64_t n1,
                      const float *b2, int64_t ldb2, int64_t n2, int64_t k, int64_t k_per_split, int splits,
                      float *part, float *colsum_part, cudaStream_t st) {
  if ((lda % 4) || (ldb1 % 4) || (ldb2 % 4) || (((uintptr_t)a | (uintptr_t)b1 | (uintptr_t)b2) & 15) ||
      (n2 > 0 && (n1 % TC_BN)) || ((n1 + n2) % 4) || (k_per_split % TC_BK)) {
    set_error("gemm_tc_tn: operands must be 16-byte aligned, leading dimensions %% 4 == 0, first N source %% %d == 0", TC_BN);
    return GDR_E_SHAPE;
  }
  int rc;
  CUtensorMap ma, mb1, mb2;
  if ((rc = make_mn_map(&ma, a, k, m, lda, TC_BM, false))) return rc;
  if ((rc = make_mn_map(&mb1, b1, k, n1, ldb1, 32, true))) return rc;
  if ((rc = make_mn_map(&mb2, n2 > 0 ? b2 : b1, k, n2 > 0 ? n2 : n1, n2 > 0 ? ldb2 : ldb1, 32, true))) return rc;
  TcParams p{};
  p.m = m; p.n = n1 + n2; p.k = k; p.k_per_split = k_per_split; p.splits = splits; p.k_a2 = n1;
  p.g.c = part; p.g.ldc = n1 + n2; p.colsum_part = colsum_part; p.trace = nullptr;
  constexpr int smem_bytes = tc_smem_bytes(true);
  static std::atomic<bool> attr_set[TC_MAX_DEVICES];  // the attribute is per device, not per process
  int dev = 0;
  GDR_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= TC_MAX_DEVICES || !attr_set[dev].load(std::memory_order_acquire)) {
    GDR_CUDA(cudaFuncSetAttribute(gemm_tc_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    if (dev >= 0 && dev < TC_MAX_DEVICES) attr_set[dev].store(true, std::memory_order_release);
  }
  const int64_t tiles = ceil_div<int64_t>(p.n, TC_BN) * ceil_div<int64_t>(m, TC_BM) * splits;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  gemm_tc_tn_kernel<<<grid, TN_THREADS, smem_bytes, st>>>(ma, mb1, mb2, p);
  GDR_LAUNCHED();
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
    if ((rc = tc_make_operand_map(&ma, ahi, m, k_a2, lda))) return rc;
    if ((rc = tc_make_operand_map(&ml, k_a2 < k ? alo : ahi, m, k_a2 < k ? k - k_a2 : k_a2, k_a2 < k ? lda2 : lda))) return rc;
  } else {
    if ((rc = tc_make_operand_map(&ma, ahi, m, k, lda))) return rc;
    if ((rc = tc_make_operand_map(&ml, alo, m, k, lda))) return rc;
  }
  if ((rc = tc_make_operand_map(&mb, bhi, n, k, ldb))) return rc;
  if ((rc = tc_make_operand_map(&mbl, blo, n, k, ldb))) return rc;
  TcParams p{};
  p.m = m; p.n = n; p.k = k; p.k_per_split = k_per_split; p.splits = splits > 0 ? splits : 1;
  p.k_a2 = RAW_A ? k_a2 : k; p.g = g; p.trace = g_tc_trace;
  constexpr int smem_bytes = tc_smem_bytes(EPI != EPI_HIST && EPI != EPI_MINGT);
  static std::atomic<bool> attr_set[TC_MAX_DEVICES];  // per template instance AND per device (the attribute is)
  int dev = 0;
  GDR_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= TC_MAX_DEVICES || !attr_set[dev].load(std::memory_order_acquire)) {
    GDR_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<EPI, RAW_A>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  smem_bytes));
    if (dev >= 0 && dev < TC_MAX_DEVICES) attr_set[dev].store(true, std::memory_order_release);
  }
  const int64_t tiles = ceil_div<int64_t>(n, TC_BN) * ceil_div<int64_t>(m, TC_BM) * p.splits;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);  // persistent: one CTA per SM walks the tiles
  gemm_tc_kernel<EPI, RAW_A><<<grid, RAW_A ? TC_THREADS_RAW : TC_THREADS, smem_bytes, st>>>(ma, ml, mb, mbl, p);
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

// split-K form of the above for products with few output tiles and a long reduction (dE[panel] = 2dG[panel, :]·E with
// K = B): part[z][m][n] = A[:, K range z] · B[:, K range z]ᵀ, reduced by the caller in a fixed order
int tc_gemm_raw_a_splitk(const float *a, int64_t lda, int64_t k, const float *bhi, const float *blo, int64_t ldb,
                         int64_t m, int64_t n, int64_t k_per_split, int splits, float *part, cudaStream_t st) {
  if (k_per_split % TC_BK) {
    set_error("gemm_tc: split-K ranges must be multiples of %d", TC_BK);
    return GDR_E_SHAPE;
  }
  GemmParams g{};
  g.c = part; g.ldc = n;
  return launch_gemm_tc<EPI_SPLITK, true>(a, nullptr, bhi, blo, m, n, k, lda, ldb, k_per_split, splits, g, st, k, 0);
}

}  // namespace gdr

using namespace gdr;

extern "C" {

// Diagnostic hook, deliberately NOT part of include/gdr.h (process-global state has no place in the product ABI;
// profiles/tc_pipeline_trace.py binds it by name): while `buf` (device, 4*128*4 int64, zeroed by the caller) is set,
// CTA 0 of every tensor-core GEMM launch records clock64() at its pipeline hand-offs into it — buf[role][index][event],
// roles 0 TMA producer, 1 MMA issuer, 2 converter warp, 3 epilogue warp; nullptr switches the recording off.
int gdr_debug_tc_trace(long long *buf) {
  g_tc_trace = buf;
  return GDR_OK;
}

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
