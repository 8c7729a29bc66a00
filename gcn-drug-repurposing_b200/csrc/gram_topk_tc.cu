// gram_topk_tc.cu — gen_graph's kNN selection (helpers/helper.py:39,44: x_sim = XᵀX in fp64, then
// np.argpartition(x_sim, -k, 1)[:, -k:]) with the N x N contraction on the tcgen05 tensor cores and the top-k
// kept in the epilogue, so that neither the Gram matrix nor any tile of it ever leaves the SM.
//
// The tensor cores have no fp64 path, so the selection is done in two steps whose result is the fp64 one:
//   1. candidates:  S~ = Xh·Xhᵀ with Xh = rn_tf32(X), one TF32 MMA per K step, fp32 accumulation in TMEM.  Every
//      epilogue thread owns one row of the tile and keeps a running list of the C largest S~ it has seen among the
//      columns it serves (C = 16 for k <= 8, else 32), with their column ids, in shared memory.
//   2. exact re-rank: the C x (lists per row) candidates of a row get their exact fp64 inner products (sequential
//      fma over the d features, the summation order of gdr_gram_topk_f64) and the k best of them are the answer,
//      PROVIDED no column that was dropped in step 1 can beat the k-th of them:
//          |S~_ij - S_ij| <= delta * |x_i| * |x_j|,   delta = 1.1e-3
//      (|rn_tf32(x) - x| <= 2^-11 |x| (1 + 2^-12) on both operands -> 2^-10 (1 + ...) of sum_k |x_ik x_jk|, the products
//      of two 11-bit significands are exact in fp32, the accumulation error of the tensor core is < 1e-6 of
//      sum |x_ik x_jk| (measured, tests/test_gpu_tc.py), and sum_k |x_ik x_jk| <= |x_i| |x_j|.)  A dropped column j of
//      a list has S~_ij <= theta (the list's smallest kept value), hence S_ij <= theta + delta |x_i| max_j |x_j|.
//      A row is CERTIFIED when its k-th exact value is strictly above that bound for all its lists; every row that
//      is not certified (near-ties inside the margin deeper than the list) is recomputed by the exact fp64 kernel
//      (gram_topk_f64_rows_kernel).  The index sets therefore equal those of gdr_gram_topk_f64 bit for bit, genuine
//      fp64 ties excepted, whatever the data; the margin only decides how many rows take the slow path.
//      Range: the bound above is a RELATIVE one and holds only while every operand and product is a normal fp32
//      number.  Step 0 therefore multiplies X by 2^-e, e = exponent of max |x_ik| (exact, and a common positive
//      factor changes no ranking), so that the largest entry lies in [1, 2): nothing can overflow fp32, and whatever
//      still falls below the fp32 normal range (entries or products under 2^-126 relative to the largest entry) costs
//      at most 2^-126 of absolute error per term.  The margin carries that as an absolute addend d * 2^-120; rows
//      whose scale is so small that the addend matters are simply not certified and take the fp64 path.  (|e| is
//      clamped to 500; data whose Gram entries overflow fp64 itself is outside any guarantee, here and upstream.)
//
// Kernel structure (persistent, one CTA per SM, warp-specialised; operand A never is a shared-memory operand):
//   a work item = (128-row block, range of 128-column tiles).  The row block's operand (128 x d TF32, d <= 256) is
//   loaded ONCE per item into TENSOR MEMORY (TMA -> smem -> 4 converter warps -> tcgen05.st, columns 256..256+d) and
//   stays there while the item sweeps its column tiles, so per k-block only the 16 KB B tile travels L2 -> smem
//   (the SS form would need 32 KB and is L2-fill bound at this arithmetic intensity).  Accumulators: 2 sets of 128
//   TMEM columns, the epilogue of tile t overlaps the MMAs of tile t+1.
//     warp 0       TMA producer (A blocks of the item, then the B tile stream, 6-stage ring)
//     warp 1       TMEM allocation, single-thread MMA issue (tcgen05.mma kind::tf32, A from TMEM)
//     warps 2-9    epilogue: tcgen05.ld of 64 accumulator columns per thread, threshold test, list insertion
//     warps 10-13  A converter
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace gdr {

constexpr int GT_BM = 128, GT_BN = 128, GT_BK = 32;
constexpr int GT_B_STAGES = 6, GT_A_STAGES = 2;
constexpr int GT_TILE_BYTES = GT_BM * GT_BK * 4;  // 16 KB
constexpr int GT_EPI_WARPS = 8;
constexpr int GT_LISTS_THREADS = 32 * GT_EPI_WARPS;  // 256 (row, column-half) lists per CTA
constexpr int GT_THREADS = 64 + GT_LISTS_THREADS + 128;
constexpr int GT_MAX_D = 256;                     // A resident in TMEM: 256 accumulator + d operand columns <= 512
constexpr int GT_MAX_C = 32;
constexpr uint32_t GT_A_COL = 256;
constexpr int GT_MAX_LISTS = 16;                  // lists per row = 2 * column chunks
constexpr double GT_DELTA = 1.1e-3;
constexpr int gt_smem_bytes(int c) {
  return GT_B_STAGES * GT_TILE_BYTES + GT_A_STAGES * GT_TILE_BYTES + GT_LISTS_THREADS * c * 8 + 1024 + 256;
}

struct GtParams {
  int64_t n;                // rows = columns of the Gram matrix
  int nka;                  // k-blocks = ceil(d / 32)
  int chunks;               // column ranges per row block
  int64_t tiles_per_chunk;  // column tiles per range
  int c;                    // list length (16 or 32)
  int32_t *cand_idx;        // [n][2 * chunks][c]   (-1: empty slot)
  float *cand_thr;          // [n][2 * chunks]      smallest kept value of a full list, -inf otherwise
  long long *trace;         // diagnostic (gdr_debug_tc_trace): CTA 0 stamps clock64() at the pipeline hand-offs
  uint32_t trace_skip;      // first traced tile of CTA 0 (GDR_TRACE_SKIP, default 64)
};

// trace[role][index < 128][event < 4], recorded from tile `trace_skip` of CTA 0 on (k-blocks: trace_skip * nka);
// roles: 0 TMA producer, 1 MMA issuer, 3 epilogue (warp 2)
__device__ __forceinline__ void gt_trace(const GtParams &p, int role, uint32_t idx, uint32_t skip, int ev) {
  if (p.trace != nullptr && blockIdx.x == 0 && idx >= skip && idx < skip + 128u)
    p.trace[(role * 128 + (int)(idx - skip)) * 4 + ev] = clock64();
}

// ---- step 0a: max |x_ik| over the whole matrix (bit pattern of a non-negative double orders like an integer) --------
__global__ void __launch_bounds__(256)
gram_absmax_kernel(const double *__restrict__ x, int64_t count, unsigned long long *__restrict__ absmax_bits) {
  unsigned long long m = 0ull;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(fabs(__ldg(x + i)));
    m = b > m ? b : m;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, m, o);
    m = other > m ? other : m;
  }
  if ((threadIdx.x & 31) == 0 && m != 0ull) atomicMax(absmax_bits, m);
}

// exponent e with max |x| in [2^e, 2^(e+1)), clamped; 0 for an all-zero (or non-finite) matrix
__device__ __forceinline__ int gt_scale_exponent(unsigned long long absmax_bits) {
  const double m = __longlong_as_double((long long)absmax_bits);
  if (!(m > 0.0) || isinf(m)) return 0;
  const int e = ilogb(m);
  return e < -500 ? -500 : (e > 500 ? 500 : e);
}

// ---- step 0b: Xh = rn_tf32(X * 2^-e) as fp32 [n, ldp] (zero padded), |x_i| 2^-e rounded up, max_i of it --------------
__global__ void __launch_bounds__(256)
gram_prep_kernel(const double *__restrict__ x, int64_t n, int64_t d, int64_t ldp, float *__restrict__ xh,
                 float *__restrict__ norms, unsigned int *__restrict__ maxnorm_bits,
                 const unsigned long long *__restrict__ absmax_bits, int *__restrict__ exp_out) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int e = gt_scale_exponent(*absmax_bits);
  if (blockIdx.x == 0 && threadIdx.x == 0) *exp_out = e;
  if (row >= n) return;
  const double scale = ldexp(1.0, -e);
  double s = 0.0;
  for (int64_t c = lane; c < ldp; c += 32) {
    float h = 0.f;
    if (c < d) {
      const double v = __ldg(x + row * d + c) * scale;   // exact: a power of two
      s = fma(v, v, s);
      h = __uint_as_float(tc::cvt_tf32((float)v));
    }
    xh[row * ldp + c] = h;
  }
  s = warp_sum(s);
  if (lane == 0) {
    const float nr = __double2float_ru(sqrt(s) * (1.0 + 1e-12));
    norms[row] = nr;
    atomicMax(maxnorm_bits, __float_as_uint(nr));  // non-negative floats order like their bit patterns
  }
}

// v[idx / 32][idx % 32] for a run-time idx without local memory: a 6-level tree of selects (63 SEL)
__device__ __forceinline__ float gt_select64(const float (&v)[2][32], int idx) {
  float s5[32], s4[16], s3[8], s2[4], s1[2];
#pragma unroll
  for (int i = 0; i < 32; ++i) s5[i] = (idx & 32) ? v[1][i] : v[0][i];
#pragma unroll
  for (int i = 0; i < 16; ++i) s4[i] = (idx & 16) ? s5[i + 16] : s5[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s3[i] = (idx & 8) ? s4[i + 8] : s4[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) s2[i] = (idx & 4) ? s3[i + 4] : s3[i];
#pragma unroll
  for (int i = 0; i < 2; ++i) s1[i] = (idx & 2) ? s2[i + 2] : s2[i];
  return (idx & 1) ? s1[1] : s1[0];
}

// sorted (descending) insertion into the calling thread's list; lists are position-major ([pos][256 threads]) so the
// lanes of a warp, which shift the same positions, hit 32 different banks
__device__ __noinline__ float gt_insert(float *lv, int32_t *lx, int c, float v, int32_t col) {
  int pos = c - 1;
  while (pos > 0) {
    const float pv = lv[(pos - 1) * GT_LISTS_THREADS];
    if (pv >= v) break;
    lv[pos * GT_LISTS_THREADS] = pv;
    lx[pos * GT_LISTS_THREADS] = lx[(pos - 1) * GT_LISTS_THREADS];
    --pos;
  }
  lv[pos * GT_LISTS_THREADS] = v;
  lx[pos * GT_LISTS_THREADS] = col;
  return lv[(c - 1) * GT_LISTS_THREADS];
}

__global__ void __launch_bounds__(GT_THREADS, 1)
gram_topk_tc_kernel(const __grid_constant__ CUtensorMap map_x, const GtParams p) {
  extern __shared__ unsigned char gt_smem_raw[];
  unsigned char *smem = gt_smem_raw + ((1024u - (tc::smem_u32(gt_smem_raw) & 1023u)) & 1023u);
  unsigned char *b_ring = smem;
  unsigned char *a_ring = smem + GT_B_STAGES * GT_TILE_BYTES;
  float *list_val = reinterpret_cast<float *>(a_ring + GT_A_STAGES * GT_TILE_BYTES);
  int32_t *list_idx = reinterpret_cast<int32_t *>(list_val + GT_LISTS_THREADS * p.c);
  uint64_t *full = reinterpret_cast<uint64_t *>(list_idx + GT_LISTS_THREADS * p.c);
  uint64_t *empty = full + GT_B_STAGES;
  uint64_t *afull = empty + GT_B_STAGES;
  uint64_t *aempty = afull + GT_A_STAGES;
  uint64_t *acc_full = aempty + GT_A_STAGES;
  uint64_t *acc_empty = acc_full + 2;
  uint64_t *a_ready = acc_empty + 2;  // the item's operand A is in tensor memory
  uint64_t *a_free = a_ready + 1;     // every MMA of the item has read it
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(a_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t tiles_n = ceil_div<int64_t>(p.n, GT_BN), tiles_m = ceil_div<int64_t>(p.n, GT_BM);
  const int64_t n_items = tiles_m * p.chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < GT_B_STAGES; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < GT_A_STAGES; ++s) {
      tc::mbar_init(&afull[s], 1);
      tc::mbar_init(&aempty[s], 128);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(&acc_full[a], 1);
      tc::mbar_init(&acc_empty[a], GT_EPI_WARPS);
    }
    tc::mbar_init(a_ready, 128);
    tc::mbar_init(a_free, 1);
    tc::mbar_fence_init();
    tc::tma_prefetch_desc(&map_x);
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto item_coords = [&](int64_t item, int64_t &mb, int &ch, int64_t &nb_begin, int64_t &nb_end) {
    mb = item / p.chunks;
    ch = (int)(item - mb * p.chunks);
    nb_begin = ch * p.tiles_per_chunk;
    nb_end = min(tiles_n, nb_begin + p.tiles_per_chunk);
  };

  if (warp == 0) {
    // ===== TMA producer (whole warp walks the loop, one elected lane issues) =====
    {
      const bool tracing = lane == 0;
      uint32_t ita = 0, itb = 0;
      for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
        int64_t mb, nb0, nb1;
        int ch;
        item_coords(item, mb, ch, nb0, nb1);
        for (int kb = 0; kb < p.nka; ++kb, ++ita) {
          const int s = ita % GT_A_STAGES;
          tc::mbar_wait(&aempty[s], ((ita / GT_A_STAGES) & 1u) ^ 1u);
          if (tc::elect_one()) {
            tc::mbar_expect_tx(&afull[s], GT_TILE_BYTES);
            tc::tma_load_2d(a_ring + s * GT_TILE_BYTES, &map_x, &afull[s], kb * GT_BK, (int32_t)(mb * GT_BM));
          }
          __syncwarp();
        }
        for (int64_t nb = nb0; nb < nb1; ++nb) {
          for (int kb = 0; kb < p.nka; ++kb, ++itb) {
            const int s = itb % GT_B_STAGES;
            if (tracing) gt_trace(p, 0, itb, p.trace_skip * (uint32_t)p.nka, 0);
            tc::mbar_wait(&empty[s], ((itb / GT_B_STAGES) & 1u) ^ 1u);
            if (tracing) gt_trace(p, 0, itb, p.trace_skip * (uint32_t)p.nka, 1);
            if (tc::elect_one()) {
              tc::mbar_expect_tx(&full[s], GT_TILE_BYTES);
              tc::tma_load_2d(b_ring + s * GT_TILE_BYTES, &map_x, &full[s], kb * GT_BK, (int32_t)(nb * GT_BN));
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the loop (uniform control flow), one elected lane issues =====
    {
      constexpr uint32_t idesc = tc::idesc_tf32(GT_BM, GT_BN);
      uint32_t itb = 0, tl = 0, it_item = 0;
      for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++it_item) {
        int64_t mb, nb0, nb1;
        int ch;
        item_coords(item, mb, ch, nb0, nb1);
        tc::mbar_wait(a_ready, it_item & 1u);
        tc::tc_fence_after();
        for (int64_t nb = nb0; nb < nb1; ++nb, ++tl) {
          const uint32_t a = tl & 1u;
          tc::mbar_wait(&acc_empty[a], ((tl >> 1) & 1u) ^ 1u);
          if (lane == 0) gt_trace(p, 1, itb, p.trace_skip * (uint32_t)p.nka, 3);
          tc::tc_fence_after();
          const uint32_t acc = tmem + a * GT_BN;
          for (int kb = 0; kb < p.nka; ++kb, ++itb) {
            const int s = itb % GT_B_STAGES;
            if (lane == 0) gt_trace(p, 1, itb, p.trace_skip * (uint32_t)p.nka, 0);
            tc::mbar_wait(&full[s], (itb / GT_B_STAGES) & 1u);
            if (lane == 0) gt_trace(p, 1, itb, p.trace_skip * (uint32_t)p.nka, 1);
            tc::tc_fence_after();
            const uint64_t d_b = tc::smem_desc_k_sw128(b_ring + s * GT_TILE_BYTES);
            const uint32_t t_a = tmem + GT_A_COL + (uint32_t)kb * GT_BK;
            if (tc::elect_one()) {
#pragma unroll
              for (int k = 0; k < GT_BK / 8; ++k)
                tc::mma_tf32_ts(acc, t_a + (uint32_t)k * 8u, tc::smem_desc_advance(d_b, (uint32_t)k * 32u), idesc,
                                (kb == 0 && k == 0) ? 0u : 1u);
              tc::mma_commit(&empty[s]);
            }
            __syncwarp();
            if (lane == 0) gt_trace(p, 1, itb, p.trace_skip * (uint32_t)p.nka, 2);
          }
          if (tc::elect_one()) tc::mma_commit(&acc_full[a]);
          __syncwarp();
        }
        if (tc::elect_one()) tc::mma_commit(a_free);
        __syncwarp();
      }
    }
  } else if (warp < 2 + GT_EPI_WARPS) {
    // ===== epilogue: thread = (row q*32+lane of the tile, column half) =====
    const int q = warp & 3, half = (warp - 2) >> 2;
    const int li = (warp - 2) * 32 + lane;
    float *lv = list_val + li;
    int32_t *lx = list_idx + li;
    const int lists = 2 * p.chunks;
    uint32_t tl = 0;
    for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
      int64_t mb, nb0, nb1;
      int ch;
      item_coords(item, mb, ch, nb0, nb1);
      const int64_t m_row = mb * GT_BM + q * 32 + lane;
      const bool row_ok = m_row < p.n;
      for (int pos = 0; pos < p.c; ++pos) {
        lv[pos * GT_LISTS_THREADS] = -INFINITY;
        lx[pos * GT_LISTS_THREADS] = -1;
      }
      float thr = -INFINITY;
      for (int64_t nb = nb0; nb < nb1; ++nb, ++tl) {
        const uint32_t a = tl & 1u;
        if (threadIdx.x == 64) gt_trace(p, 3, tl, p.trace_skip, 0);
        tc::mbar_wait(&acc_full[a], (tl >> 1) & 1u);
        if (threadIdx.x == 64) gt_trace(p, 3, tl, p.trace_skip, 1);
        tc::tc_fence_after();
        __syncwarp();
        const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + a * GT_BN + (uint32_t)(half * 64);
        float v[2][32];
        tc::tmem_ld_32x32(lane_base, v[0]);
        tc::tmem_ld_32x32(lane_base + 32u, v[1]);
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&acc_empty[a]);
        if (threadIdx.x == 64) gt_trace(p, 3, tl, p.trace_skip, 2);
        if (!row_ok) continue;
        // Fast reject without branches per value: the row's 64 new similarities against its threshold.  In steady
        // state (the threshold is the C-th largest of everything seen) almost nothing passes.
        float mx = v[0][0];
#pragma unroll
        for (int j = 1; j < 32; ++j) mx = fmaxf(mx, v[0][j]);
#pragma unroll
        for (int j = 0; j < 32; ++j) mx = fmaxf(mx, v[1][j]);
        if (mx > thr) {
          uint32_t m0 = 0u, m1 = 0u;  // bit j: value j of the block is above the threshold
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            m0 |= v[0][j] > thr ? (1u << j) : 0u;
            m1 |= v[1][j] > thr ? (1u << j) : 0u;
          }
          const int64_t c0 = nb * GT_BN + half * 64;
          const int64_t valid = p.n - c0;  // columns of this thread's 64 that exist (the last tile is ragged)
          if (valid < 64) {
            m0 &= valid >= 32 ? 0xffffffffu : (valid > 0 ? (1u << valid) - 1u : 0u);
            m1 &= valid > 32 ? (1u << (valid - 32)) - 1u : 0u;
          }
          // each lane walks ITS OWN hits: the loop runs max-over-lanes(#hits) times, not once per distinct position
          while (m0 | m1) {
            const int b = m0 ? 0 : 1;
            const uint32_t m = b ? m1 : m0;
            const int j = __ffs(m) - 1;
            if (b) m1 = m & (m - 1u); else m0 = m & (m - 1u);
            const float val = gt_select64(v, b * 32 + j);
            if (val > thr) thr = gt_insert(lv, lx, p.c, val, (int32_t)(c0 + b * 32 + j));  // thr may have risen meanwhile
          }
        }
        if (threadIdx.x == 64) gt_trace(p, 3, tl, p.trace_skip, 3);
      }
      if (row_ok) {
        const int64_t l = m_row * lists + ch * 2 + half;
        for (int pos = 0; pos < p.c; ++pos) p.cand_idx[l * p.c + pos] = lx[pos * GT_LISTS_THREADS];
        p.cand_thr[l] = lx[(p.c - 1) * GT_LISTS_THREADS] >= 0 ? lv[(p.c - 1) * GT_LISTS_THREADS] : -INFINITY;
      }
    }
  } else {
    // ===== A converter (warps 10..13): row r of the raw [128 x 32] block -> TMEM lane r, columns 256 + 32 kb .. =====
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
    const uint32_t a_base = tc::smem_u32(a_ring);
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + GT_A_COL;
    uint32_t ita = 0, it_item = 0;
    for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++it_item) {
      for (int kb = 0; kb < p.nka; ++kb, ++ita) {
        const int s = ita % GT_A_STAGES;
        tc::mbar_wait(&afull[s], (ita / GT_A_STAGES) & 1u);
        const uint32_t src = a_base + (uint32_t)s * GT_TILE_BYTES + row_off;
        uint32_t w[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 x = tc::lds_f4(src + ((((uint32_t)j) ^ sw) << 4));
          w[4 * j + 0] = __float_as_uint(x.x);
          w[4 * j + 1] = __float_as_uint(x.y);
          w[4 * j + 2] = __float_as_uint(x.z);
          w[4 * j + 3] = __float_as_uint(x.w);
        }
        if (kb == 0 && it_item > 0) tc::mbar_wait(a_free, (it_item - 1) & 1u);  // the previous item's MMAs are done
        tc::tc_fence_after();
        __syncwarp();
        tc::tmem_st_32x32(t_lane + (uint32_t)kb * GT_BK, w);
        tc::tmem_st_wait();
        // only now may TMA refill the slot: the st has consumed the registers, hence the LDS that filled them are
        // complete (an arrive right after the loads can overtake them — nothing else depends on their data)
        tc::mbar_arrive(&aempty[s]);
      }
      tc::tc_fence_before();
      tc::mbar_arrive(a_ready);
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem, 512);
  }
}

// ---- step 2: exact fp64 re-rank of the candidates, certification --------------------------------------------------
constexpr int RR_WARPS = 4;

__global__ void __launch_bounds__(32 * RR_WARPS)
gram_topk_rerank_kernel(const double *__restrict__ x, int64_t n, int64_t d, int k, int lists, int c,
                        const int32_t *__restrict__ cand_idx, const float *__restrict__ cand_thr,
                        const float *__restrict__ norms, const unsigned int *__restrict__ maxnorm_bits,
                        const int *__restrict__ scale_exp, int32_t *__restrict__ idx_out, double *__restrict__ val_out, int32_t *__restrict__ fail_count,
                        int32_t *__restrict__ fail_rows) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ncand = lists * c;
  double *xi = reinterpret_cast<double *>(rr_smem) + (int64_t)warp * d;
  double *cv = reinterpret_cast<double *>(rr_smem) + (int64_t)RR_WARPS * d + (int64_t)warp * ncand;
  int32_t *cx = reinterpret_cast<int32_t *>(reinterpret_cast<double *>(rr_smem) + (int64_t)RR_WARPS * (d + ncand)) +
                (int64_t)warp * ncand;
  const int64_t row = (int64_t)blockIdx.x * RR_WARPS + warp;
  if (row >= n) return;
  for (int64_t t = lane; t < d; t += 32) xi[t] = __ldg(x + row * d + t);
  __syncwarp();
  for (int t = lane; t < ncand; t += 32) {
    const int32_t j = cand_idx[row * ncand + t];
    double acc = -INFINITY;
    if (j >= 0) {
      const double *xj = x + (int64_t)j * d;
      acc = 0.0;
      for (int64_t kk = 0; kk < d; ++kk) acc = fma(xi[kk], __ldg(xj + kk), acc);  // the order of gram_topk_f64_kernel
    }
    cv[t] = acc;
    cx[t] = j;
  }
  __syncwarp();
  double vk = -INFINITY;
  for (int r = 0; r < k; ++r) {
    double bv = -INFINITY;
    int32_t bi = INT32_MAX;
    int bt = -1;
    for (int t = lane; t < ncand; t += 32) {
      const int32_t j = cx[t];
      if (j < 0) continue;
      const double v = cv[t];
      if (v > bv || (v == bv && j < bi)) {
        bv = v; bi = j; bt = t;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
      const int ot = __shfl_xor_sync(0xffffffffu, bt, o);
      if (ot >= 0 && (bt < 0 || ov > bv || (ov == bv && oi < bi))) {
        bv = ov; bi = oi; bt = ot;
      }
    }
    if (lane == 0) {
      idx_out[row * k + r] = bt >= 0 ? bi : -1;
      val_out[row * k + r] = bv;
      if (bt >= 0) cx[bt] = -1;
    }
    vk = bv;
    __syncwarp();
  }
  if (lane == 0) {
    float theta = -INFINITY;
    for (int l = 0; l < lists; ++l) theta = fmaxf(theta, cand_thr[row * lists + l]);
    // theta, the norms and the margin live in the scaled domain (X * 2^-e); vk is exact and unscaled
    const double vk_scaled = vk * ldexp(1.0, -2 * *scale_exp);
    const double margin = GT_DELTA * (double)norms[row] * (double)__uint_as_float(*maxnorm_bits) +
                          (double)d * 7.52316384526264e-37;  // d * 2^-120: terms below the fp32 normal range
    const bool certified = (theta == -INFINITY) || ((double)theta + margin < vk_scaled);
    if (!certified) fail_rows[atomicAdd(fail_count, 1)] = (int32_t)row;
  }
}

// ---- exact fallback for the rows that could not be certified: the fp64 SIMT kernel of gram.cu on a row list -------
constexpr int TKR_BM = 64, TKR_BN = 64, TKR_BK = 16, TKR_THREADS = 256, TKR_PAD = 2, TKR_MAXK = 32, TKR_QCAP = 64;

__global__ void __launch_bounds__(TKR_THREADS)
gram_topk_f64_rows_kernel(const double *__restrict__ x, int64_t n, int64_t d, int k, const int32_t *__restrict__ rows,
                          const int32_t *__restrict__ n_rows_dev, int32_t *__restrict__ idx_out,
                          double *__restrict__ val_out) {
  const int n_rows = *n_rows_dev;
  const int64_t m0 = (int64_t)blockIdx.x * TKR_BM;
  if (m0 >= n_rows) return;
  __shared__ __align__(16) double As[TKR_BK][TKR_BM + TKR_PAD];
  __shared__ __align__(16) double Bs[TKR_BK][TKR_BN + TKR_PAD];
  __shared__ double thr[TKR_BM];
  __shared__ int q_cnt[TKR_BM];
  __shared__ int32_t row_of[TKR_BM];
  extern __shared__ __align__(16) unsigned char tk_dyn[];
  double(*top_val)[TKR_MAXK] = reinterpret_cast<double(*)[TKR_MAXK]>(tk_dyn);
  double(*q_val)[TKR_QCAP] = reinterpret_cast<double(*)[TKR_QCAP]>(tk_dyn + sizeof(double) * TKR_BM * TKR_MAXK);
  int32_t(*top_idx)[TKR_MAXK] =
      reinterpret_cast<int32_t(*)[TKR_MAXK]>(tk_dyn + sizeof(double) * TKR_BM * (TKR_MAXK + TKR_QCAP));
  int32_t(*q_idx)[TKR_QCAP] = reinterpret_cast<int32_t(*)[TKR_QCAP]>(
      tk_dyn + sizeof(double) * TKR_BM * (TKR_MAXK + TKR_QCAP) + sizeof(int32_t) * TKR_BM * TKR_MAXK);

  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  for (int i = t; i < TKR_BM * TKR_MAXK; i += TKR_THREADS) {
    top_val[i / TKR_MAXK][i % TKR_MAXK] = -INFINITY;
    top_idx[i / TKR_MAXK][i % TKR_MAXK] = -1;
  }
  if (t < TKR_BM) {
    thr[t] = -INFINITY;
    q_cnt[t] = 0;
    row_of[t] = (m0 + t < n_rows) ? rows[m0 + t] : -1;
  }
  __syncthreads();
  const int lr = t >> 2, lk = (t & 3) * 4;
  const int64_t ar = row_of[lr];
  for (int64_t n0 = 0; n0 < n; n0 += TKR_BN) {
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
    for (int64_t k0 = 0; k0 < d; k0 += TKR_BK) {
      double a[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
      const int64_t br = n0 + lr;
#pragma unroll
      for (int qq = 0; qq < 4; ++qq) {
        const int64_t kk = k0 + lk + qq;
        if (kk < d) {
          if (ar >= 0) a[qq] = __ldg(x + ar * d + kk);
          if (br < n) b[qq] = __ldg(x + br * d + kk);
        }
      }
      __syncthreads();
#pragma unroll
      for (int qq = 0; qq < 4; ++qq) {
        As[lk + qq][lr] = a[qq];
        Bs[lk + qq][lr] = b[qq];
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < TKR_BK; ++kk) {
        double av[4], bv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          av[i] = As[kk][ty * 4 + i];
          bv[i] = Bs[kk][tx * 4 + i];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = ty * 4 + i;
      if (row_of[r] < 0) continue;
      const double th = thr[r];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t c = n0 + tx * 4 + j;
        if (c < n && acc[i][j] > th) {
          const int slot = atomicAdd(&q_cnt[r], 1);
          q_val[r][slot] = acc[i][j];
          q_idx[r][slot] = (int32_t)c;
        }
      }
    }
    __syncthreads();
    if (t < TKR_BM) {
      const int cnt = q_cnt[t];
      for (int s = 0; s < cnt; ++s) {
        const double v = q_val[t][s];
        const int32_t c = q_idx[t][s];
        // (value descending, column ascending): independent of the order in which the queue was filled
        if (!(v > top_val[t][k - 1] || (v == top_val[t][k - 1] && c < top_idx[t][k - 1]))) continue;
        int pos = k - 1;
        while (pos > 0 && (top_val[t][pos - 1] < v || (top_val[t][pos - 1] == v && top_idx[t][pos - 1] > c))) {
          top_val[t][pos] = top_val[t][pos - 1];
          top_idx[t][pos] = top_idx[t][pos - 1];
          --pos;
        }
        top_val[t][pos] = v;
        top_idx[t][pos] = c;
      }
      q_cnt[t] = 0;
      thr[t] = top_val[t][k - 1];
    }
    __syncthreads();
  }
  for (int i = t; i < TKR_BM * k; i += TKR_THREADS) {
    const int r = i / k, j = i % k;
    if (row_of[r] >= 0) {
      idx_out[(int64_t)row_of[r] * k + j] = top_idx[r][j];
      val_out[(int64_t)row_of[r] * k + j] = top_val[r][j];
    }
  }
}

static void gt_plan(int64_t n, int64_t k, int &c, int &chunks, int64_t &tiles_per_chunk) {
  c = k <= 8 ? 16 : GT_MAX_C;
  const int64_t tiles = ceil_div<int64_t>(n, GT_BM);
  // enough work items for ~5 waves of the 148 persistent CTAs, at most 8 column ranges
  int64_t want = ceil_div<int64_t>(5 * kNumSMs, tiles);
  want = std::max<int64_t>(1, std::min<int64_t>({want, (int64_t)GT_MAX_LISTS / 2, tiles}));
  tiles_per_chunk = ceil_div<int64_t>(tiles, want);
  chunks = (int)ceil_div<int64_t>(tiles, tiles_per_chunk);
}

}  // namespace gdr

using namespace gdr;

extern "C" {

int64_t gdr_gram_topk_tc_workspace_bytes(int64_t n, int64_t d, int64_t k) {
  if (n <= 0 || d <= 0 || k <= 0) return GDR_E_BADARG;
  int c, chunks;
  int64_t tpc;
  gt_plan(n, k, c, chunks, tpc);
  const int64_t ldp = align_up(d, 4);
  return align_up(n * ldp * 4, 256) + align_up(n * 4, 256) * 2 + align_up(n * 2 * chunks * c * 4, 256) +
         align_up(n * 2 * chunks * 4, 256) + 4096;
}

int gdr_gram_topk_tc_f64(const double *x, int64_t n, int64_t d, int64_t k, int32_t *idx_out, double *val_out,
                         int32_t *n_fallback_out, void *workspace, int64_t workspace_bytes, gdr_stream_t stream) {
  GDR_CHECK_ARG(n > 0 && d > 0 && x && idx_out && val_out && workspace, GDR_E_BADARG, "gram_topk_tc: bad args");
  GDR_CHECK_ARG(k >= 1 && k <= 16, GDR_E_SHAPE, "gram_topk_tc: k=%lld outside 1..16 (use gdr_gram_topk_f64)", (long long)k);
  GDR_CHECK_ARG(k <= n, GDR_E_SHAPE, "gram_topk_tc: k=%lld exceeds n=%lld (np.argpartition raises too)", (long long)k,
                (long long)n);
  GDR_CHECK_ARG(d <= GT_MAX_D, GDR_E_SHAPE, "gram_topk_tc: d=%lld exceeds %d (operand A is held in tensor memory; use gdr_gram_topk_f64)",
                (long long)d, GT_MAX_D);
  GDR_CHECK_ARG(n < (int64_t)INT32_MAX, GDR_E_SHAPE, "gram_topk_tc: int32 index range exceeded");
  GDR_CHECK_ARG(workspace_bytes >= gdr_gram_topk_tc_workspace_bytes(n, d, k), GDR_E_WORKSPACE,
                "gram_topk_tc: workspace too small");
  cudaStream_t st = as_stream(stream);
  int c, chunks;
  int64_t tpc;
  gt_plan(n, k, c, chunks, tpc);
  const int lists = 2 * chunks;
  const int64_t ldp = align_up(d, 4);
  Carver cv(workspace);
  float *xh = cv.take<float>(n * ldp);
  float *norms = cv.take<float>(n);
  int32_t *fail_rows = cv.take<int32_t>(n);
  int32_t *cand_idx = cv.take<int32_t>(n * lists * c);
  float *cand_thr = cv.take<float>(n * lists);
  unsigned int *scal = cv.take<unsigned int>(64);  // [0] max norm bits, [1] fail count, [2] scale exponent e,
  GDR_CUDA(cudaMemsetAsync(scal, 0, 256, st));     // [4..5] max |x| bits (fp64)
  unsigned long long *absmax_bits = reinterpret_cast<unsigned long long *>(scal + 4);
  int *scale_exp = reinterpret_cast<int *>(scal + 2);

  gram_absmax_kernel<<<(unsigned)std::min<int64_t>(ceil_div<int64_t>(n * d, 256 * 8), 8 * kNumSMs), 256, 0, st>>>(
      x, n * d, absmax_bits);
  GDR_LAUNCHED();
  gram_prep_kernel<<<(unsigned)ceil_div<int64_t>(n, 8), 256, 0, st>>>(x, n, d, ldp, xh, norms, scal, absmax_bits,
                                                                     scale_exp);
  GDR_LAUNCHED();

  CUtensorMap map;
  int rc;
  if ((rc = tc_make_operand_map(&map, xh, n, ldp, ldp))) return rc;
  GtParams p{};
  p.n = n; p.nka = (int)ceil_div<int64_t>(d, GT_BK); p.chunks = chunks; p.tiles_per_chunk = tpc; p.c = c;
  p.cand_idx = cand_idx; p.cand_thr = cand_thr; p.trace = tc_trace_buffer();
  p.trace_skip = 64u;
  if (p.trace != nullptr && getenv("GDR_TRACE_SKIP")) p.trace_skip = (uint32_t)atoi(getenv("GDR_TRACE_SKIP"));
  const int smem_bytes = gt_smem_bytes(c);
  GDR_CUDA(cudaFuncSetAttribute(gram_topk_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gt_smem_bytes(GT_MAX_C)));
  const int64_t items = ceil_div<int64_t>(n, GT_BM) * chunks;
  gram_topk_tc_kernel<<<(unsigned)std::min<int64_t>(items, kNumSMs), GT_THREADS, smem_bytes, st>>>(map, p);
  GDR_LAUNCHED();

  int32_t *fail_count = reinterpret_cast<int32_t *>(scal + 1);
  const int rr_smem = (int)(RR_WARPS * (d * 8 + lists * c * 12));
  GDR_CUDA(cudaFuncSetAttribute(gram_topk_rerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, rr_smem));
  gram_topk_rerank_kernel<<<(unsigned)ceil_div<int64_t>(n, RR_WARPS), 32 * RR_WARPS, rr_smem, st>>>(
      x, n, d, (int)k, lists, c, cand_idx, cand_thr, norms, scal, scale_exp, idx_out, val_out, fail_count, fail_rows);
  GDR_LAUNCHED();

  // every CTA beyond the (device-side) number of failed rows exits at once: no host synchronisation
  const int dyn = (int)((sizeof(double) + sizeof(int32_t)) * TKR_BM * (TKR_MAXK + TKR_QCAP));
  GDR_CUDA(cudaFuncSetAttribute(gram_topk_f64_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  gram_topk_f64_rows_kernel<<<(unsigned)ceil_div<int64_t>(n, TKR_BM), TKR_THREADS, dyn, st>>>(
      x, n, d, (int)k, fail_rows, fail_count, idx_out, val_out);
  GDR_LAUNCHED();
  if (n_fallback_out) GDR_CUDA(cudaMemcpyAsync(n_fallback_out, fail_count, 4, cudaMemcpyDeviceToDevice, st));
  return GDR_OK;
}

}  // extern "C"
