// gemm_tc.cuh — host entry points of the tcgen05 3xTF32 GEMM (gemm_tc.cu) for the other translation units.
#pragma once
#include <cuda.h>  // CUtensorMap

#include "gemm_simt.cuh"

namespace gdr {

// hi = rn_tf32(x), lo = rn_tf32(x - hi) of [rows, k0 + k1] built from two row-major sources side by side
// (k1 = 0: one source).  k0, k1 multiples of 4.
int tc_split(const float *s0, int64_t ld0, int64_t k0, const float *s1, int64_t ld1, int64_t k1, int64_t rows,
             float *hi, float *lo, cudaStream_t st);
// the same for the transpose of s[rows, cols]: out[cols, ld_out], columns >= rows zero-filled
int tc_split_transpose(const float *s, int64_t ld, int64_t rows, int64_t cols, int64_t ld_out, float *hi, float *lo,
                       cudaStream_t st);
// C[m,n] = A[m,k]·B[n,k]ᵀ from pre-split operands; `epi` is a GemmEpi, `g` carries the epilogue operands
int tc_gemm(int epi, const float *ahi, const float *alo, const float *bhi, const float *blo, int64_t m, int64_t n,
            int64_t k, int64_t lda, int64_t ldb, int64_t k_per_split, int splits, const GemmParams &g,
            cudaStream_t st);

// the same with operand A as plain fp32, split inside the kernel: A = [a1 (k1 columns) | a2 (k2 columns)] along K
// (a2 == nullptr, k2 == 0: one source).  Epilogues: EPI_STORE, EPI_LAYER.  k1 % 32 == 0 when a2 is given.
int tc_gemm_raw_a(int epi, const float *a1, int64_t lda1, int64_t k1, const float *a2, int64_t lda2, int64_t k2,
                  const float *bhi, const float *blo, int64_t ldb, int64_t m, int64_t n, const GemmParams &g,
                  cudaStream_t st);

// raw-A product cut along K: part[z][m][n] over K ranges of k_per_split columns (multiple of 32); n % 4 == 0
int tc_gemm_raw_a_splitk(const float *a, int64_t lda, int64_t k, const float *bhi, const float *blo, int64_t ldb,
                         int64_t m, int64_t n, int64_t k_per_split, int splits, float *part, cudaStream_t st);

// TN form on raw operands: part[z][m][n1 + n2] = A[K range z, 0:m]ᵀ · [B1 | B2][K range z, :], A [k, m] and B1/B2
// [k, n1] / [k, n2] row-major (n1, n2 multiples of 128; n2 == 0: one source), K ranges of k_per_split rows
// (a multiple of 32).  colsum_part[z][m] (optional) receives the column sums of A over each K range.
int tc_gemm_tn_splitk(const float *a, int64_t lda, int64_t m, const float *b1, int64_t ldb1, int64_t n1,
                      const float *b2, int64_t ldb2, int64_t n2, int64_t k, int64_t k_per_split, int splits,
                      float *part, float *colsum_part, cudaStream_t st);

// TMA map of a K-major fp32 operand [rows, k] (leading dimension ld): boxes of 128 rows x 32 columns, SWIZZLE_128B,
// out-of-range elements read as zero
int tc_make_operand_map(CUtensorMap *map, const float *base, int64_t rows, int64_t k, int64_t ld);

// the buffer set by gdr_debug_tc_trace (nullptr: tracing off)
long long *tc_trace_buffer();

constexpr int TC_K_ALIGN = 32;  // split-K ranges must be multiples of the 32-column TMA box

}  // namespace gdr
