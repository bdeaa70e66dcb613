// Blocked dense factorisations built on the DMMA GEMM (linalg.cu).
#pragma once
#include "common.cuh"

#define BO_NB 64  // diagonal block size of the blocked Cholesky / triangular inverse

// Scratch (in doubles) the two routines below need for a batch of `batch` n x n matrices.
long bo_linalg_scratch_doubles(int n, int batch);

// In-place blocked right-looking Cholesky of the lower triangle of A[z] (n x n, row-major, lda even).
// On return A holds L (strict upper triangle zeroed).  If Xinv != nullptr its diagonal NB-blocks
// receive inv(L_kk) and the rest of Xinv is zeroed, ready for bo_trtri_blocked.
// info[z] (device int, must be zeroed by the caller) = 1-based index of the first non-positive pivot.
int bo_potrf_blocked(double* A, long lda, long strideA, int n, int batch, double* Xinv, long ldx, long strideX,
                     double* scratch, int* info, cudaStream_t stream);

// X = inv(L) for lower-triangular L, block-row forward substitution.  X's diagonal blocks must already
// hold inv(L_kk) (bo_potrf_blocked does that) and its strict upper part must be zero.
int bo_trtri_blocked(const double* L, long ldl, long strideL, int n, int batch, double* X, long ldx, long strideX,
                     double* scratch, cudaStream_t stream);

// out[z][j][i] = in[z][i][j]
int bo_transpose(const double* in, long ldi, long strideI, double* out, long ldo, long strideO, int rows, int cols,
                 int batch, cudaStream_t stream);
// dst[z][i][j] = src[z][i][j]
int bo_copy2d(const double* src, long lds, long strideS, double* dst, long ldd, long strideD, int rows, int cols,
              int batch, cudaStream_t stream);
// A[z][i][i] += add[z]  (add may be null -> no-op)
int bo_add_diag(double* A, long lda, long strideA, int n, int batch, const double* add, cudaStream_t stream);
// A[z][i][i] += vec[z][i]  (vec may be null -> no-op)
int bo_add_diag_vec(double* A, long lda, long strideA, int n, int batch, const double* vec, long strideVec,
                    cudaStream_t stream);
// dst[z][i][:] = (vec[z][i] + add[z]) * src[z][i][:]   (add may be null)
int bo_scale_rows(const double* src, double* dst, long ld, long stride, int rows, int cols, int batch, const double* vec,
                  long strideVec, const double* add, cudaStream_t stream);
int bo_fill(double* p, long count, double v, cudaStream_t stream);
