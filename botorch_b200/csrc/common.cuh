// Shared declarations for the botorch_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define BO_OK 0
#define BO_ERR_INVALID 1
#define BO_ERR_CUDA 2
#define BO_ERR_WORKSPACE 3
#define BO_ERR_UNSUPPORTED 4
#define BO_ERR_STATIC_SCRATCH 5

#define BO_MAX_Q 32   // largest q-batch (incl. pending points) the per-candidate kernels accept
#define BO_MAX_D 128  // largest input dimension

#define BO_KERNEL_RBF 0
#define BO_KERNEL_MATERN52 1
#define BO_REDUCE_LOG 0    // log_fatplus / fatmax / logmeanexp  (qLogNEI, qLogEI)
#define BO_REDUCE_PLAIN 1  // clamp_min(0) / amax / mean           (qNEI, qEI)
#define BO_REDUCE_LOG_SOFT 2  // log_softplus / smooth_amax / logmeanexp (fat = False)

#define BO_CUDA_TRY(expr)                                                                    \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            bo_set_last_cuda_error(_e, __FILE__, __LINE__);                                  \
            return BO_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

#define BO_TRY(expr)                 \
    do {                             \
        int _r = (expr);             \
        if (_r != BO_OK) return _r;  \
    } while (0)

void bo_set_last_cuda_error(cudaError_t e, const char* file, int line);

static inline long bo_round_up(long x, long m) { return (x + m - 1) / m * m; }

// ---------------------------------------------------------------------------------------------
// Batched FP64 GEMM on DMMA (gemm.cu).   C[z] = epilogue( A[z] (M x K) * B[z] (K x N) )
// All matrices row-major; leading dimensions must be even and base pointers 16-byte aligned.
// ---------------------------------------------------------------------------------------------
enum BoEpilogue : int {
    BO_EPI_LINEAR = 0,  // C = alpha*acc + beta*D
    BO_EPI_RBF = 1,     // acc is a squared distance: C = os[z]*exp(-0.5*max(acc,0))   (+ diag options)
    BO_EPI_MATERN52 = 2,// C = os[z]*(1+sqrt5 r+5/3 r^2) exp(-sqrt5 r), r = sqrt(max(acc,1e-30))
    BO_EPI_SINCOS = 3   // random Fourier features: C = alpha*sqrt(os[z])*sin(acc), C2 = alpha*sqrt(os[z])*cos(acc)
};

enum BoTri : int {
    BO_TRI_NONE = 0,
    BO_TRI_A_LOWER = 1,  // A[m][k] == 0 for k > m        -> k-loop stops at the row tile's last row
    BO_TRI_A_UPPER = 2,  // A[m][k] == 0 for k < m        -> k-loop starts at the row tile's first row
    BO_TRI_B_LOWER = 3   // B[k][n] == 0 for k < n        -> k-loop starts at the column tile's first col
};

struct BoGemm {
    const double* A; long lda; long strideA;
    const double* B; long ldb; long strideB;
    double* C; long ldc; long strideC;
    const double* D; long ldd; long strideD;   // optional addend (LINEAR only), may alias C
    double* CT; long ldct; long strideCT;      // optional transposed copy of C
    double alpha, beta;
    int M, N, K, batch;
    int tri;          // BoTri
    int tri_off;      // A_LOWER: A[m][k] == 0 for k > m + tri_off (tri_off dense leading columns);
                      // A_UPPER: rows < tri_off are dense, row m >= tri_off has A[m][k] == 0 for k < m - tri_off
    int out_lower;    // 1: skip output tiles strictly above the block diagonal (syrk-style)
    int epi;          // BoEpilogue
    const double* os;        // [batch] outputscale (kernel epilogues), may be null (=1)
    const double* diag_add;  // [batch] value added where row==col (noise), may be null
    int diag_exact;          // 1: squared distance forced to 0 where row==col (x1_eq_x2)
    int a_bcast, b_bcast;    // unused (strides of 0 already broadcast); kept for clarity
    int tile64;              // 1: force the 64 x 64 tile configuration (lets C alias A when N <= 64: a CTA then reads exactly
                             // the rows it writes, and all of its loads complete before its epilogue)
    const double* alpha_vec; // [batch] optional per-member factor multiplying alpha (LINEAR only)
    const double* E; long lde; long strideE;    // optional elementwise factor: C = (alpha*acc + beta*D) * E  (LINEAR only)
    double* C2; long ldc2; long strideC2;       // kernel epilogues: optional second output g(r), d k/dx_a = -g (x_a-x_b)/l^2
    // Fused per-column-group pieces of the combined-root qLogNEI path.  Honoured by the TMA 64 x 64 kernel only: set them
    // only when bo_gemm_fused_ok(g) says the launch will take it (LINEAR epilogue, alpha = 1, no D).
    //   gram (forward): for every 64-row tile rt and every group of 8 adjacent columns, the 8 x 8 Gram
    //     sum_r C[r][c1] C[r][c2] over the tile's rows r >= gram_row0, written to gram[z][rt][group][c1 * 8 + c2];
    //     summed over rt in order by the consumer: the q x q Gram U^T U of a t-batch without a second pass over U.
    //   post (backward): the k-range [post_k0, K) is accumulated FIRST, every 8-column group of that partial product is
    //     multiplied from the right by post[z][group] (8 x 8, row-major), then the k-range [0, post_k0) is added:
    //     A[:, k0:] (B[k0:, :] P) = (A[:, k0:] B[k0:, :]) P for block-diagonal P -- the adjoint U_bar = -2 U S_bar is
    //     never materialised.
    double* gram; long strideGram; int gram_row0;
    const double* post; long stridePost; int post_k0;
    //   hfuse (backward, together with post): the (d+1)-row contraction H = hA C that follows this GEMM on the gradient
    //     path is formed from the accumulators of every 64-row tile on the DMMA pipe and written as per-row-tile partials
    //     hpart[z][row tile][BO_HFUSE_ROWS][ldc] (rows >= hrows are zero); summed over the tiles in order by
    //     kgrad_finalize.  C itself is not stored when g.C is null: nobody reads G any more.
    const double* hA; long ldhA; int hrows;   // [hrows x ldhA], shared by the batch: H = hA[:, 0:M] C
    double* hpart; long strideHpart;
};
#define BO_HFUSE_ROWS 24  // padded row count of the fused gradient contraction (d + 1 <= 24)

#define BO_GRAM_BM 64  // row-tile height of the gram partials
// 1 if bo_gemm_launch(g) will run the kernel that honours g.gram / g.post
int bo_gemm_fused_ok(const BoGemm& g);

int bo_gemm_launch(const BoGemm& g, cudaStream_t stream);
// number of kernels launched by this library since the last reset (bench.py's gpu_launches claim)
extern long g_bo_launch_count;
