// Per-candidate (t-batch element) kernels of the qLogNEI path (cand.cu).
#pragma once
#include "common.cuh"

struct BoKernelHyper {
    int kernel_id, d, M;
    const double* inv_ls;       // [M x d]
    const double* outputscale;  // [M] or null
    const double* center;       // [d]
};

// Augmented distance operands for the DMMA contraction  r^2 = [-2x~, |x~|^2, 1] . [y~, 1, |y~|^2],
// x~ = (x - center)/lengthscale (gpytorch sq_dist formulation).
// rows: out[z][i][0..Kp)  (A operand, one row per point)
int bo_aug_rows(const BoKernelHyper& h, const double* X, long npts, double* out, long Kp, long strideOut,
                cudaStream_t stream);
// cols: out[z][k][c], pitch ldo (B operand, one column per point)
int bo_aug_cols(const BoKernelHyper& h, const double* X, long npts, double* out, long Kp, long ldo, long strideOut,
                cudaStream_t stream);

struct BoCandArgs {
    BoKernelHyper h;
    int n, n_b, S, B, q;
    double y_mean, y_std;
    const double* mean_const;  // [M]
    const double* X;           // [B x q x d]
    const double* Zq;          // [S x q]
    const double* best_f;      // [M][Sp]
    long Sp;
    double tau_relu, tau_max;
    long ldn;                  // pitch of every [rows x N] workspace matrix (N = B*q padded)
    // per-model strides (doubles) of the workspace matrices
    const double* V;  long strideV;    // [(n+1) x ldn]  rows 0..n-1 inv(L) K(X_tr, X); row n: alpha . K
    const double* BL; long strideBL;   // [n_b x ldn]    bl_chol^T  (combined-root path: U = L_W^T K(X_tr, X), n rows, V null)
    const double* mu_row; long strideMu; // optional [ldn] row holding alpha . K(X_tr, x) (else row n of V)
    double* F;        long strideF;    // [S x ldn]      in: Z_b BL ; out: samples f (fwd) / df (bwd)
    int f_zero;                        // 1: no baseline part (qEI / qLogEI): F is write-only scratch for the samples
    double obj_w;                      // objective = obj_w * f (LinearMCObjective on the single outcome); best_f lives in objective space
    int reduction;                     // BO_REDUCE_LOG: log_fatplus / fatmax / logmeanexp; BO_REDUCE_PLAIN: relu / amax / mean
    double* Lq;                        // [M][B][q*q]    br_chol
    double* lse;                       // [M][B]         logsumexp over MC samples
    double* acq;                       // [M][B]
    int32_t* info;                     // [M][B]
    // backward only
    const double* grad_acq;            // [M][B]
    double* Sbar;                      // [M][B][q*q]    adjoint of the q x q Schur block
    double* Vbar_mu; long strideVbar;  // row n of the adjoint of V (receives y_std * dmu)
    double* gradX;                     // [M][B][q][d]
    int df_external;                   // 1: F already holds d loss / d f (constraint-weighted path: the reductions ran outside)
    double* Uinplace; long strideU;    // optional (QT <= 8 kernels): [n_b x ldn] block that receives u_coef * U Sbar in place
    double u_coef;
    // fused forms (8 % q == 0; BoGemm::gram / BoGemm::post):
    const double* gram; long strideGram; int gram_tiles;  // forward: [row tile][ceil(N/8)][8 x 8] partial Grams of the BL block
                                                           // produced by the GEMM epilogue; BL itself is not read
    double schur_jitter; int schur_fixed;  // forward: one factorisation attempt of the q x q block with this jitter (bo_nei)
    // bo_nei::local_derivs: the forward leaves d qmax_s / d f[s][j] in the sample rows and qmax_s in msave [M][B][Sp]
    double* msave; int local_derivs;
    double* post; long stridePost;     // backward: [ceil(N/8)][8 x 8] <- u_coef * Sbar (block diagonal), instead of Uinplace
};

int bo_cand_fwd(const BoCandArgs& a, cudaStream_t stream);

// Sample / q reductions of SampleReducingMCAcquisitionFunction on a materialised sample matrix (monte_carlo.py:253-330):
// F [S x ldn] objective-output samples of B t-batches of q points, lw [S x ldn] optional per-sample log feasibility weights
// (added to the log improvement before the q-reduction: _apply_constraints with log=True), best_f [S].
struct BoReduceArgs {
    int S, B, q, reduction;
    long ldn;
    const double* F;
    const double* lw;        // or null
    const double* best_f;    // [S]
    double obj_w, tau_relu, tau_max;
    double* acq;             // [B]
    double* lse;             // [B]
    // backward only
    const double* grad_acq;  // [B]
    double* dF;              // [S x ldn]  d loss / d f
    double* lw_bar;          // [S x ldn]  d loss / d lw (or null)
};
int bo_mc_reduce_fwd_launch(const BoReduceArgs& a, cudaStream_t stream);
int bo_mc_reduce_bwd_launch(const BoReduceArgs& a, cudaStream_t stream);
int bo_cand_bwd(const BoCandArgs& a, cudaStream_t stream);

// Y[z][i][b*q+j] (+)= coef * sum_k Xm[z][i][b*q+k] * T[z][b][k][j]     (rows x (B*q) matrices, pitch ldn)
int bo_colblock_update(double* Y, long strideY, const double* Xm, long strideX, const double* T, int rows, int B,
                       int q, int M, long ldn, double coef, int accumulate, cudaStream_t stream);

// Y[z][i][b*q+:] <- coef * Y[z][i][b*q+:] T[z][b]   (in place)
int bo_colblock_inplace(double* Y, long strideY, const double* T, int rows, int B, int q, int M, long ldn, double coef,
                        cudaStream_t stream);

// last step of the GEMM-shaped kernel gradient: gradX[z][c][t] += scale/l_t^2 (H[t][c] - x_c[t] H[d][c])
// H holds `ksplit` partial sums (slice stride strideKs) that are added in slice order
int bo_kgrad_finalize(const BoKernelHyper& h, const double* H, long strideH, long ldn, const double* X, long N,
                      double scale, double* gradX, int ksplit, long strideKs, cudaStream_t stream);

// gradX[z][b][j][:] (+)= scale * sum_i Kbar[z][i][b*q+j] * d k(x_i, x_bj) / d x_bj
int bo_kgrad(const BoKernelHyper& h, const double* Xpts, int npts, const double* Kbar, long strideK, long ldn,
             const double* X, int B, int q, double scale, double* gradX, cudaStream_t stream);

// joint posterior blocks from V: mean[z][b][j], cov[z][b][j][k] = ys^2 (k(x_j,x_k) - sum_i V_ij V_ik)
int bo_cand_posterior(const BoCandArgs& a, double* mean, double* cov, cudaStream_t stream);

// reverse of bo_cand_posterior: Tout[z][b] = cov_bar + cov_bar^T, Vbar_mu row, gradX <- k(X,X) part
int bo_cand_posterior_bwd(const BoCandArgs& a, const double* mean_bar, const double* cov_bar, double* Tout,
                          cudaStream_t stream);

// marginal: mean[z][c] = y_mean + y_std*(m_z + V[n][c]); var[z][c] = ys^2 (os_z - sum_i V[i][c]^2)
int bo_col_mean_var(const double* V, long strideV, long ldn, int n, long Nc, int M, double y_mean, double y_std,
                    const double* mean_const, const double* outputscale, double* mean, double* var,
                    cudaStream_t stream);

// best_f[z][s] = max_i obj_w * (mean_b[z][i] + Fb[z][i][s])
int bo_col_max(const double* Fb, long strideF, long ldf, const double* mean_b, long strideMb, int rows, int cols,
               int M, double* out, long strideOut, double obj_w, cudaStream_t stream);

int bo_ts_argmax_launch(int M, long Nc, const double* mean, const double* var, const double* z, double* out_val,
                        long long* out_idx, cudaStream_t stream);
