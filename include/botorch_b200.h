/* botorch_b200 -- C ABI of the B200-native qLogNEI / exact-GP posterior hot path.
 *
 * The reference (anand-12/botorch) is pure Python and has no FFI of its own:
 * the drop-in boundary is its Python class contracts, implemented in botorch_b200/*.py.  This header
 * is the C ABI those classes sit on -- every arithmetic step of the path runs behind these entry
 * points.  Each entry point names the reference code it replaces (paths under /root/reference).
 *
 * Conventions
 *   - all pointers are DEVICE pointers to float64 unless marked [host]; matrices are row-major;
 *   - the caller owns every buffer (state and workspace sizes come from the *_bytes queries);
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no hidden synchronisation
 *     and no allocation inside the library;
 *   - return value: BO_OK (0) or a BO_ERR_* code, text via bo_strerror(); numerical failures are
 *     reported per batch element through `info` arrays, never by exception.
 *   - M = number of models in the batch (1 for SingleTaskGP, #MCMC samples for SAAS).
 */
#ifndef BOTORCH_B200_H
#define BOTORCH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BO_OK 0
#define BO_ERR_INVALID 1
#define BO_ERR_CUDA 2
#define BO_ERR_WORKSPACE 3
#define BO_ERR_UNSUPPORTED 4
#define BO_ERR_STATIC_SCRATCH 5 /* problem larger than the built-in scratch of a *_static entry point */

#define BO_KERNEL_RBF 0      /* gpytorch RBFKernel (ARD), models/utils/gpytorch_modules.py:100-127 */
#define BO_KERNEL_MATERN52 1 /* gpytorch MaternKernel(nu=2.5, ARD), models/fully_bayesian.py:81-85,275-282 */

/* sample / q reductions of SampleReducingMCAcquisitionFunction (acquisition/monte_carlo.py:180-252) */
#define BO_REDUCE_LOG 0   /* log_fatplus, fatmax, logmeanexp: qLogNEI, qLogEI (acquisition/logei.py:73-135) */
#define BO_REDUCE_PLAIN 1 /* clamp_min(0), amax, mean: qNEI, qEI (acquisition/monte_carlo.py:333-414,417-560) */
#define BO_REDUCE_LOG_SOFT 2 /* fat = False: log_softplus, smooth_amax, logmeanexp (acquisition/logei.py:122,532; utils/safe_math.py:226-273) */

/* info codes written per (model, t-batch element) by bo_qlognei_fwd */
#define BO_INFO_OK 0
#define BO_INFO_NAN (-1)      /* NaN/Inf in samples: low_rank.py:163-172 NanError */
#define BO_INFO_NOT_PSD (-2)  /* q x q Schur block not PD after 6 jitter tries: NotPSDError */
/* info > 0: number of jitter retries psd_safe_cholesky needed (NumericalWarning) */

/* Fitted exact GP: inputs + factor buffers.  Replaces the state gpytorch's DefaultPredictionStrategy
 * caches on first use (mean_cache, covar_cache) behind models/gpytorch.py:446 `mvn = self(X)`. */
typedef struct bo_gp {
    int32_t kernel_id; /* BO_KERNEL_* */
    int32_t n, d, M;
    double y_mean, y_std;      /* Standardize buffers (models/transforms/outcome.py:244-247); 0/1 if none */
    /* inputs */
    const double* train_X;     /* [n x d] */
    const double* center;      /* [d]  centring point of the distance contraction (train_X.mean(0)) */
    const double* inv_ls;      /* [M x d] 1/lengthscale */
    const double* outputscale; /* [M] or NULL (=1) */
    const double* noise;       /* [M] */
    const double* mean;        /* [M] constant mean */
    const double* targets;     /* [n] standardised train targets */
    const double* jitter;      /* [M] extra diagonal jitter (psd_safe_cholesky retry) or NULL */
    /* state filled by bo_gp_setup (caller-allocated, sizes from bo_gp_state_doubles) */
    double* aug_train;         /* [M][n x Kp]        augmented distance operand, Kp = round_up(d+2,4) */
    double* linv_aug;          /* [M][(n+1) x np]    rows 0..n-1: inv(L); row n: alpha = (K+s2 I)^-1 (y-m) */
    double* linv_aug_t;        /* [M][n x n1p]       transpose of linv_aug */
    double* chol;              /* [M][n x np]        L, K + s2 I = L L^T */
    double* xt_aug;            /* [(d+1) x np]       rows 0..d-1: train_X^T, row d: ones (gradient contraction operand) */
    /* optional: fixed per-observation noise (SingleTaskGP with train_Yvar -> FixedNoiseGaussianLikelihood,
     * models/gp_regression.py:186-195): K + diag(noise_vec); `noise` is then ignored by the setup routines */
    const double* noise_vec;   /* [M x n] or NULL */
} bo_gp;

/* qLogNEI baseline state.  Replaces qLogNoisyExpectedImprovement._init_baseline
 * (acquisition/logei.py:364-420) and CachedCholeskyMCSamplerMixin._compute_root_decomposition
 * (acquisition/cached_cholesky.py:94-120). */
typedef struct bo_nei {
    int32_t n_b, S;
    int32_t baseline_is_train; /* 1: X_baseline is train_X (shares the cross-covariance) */
    const double* X_b;         /* [n_b x d] */
    const double* Z_b;         /* [S x n_b] base samples of the baseline points (sampling/normal.py:178-209) */
    const double* jitter;      /* [M] extra jitter for chol(Sigma_bb) or NULL */
    /* state filled by bo_nei_setup */
    double* aug_b;             /* [M][n_b x Kp]   (unused when baseline_is_train) */
    double* vb;                /* [M][n x nbp]    V_b = inv(L) K(X, X_b) */
    double* vb_t;              /* [M][n_b x np]   */
    double* lb;                /* [M][n_b x nbp]  L_b = chol(Sigma(X_b,X_b)) on the outcome scale (_baseline_L) */
    double* lb_inv;            /* [M][n_b x nbp]  */
    double* lb_inv_t;          /* [M][n_b x nbp]  */
    double* zb_pad;            /* [S x nbp]       Z_b with padded pitch */
    double* zb_t;              /* [n_b x Sp]      */
    double* mean_b;            /* [M][nbp]        posterior mean at X_b (outcome scale) */
    double* best_f;            /* [M][Sp]         max_j baseline sample (acquisition/utils.py:125-129) */
    /* only when baseline_is_train: Sigma(X_b, X) = ys^2 s2 (K+s2 I)^-1 K(X_tr, X), so bl_chol^T = Q K(X_tr, X) */
    double* qmat;              /* [M][n_b x np]   Q = ys^2 (s2+jitter) inv(L_b) (K+s2 I)^-1 */
    double* qmat_t;            /* [M][n x nbp]    Q^T */
    /* combined root: ys^2 (K+s2 I)^-1 + Q^T Q = L_W L_W^T, so the q x q Schur block of a candidate is
     * ys^2 k(x,x') - U^T U with U = L_W^T K(X_tr, X): one triangular GEMM replaces V, Sigma(X_b,X) and bl_chol */
    /* stacked operands, S1p = pad8(S+1): */
    double* fwd_cat;           /* [M][(S1p+n) x np]   rows 0..S-1: Z_b Q (baseline part of the samples = (Z_b Q) K(X_tr,X)); row S: alpha;
                                *                      rows S+1..S1p-1: 0; rows S1p+i: row i of L_W^T  -> one forward GEMM */
    double* bwd_cat;           /* [M][n x (S1p+np)]   [ fwd rows 0..S1p-1 transposed | L_W ]             -> one backward GEMM */
    /* the other members of the family on the same kernels */
    int32_t reduction;         /* BO_REDUCE_LOG (qLogNEI, qLogEI) or BO_REDUCE_PLAIN (qNEI, qEI) */
    int32_t no_baseline;       /* 1: qEI / qLogEI (acquisition/logei.py:138-234, monte_carlo.py:333-414): posterior over the q
                                *    new points only, best_f [M][Sp] supplied by the caller; n_b = 0, only S, best_f, fwd_cat,
                                *    bwd_cat are used, with S1p = 8: fwd_cat = [alpha; 0 x 7; L_W^T], L_W L_W^T = ys^2 (K+s2 I)^-1
                                *    (filled by bo_ei_setup) */
    double obj_weight;         /* objective = obj_weight * f (LinearMCObjective with one outcome, acquisition/objective.py);
                                * 0 means 1.  best_f (computed or supplied) is in objective space. */
    int32_t external_df;       /* bo_qlognei_bwd only.  1: the sample rows of the workspace (bo_qlognei_samples_view) already hold
                                * d loss / d f, written by the caller -- the reductions ran outside (bo_mc_reduce_*: outcome
                                * constraints couple the outputs of several models per sample); grad_acq is ignored. */
    /* The reference's fallback when the cached update fails (acquisition/cached_cholesky.py:139-165): standard sampling
     * factors the JOINT (n_b+q) covariance with psd_safe_cholesky, i.e. the SAME jitter delta on the baseline block and on
     * the q x q block: chol(Sigma_bb + delta I) (bo_nei.jitter at setup) and chol(Schur + delta I) here.
     * schur_fixed = 1: the q x q factorisation makes ONE attempt with schur_jitter on the diagonal (no escalation);
     * failure is reported as BO_INFO_NOT_PSD and the host moves on to the next delta of the schedule. */
    double schur_jitter;
    int32_t schur_fixed;
    /* local_derivs = 1 (log reductions only): a gradient-carrying bo_qlognei_fwd leaves in the sample rows of the workspace
     * NOT the samples f[s][j] but the per-sample local derivative d qmax_s / d f[s][j] (fatmax weight x d log_fatplus), and
     * the per-sample maxima next to them; bo_qlognei_bwd then only scales them by the logmeanexp weights instead of
     * re-evaluating log_fatplus / fatmax for every (sample, point).  Same numbers up to the association of one product.
     * Must be 0 when the caller reads the samples (bo_qlognei_samples_view: outcome constraints, generic objectives) and
     * must not change between a forward and its backward. */
    int32_t local_derivs;
    /* bo_nei_setup only.  0: everything.  1: every factor that does not read the base samples Z_b (cross covariance, the
     * baseline posterior covariance, its root and inverse, the combined root) -- the expensive part.  2: only what reads
     * Z_b (zb_pad / zb_t, the baseline samples and best_f, the Z_b Q rows of fwd_cat / columns of bwd_cat), on buffers that
     * a phase-1 call filled.  The host draws Z_b (a scrambled Sobol engine of dimension n_b: ~10 ms at n_b = 2000) while
     * the device runs phase 1.  1 then 2 leaves exactly what 0 leaves. */
    int32_t setup_phase;
} bo_nei;

typedef struct bo_ctx bo_ctx;

const char* bo_strerror(int code);
const char* bo_last_cuda_error(void);
int bo_ctx_create(int device, bo_ctx** out);
void bo_ctx_destroy(bo_ctx* ctx);
/* kernels launched by this library so far (bench.py's gpu_launches) */
long bo_launch_count(void);
void bo_launch_count_reset(void);

/* per-launch CUDA-event timing of the DMMA GEMM kernel (bench.py's roofline leg): enable, run, synchronise, read
 * total milliseconds / launches / algorithmic flops (2MNK, halved for triangular operands) since enable. */
void bo_prof_enable(int on); /* 0 off, 1 every GEMM launch, 2 only the TMA-staged kernel (the dominant one) */
int bo_prof_read(double* ms, long* launches, double* flops);
/* FP64 tensor roofline denominator measured in place: launches a DMMA.8x8x4 issue-rate microbenchmark (8 independent
 * accumulators per warp, 8 warps per CTA, 4 CTAs per SM) on `stream`; the caller times it with events.  scratch: one device
 * double; *flops [host] receives the flop count of the launch. */
int bo_dmma_peak_launch(int iters, double* scratch, double* flops, void* stream);
/* sha256 over the library's sources and compiler flags at build time (botorch_b200/build.py): lets a loader detect a stale
 * binary and lets bench.py tie an ncu capture to the code it was taken from. */
const char* bo_source_digest(void);

/* padded pitches used by every buffer of the ABI */
long bo_pad8(long x);
long bo_kp(int d);

/* ---- setup -------------------------------------------------------------------------------- */
/* stacked-operand geometry shared by setup, forward and backward: number of leading dense rows of fwd_cat */
long bo_nei_dense_rows(const bo_nei* nei);
size_t bo_gp_setup_workspace_bytes(const bo_gp* gp);
/* K = k(X,X) + (noise+jitter) I; L = chol(K) (blocked right-looking, DMMA trailing updates);
 * inv(L); alpha.  info[M] (device int32): 0 or 1-based index of the first bad pivot. */
int bo_gp_setup(bo_ctx* ctx, const bo_gp* gp, void* work, size_t work_bytes, int32_t* info, void* stream);

size_t bo_nei_setup_workspace_bytes(const bo_gp* gp, const bo_nei* nei);
/* V_b, Sigma_bb, L_b (+inverse), baseline samples and best_f. info[M] as above for chol(Sigma_bb). */
int bo_nei_setup(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, void* work, size_t work_bytes, int32_t* info,
                 void* stream);

/* qEI / qLogEI (nei->no_baseline = 1): fills fwd_cat = [alpha; 0; L_W^T] and bwd_cat = [alpha | 0 | L_W] with
 * L_W L_W^T = ys^2 (K + s2 I)^-1, so that the q x q posterior covariance of a candidate set is ys^2 k(X,X) - U^T U,
 * U = L_W^T K(X_tr, X) -- the joint posterior of model.posterior(X) (models/gpytorch.py:405-466) sampled as
 * mean + chol(Sigma) z (posteriors/gpytorch.py:121-165).  best_f [M][Sp] is an INPUT here.  info[M] as above. */
size_t bo_ei_setup_workspace_bytes(const bo_gp* gp, const bo_nei* nei);
int bo_ei_setup(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, void* work, size_t work_bytes, int32_t* info,
                void* stream);

/* ---- the hot path ------------------------------------------------------------------------- */
/* workspace for B t-batch elements of q points; with_grad != 0 adds the backward regions.
 * The workspace of a with_grad forward must be kept untouched until bo_qlognei_bwd ran. */
size_t bo_qlognei_workspace_bytes(const bo_gp* gp, const bo_nei* nei, int B, int q, int with_grad);

/* acq[M x B] = qLogNEI value per model and t-batch element.  Replaces
 * SampleReducingMCAcquisitionFunction.forward (acquisition/monte_carlo.py:253-289) with
 * qLogNoisyExpectedImprovement._get_samples_and_objectives (acquisition/logei.py:448-481),
 * model.posterior (models/gpytorch.py:405-466), sample_cached_cholesky (utils/low_rank.py:85-173),
 * _log_improvement/log_fatplus, fatmax, logmeanexp (utils/safe_math.py:293-352,209-223).
 * X [B x q x d]; Z_q [S x q] base samples of the q new points; info [M x B] int32. */
int bo_qlognei_fwd(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, int B, int q, const double* X,
                   const double* Z_q, double tau_relu, double tau_max, double* acq, int32_t* info, void* work,
                   size_t work_bytes, int with_grad, void* stream);

/* grad_X[M x B x q x d] = d(sum_b grad_acq[m,b] * acq[m,b]) / dX per model.  Replaces torch.autograd
 * through the whole forward (generation/gen.py:209-211). */
int bo_qlognei_bwd(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, int B, int q, const double* X,
                   const double* Z_q, double tau_relu, double tau_max, const double* grad_acq, double* grad_X,
                   void* work, size_t work_bytes, void* stream);

/* ---- outcome constraints across several single-output models --------------------------------------
 * SampleReducingMCAcquisitionFunction with `constraints` (acquisition/monte_carlo.py:253-330, utils/objective.py:134-180):
 * the feasibility weight of a sample couples the outputs of all models, so the per-sample work splits into
 *   1. bo_qlognei_fwd per model (with_grad = 1): the cached-Cholesky samples land in its workspace (bo_qlognei_samples_view);
 *   2. the caller forms log weights lw[s][c] = sum_k log sigmoid(-constraint_k(samples) / eta_k) from those samples;
 *   3. bo_mc_reduce_fwd on the objective model's samples: log improvement + lw -> fatmax over q -> logmeanexp over samples;
 *   4. bo_mc_reduce_bwd: d loss / d f of the objective output and d loss / d lw; the caller chains d lw through its
 *      constraints into per-model sample adjoints, writes them into the sample rows and calls bo_qlognei_bwd with
 *      nei->external_df = 1 for every model.
 * Offset (in doubles, from the workspace base) and pitch of the [S x pitch] sample rows of one model (M = 1). */
int bo_qlognei_samples_view(const bo_gp* gp, const bo_nei* nei, int B, int q, long* offset, long* pitch);
/* F, lw (or NULL), dF, lw_bar (or NULL): [S x ldn]; best_f [S]; acq, lse, grad_acq: [B].  reduction: BO_REDUCE_*;
 * with BO_REDUCE_PLAIN the weight multiplies the improvement (acqval.mul(ind), monte_carlo.py:328). */
int bo_mc_reduce_fwd(bo_ctx* ctx, int S, int B, int q, long ldn, const double* F, const double* lw, const double* best_f,
                     double obj_weight, double tau_relu, double tau_max, int reduction, double* acq, double* lse,
                     void* stream);
int bo_mc_reduce_bwd(bo_ctx* ctx, int S, int B, int q, long ldn, const double* F, const double* lw, const double* best_f,
                     double obj_weight, double tau_relu, double tau_max, int reduction, const double* lse,
                     const double* grad_acq, double* dF, double* lw_bar, void* stream);

/* ---- posterior API ------------------------------------------------------------------------- */
size_t bo_posterior_workspace_bytes(const bo_gp* gp, long Nc);
/* Joint posterior of B t-batches of q points: mean [M x B x q], cov [M x B x q x q] (outcome scale).
 * Replaces BatchedMultiOutputGPyTorchModel.posterior (models/gpytorch.py:405-466). */
int bo_posterior_joint(bo_ctx* ctx, const bo_gp* gp, int B, int q, const double* X, double* mean, double* cov,
                       void* work, size_t work_bytes, void* stream);
/* grad_X[M x B x q x d] = d(sum mean_bar*mean + sum cov_bar*cov)/dX per model: the reverse pass torch.autograd
 * runs through model.posterior when a caller differentiates posterior statistics (models/gpytorch.py:405-466
 * with detached train caches, models/utils/assorted.py:294-296). */
/* Joint posterior of ONE large point set (q may exceed BO_MAX_Q; e.g. all of X_baseline for
 * prune_inferior_points, acquisition/utils.py:245-349, or a joint Thompson draw over a candidate block,
 * generation/sampling.py:95-155): mean [M x q], cov [M][q x ldq] with ldq = bo_pad8(q), all GEMM-shaped. */
size_t bo_posterior_joint_large_workspace_bytes(const bo_gp* gp, long q);
int bo_posterior_joint_large(bo_ctx* ctx, const bo_gp* gp, long q, const double* X, double* mean, double* cov,
                             void* work, size_t work_bytes, void* stream);

size_t bo_posterior_bwd_workspace_bytes(const bo_gp* gp, int B, int q);
int bo_posterior_joint_bwd(bo_ctx* ctx, const bo_gp* gp, int B, int q, const double* X, const double* mean_bar,
                           const double* cov_bar, double* grad_X, void* work, size_t work_bytes, void* stream);
/* Marginal posterior at Nc points: mean, var [M x Nc].  The cross-covariance GEMM + triangular solve
 * of MaxPosteriorSampling at scale (generation/sampling.py:95-113, config C5). */
int bo_posterior_mean_var(bo_ctx* ctx, const bo_gp* gp, long Nc, const double* Xc, double* mean, double* var,
                          void* work, size_t work_bytes, void* stream);
/* argmax_i (mean_i + sqrt(var_i) * z_i) over Nc candidates per model: out_val[M], out_idx[M] (int64).
 * Thompson draw + maximize_samples (generation/sampling.py:115-117); first occurrence wins ties. */
int bo_ts_argmax(bo_ctx* ctx, int M, long Nc, const double* mean, const double* var, const double* z,
                 double* out_val, int64_t* out_idx, void* stream);
/* first-occurrence argmax of a vector (optim/initializers.py:944, optim/optimize.py:385) */
int bo_argmax(bo_ctx* ctx, long n, const double* v, double* out_val, int64_t* out_idx, void* stream);

/* ---- hyperparameter side of the caller ----------------------------------------------
 * Exact marginal log likelihood of the training data and its gradient w.r.t. the hyperparameters: what
 * fit_gpytorch_mll's closure computes per L-BFGS-B evaluation (botorch/fit.py:117-259 ->
 * optim/fit.py:47-110 -> optim/closures/model_closures.py:171-184: `-mll(model(*train_inputs), targets)` +
 * autograd), i.e. gpytorch ExactMarginalLogLikelihood -> MultivariateNormal.log_prob with exact Cholesky
 * (botorch/__init__.py:44-49).  Here:
 *   value[z] = log N(targets | mean_z, K_z + D_z)                (NOT divided by n, no prior terms: host adds those)
 *   grad[z]  = [ d/d lengthscale_0..d-1, d/d outputscale, d/d noise, d/d mean ]   ([M x (d+3)])
 * with W = alpha alpha^T - (K + D)^-1:  d/d l_t = 1/2 sum_ij W_ij g_ij (x_it - x_jt)^2 / l_t^3  (g: derivative factor of
 * the covariance function), evaluated as one triangular GEMM for W .* g, one (d+1)-row GEMM against [x~^T; 1] and
 * deterministic reductions.  The d/d noise entry is meaningful for homoskedastic noise only (noise_vec == NULL).
 * Runs bo_gp_setup first: on return the factor buffers of `gp` are the prediction caches of these hyperparameters.
 * info[M]: as bo_gp_setup (first bad pivot; value/grad of that model are then meaningless). */
size_t bo_mll_workspace_bytes(const bo_gp* gp);
int bo_mll_value_grad(bo_ctx* ctx, const bo_gp* gp, void* work, size_t work_bytes, double* value, double* grad,
                      int32_t* info, void* stream);

/* Potential energy of the reference's SAAS model in the unconstrained space pyro's NUTS works in, for M = gp->M chains at
 * once: what fit_fully_bayesian_model_nuts (botorch/fit.py:335-396) evaluates per leapfrog step through
 * SaasPyroModel.sample (models/fully_bayesian.py:167-246) + autograd.
 *   u [M x (d+4)] = [ log kernel_tausq, log _kernel_inv_length_sq (d), mean, log (noise - 1e-4), log outputscale ]
 *   energy[z] = -( log N(targets | mean, outputscale * matern52(X / lengthscale) + noise I) + log priors + log|dx/du| )
 *   with outputscale ~ Gamma(2, 0.15), mean ~ N(0, 1), noise - 1e-4 ~ Gamma(0.9, 10), tausq ~ HalfCauchy(0.1),
 *   inv_length_sq ~ HalfCauchy(1)^d, lengthscale = (tausq * inv_length_sq)^-1/2;   grad_u [M x (d+4)] = d energy / d u.
 * The call writes the hyperparameter buffers of `gp` (inv_ls, outputscale, noise, mean: must be writable, kernel
 * BO_KERNEL_MATERN52 with outputscale, no noise_vec), runs bo_mll_value_grad (value [M], grad [M x (d+3)] are its
 * outputs, kept for the caller) and finishes with the closed-form chain rule.  A chain with a non-finite / out-of-range u
 * (wild[z] = 1), a failed factorisation (info[z] != 0; no jitter is added) or a non-finite result gets energy = +inf and
 * a zero gradient -- a divergent transition, as pyro's registered exception handlers make it (fully_bayesian.py:67-78);
 * the other chains are unaffected.  Workspace: bo_mll_workspace_bytes. */
int bo_saas_potential(bo_ctx* ctx, const bo_gp* gp, const double* u, void* work, size_t work_bytes, double* value,
                      double* grad, int32_t* wild, double* energy, double* grad_u, int32_t* info, void* stream);

/* ---- pathwise (Matheron) posterior samples (second half) -----------------------------------------
 * Replaces draw_matheron_paths (sampling/pathwise/posterior_samplers.py:143-199) for the exact single-model GP:
 *   prior path   f0(x) = m + phi(x) . w,  phi(x) = sqrt(os) sqrt(2/F) [sin(W x/l), cos(W x/l)]      (random Fourier features:
 *                sampling/pathwise/features/generators.py:66-110, prior_samplers.py:51-91, utils.py:84-117)
 *   update       v = (K + D)^-1 (y - f0(X) - sqrt(D) eps),  path(x) = f0(x) + k(x, X) v          (update_strategies.py:72-103)
 *   output       y_mean + y_std * path(x)                                                       (Standardize.untransform)
 * The random draws are INPUTS (the host makes them with the reference's generators: draw_sobol_normal_samples for W and w,
 * randn for eps), so a path is a deterministic function of them.  Unlike the marginal form of bo_posterior_mean_var, the
 * S paths are JOINT draws over any number of candidates: Thompson sampling over 1e6 points costs one covariance GEMM
 * k(X, x) and two skinny GEMMs per candidate chunk, no n^2-per-candidate triangular product.  M = 1 models only
 * (the reference rejects fully Bayesian models, acquisition/thompson_sampling.py:47-50). */
typedef struct bo_paths {
    int32_t S, F;          /* sample paths; Fourier features (even): F/2 frequencies */
    const double* W;       /* [F/2 x d]  frequencies (standard normal for RBF; the host scales them for Matern) */
    const double* w;       /* [S x F]    feature weights */
    const double* eps;     /* [S x n]    standard-normal noise draws (scaled by sqrt(noise) inside) */
    /* state filled by bo_paths_setup (caller allocated) */
    double* w_aug;         /* [F/2 x Kp]     frequencies as a distance-contraction operand: [W, W.(center/l), 0] */
    double* w_pad;         /* [S x pad8(F)]  */
    double* v;             /* [S x pad8(n)]  update weights */
    double* w_t;           /* [F x pad8(S)]  w^T  (gradient path) */
    double* v_t;           /* [n x pad8(S)]  v^T  (gradient path) */
} bo_paths;
size_t bo_paths_setup_workspace_bytes(const bo_gp* gp, const bo_paths* paths);
/* needs the factors of bo_gp_setup */
int bo_paths_setup(bo_ctx* ctx, const bo_gp* gp, const bo_paths* paths, void* work, size_t work_bytes, void* stream);
size_t bo_paths_eval_workspace_bytes(const bo_gp* gp, const bo_paths* paths, long Nc);
/* out [S x pad8(Nc)]: every path at Nc points Xc [Nc x d] (outcome scale) */
int bo_paths_eval(bo_ctx* ctx, const bo_gp* gp, const bo_paths* paths, long Nc, const double* Xc, double* out, void* work,
                  size_t work_bytes, void* stream);
/* grad_X[Nc x d] = d(sum_{s,c} out_bar[s][c] * path_s(x_c)) / dX: the reverse pass autograd runs through a path when
 * PathwiseThompsonSampling is maximised with optimize_acqf (acquisition/thompson_sampling.py:60-87, generation/gen.py:211).
 * out_bar [S x pad8(Nc)].  Feature part: ((w^T out_bar) .* [cos; -sin]) contracted with W / l; data part: the same
 * fused-epilogue GEMM + [X^T; 1] contraction as the acquisition backward. */
size_t bo_paths_grad_workspace_bytes(const bo_gp* gp, const bo_paths* paths, long Nc);
int bo_paths_grad(bo_ctx* ctx, const bo_gp* gp, const bo_paths* paths, long Nc, const double* Xc, const double* out_bar,
                  double* grad_X, void* work, size_t work_bytes, void* stream);

/* ---- building blocks exported for the parity tests ------------------------------------------ */
int bo_dgemm(bo_ctx* ctx, int M, int N, int K, int batch, double alpha, const double* A, long lda, long strideA,
             const double* B, long ldb, long strideB, double beta, double* C, long ldc, long strideC, int tri,
             void* stream);
int bo_kernel_matrix(bo_ctx* ctx, const bo_gp* gp, long Nc, const double* Xc, double* out /* [M][n x pad8(Nc)] */,
                     void* work, size_t work_bytes, void* stream);
int bo_potrf(bo_ctx* ctx, int n, int batch, double* A, long lda, long strideA, double* inv /* or NULL */,
             void* work, size_t work_bytes, int32_t* info, void* stream);
size_t bo_potrf_workspace_bytes(int n, int batch);

/* bo_kernel_matrix without a workspace argument.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same kernels and bit-identical output to bo_kernel_matrix; the augmented candidate operand
 * [M][bo_kp(d) x bo_pad8(Nc)] lives in a file-scope device array of BO_STATIC_SCRATCH_DOUBLES doubles (512 KiB, held by
 * every GPU that loads the library until the process exits) instead of caller memory.  The call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine).  Limits and rules:
 *   - M * bo_kp(d) * bo_pad8(Nc) <= BO_STATIC_SCRATCH_DOUBLES, else BO_ERR_STATIC_SCRATCH (e.g. d = 20, M = 1: Nc <= 2728);
 *   - the device that owns `out` must be the current device (its copy of the array is used), else BO_ERR_INVALID;
 *   - every call on a device shares the one array: calls that overlap on the GPU (different streams, or graph branches
 *     without a dependency) RACE -- serialise them.  The array keeps the previous call's contents; the call overwrites all
 *     of the region it uses, so nothing needs clearing in between.
 * out [M][n x bo_pad8(Nc)] as for bo_kernel_matrix. */
#define BO_STATIC_SCRATCH_DOUBLES (64 * 1024)
int bo_kernel_matrix_static(bo_ctx* ctx, const bo_gp* gp, long Nc, const double* Xc, double* out, void* stream);

/* ---- device-resident multi-start optimiser ---------------------------------------
 * Replaces the SciPy L-BFGS-B round trip of gen_candidates_scipy (generation/gen.py:46-298) for the fused path: the
 * objective is a sum over restarts (gen.py:209), so every restart is an independent box-constrained problem of
 * D = q*d variables with its own limited-memory model.  State lives in caller-owned device buffers; a round is
 * bo_lbfgs_propose -> (caller evaluates f = -acq and g = -d acq / dX at x_trial) -> bo_lbfgs_accept; no host sync.
 * f is MINIMISED. */
typedef struct bo_lbfgs {
    int32_t B, D, m;           /* restarts, variables per restart, history length (<= 64) */
    double* x;                 /* [B x D] current iterate (inside the bounds) */
    double* g;                 /* [B x D] its gradient */
    double* f;                 /* [B]     its objective */
    double* x_prev;            /* [B x D] previous iterate / gradient (curvature pair) */
    double* g_prev;
    double* S;                 /* [B x m x D] steps */
    double* Y;                 /* [B x m x D] gradient differences */
    double* rho;               /* [B x m] */
    int32_t* hist_count;       /* [B] */
    int32_t* hist_head;        /* [B] */
    double* p;                 /* [B x D] search direction (zero on the active set) */
    double* alpha;             /* [B] step length of the next trial */
    double* x_trial;           /* [B x D] out: clip(x + alpha p) */
    int32_t* state;            /* [B] 0 new direction, 1 backtrack, 2 stopped, 3 restart from steepest descent */
    const double* lower;       /* [D] */
    const double* upper;       /* [D] */
    int32_t* n_iter;           /* [B] accepted steps */
} bo_lbfgs;

/* x must be set by the caller; f0 [B], g0 [B x D] are the objective and gradient at x */
int bo_lbfgs_init(bo_ctx* ctx, const bo_lbfgs* s, const double* f0, const double* g0, void* stream);
int bo_lbfgs_propose(bo_ctx* ctx, const bo_lbfgs* s, void* stream);
/* projected Armijo test f_trial <= f + c1 g.(x_trial - x); stops a restart when the projected gradient (sup norm) is
 * <= pgtol, the decrease is <= ftol * max(|f|, |f_trial|, 1), or maxiter accepted steps were taken */
int bo_lbfgs_accept(bo_ctx* ctx, const bo_lbfgs* s, const double* f_trial, const double* g_trial, double c1, double ftol,
                    double pgtol, int maxiter, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BOTORCH_B200_H */
