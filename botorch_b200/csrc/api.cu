// C ABI of libbotorch_b200 (declared in include/botorch_b200.h): orchestration of the DMMA GEMM
// pipeline and the per-candidate kernels.  No allocation, no synchronisation: every buffer is the
// caller's, every launch goes to the caller's stream.
#include "../../include/botorch_b200.h"

#include <cstdlib>
#include <cstring>

#include "cand.cuh"
#include "common.cuh"
#include "linalg.cuh"

struct bo_ctx {
    int device;
    int sms;
};

static char g_last_cuda_err[512] = "";

void bo_set_last_cuda_error(cudaError_t e, const char* file, int line) {
    snprintf(g_last_cuda_err, sizeof(g_last_cuda_err), "%s (%s) at %s:%d", cudaGetErrorName(e),
             cudaGetErrorString(e), file, line);
}

namespace {

struct Bump {  // carve the caller's workspace; 256-byte aligned regions
    char* base;
    size_t cap, off;
    bool ok;
    Bump(void* p, size_t c) : base((char*)p), cap(c), off(0), ok(true) {}
    double* doubles(long count) {
        size_t bytes = (size_t)bo_round_up((long)(count * sizeof(double)), 256);
        if (base != nullptr && off + bytes > cap) ok = false;
        double* r = base ? (double*)(base + off) : nullptr;
        off += bytes;
        return r;
    }
};

__global__ void vec_add_kernel(const double* a, const double* b, double* out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (a ? a[i] : 0.0) + (b ? b[i] : 0.0);
}
// r[z][i][0] = targets[i] - mean[z];  r[z][i][1] = 0
__global__ void resid_kernel(const double* targets, const double* mean, double* r, int n) {
    const int z = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        r[((long)z * n + i) * 2] = targets[i] - mean[z];
        r[((long)z * n + i) * 2 + 1] = 0.0;
    }
}
// alpha (stored as [n x 2] column) -> row n of linv_aug and column n of linv_aug_t
__global__ void scatter_alpha_kernel(const double* av, double* linv_aug, long np, double* linv_aug_t, long n1p, int n) {
    const int z = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const double v = av[((long)z * n + i) * 2];
        linv_aug[(long)z * (n + 1) * np + (long)n * np + i] = v;
        linv_aug_t[(long)z * n * n1p + (long)i * n1p + n] = v;
    }
}
// mean_b[z][i] = y_mean + y_std * (m_z + vb[z][n][i])
__global__ void mean_b_kernel(const double* vb, long strideVb, long nbp, int n, int n_b, double y_mean, double y_std,
                              const double* mean, double* out) {
    const int z = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_b) out[(long)z * nbp + i] = y_mean + y_std * (mean[z] + vb[(long)z * strideVb + (long)n * nbp + i]);
}

BoKernelHyper hyper_of(const bo_gp* gp) {
    BoKernelHyper h;
    h.kernel_id = gp->kernel_id;
    h.d = gp->d;
    h.M = gp->M;
    h.inv_ls = gp->inv_ls;
    h.outputscale = gp->outputscale;
    h.center = gp->center;
    return h;
}

// covariance matrix from augmented operands: C[z] = k(A_aug[z] . B_aug[z])
int kmat(const bo_gp* gp, const double* A_aug, int rows, const double* B_aug, long ldb, long cols, double* C, long ldc,
         long strideC, int diag_exact, const double* diag_add, cudaStream_t st, double* C2 = nullptr) {
    const long Kp = bo_kp(gp->d);
    BoGemm g{};
    g.A = A_aug; g.lda = Kp; g.strideA = (long)rows * Kp;
    g.B = B_aug; g.ldb = ldb; g.strideB = Kp * ldb;
    g.C = C; g.ldc = ldc; g.strideC = strideC;
    g.M = rows; g.N = (int)cols; g.K = (int)Kp; g.batch = gp->M;
    g.epi = gp->kernel_id == BO_KERNEL_RBF ? BO_EPI_RBF : BO_EPI_MATERN52;
    g.os = gp->outputscale; g.diag_add = diag_add; g.diag_exact = diag_exact;
    g.C2 = C2; g.ldc2 = ldc; g.strideC2 = strideC;
    return bo_gemm_launch(g, st);
}

int gemm(int M, int N, int K, int batch, double alpha, const double* A, long lda, long sA, const double* B, long ldb,
         long sB, double beta, const double* D, long ldd, long sD, double* C, long ldc, long sC, int tri, int out_lower,
         cudaStream_t st) {
    BoGemm g{};
    g.A = A; g.lda = lda; g.strideA = sA;
    g.B = B; g.ldb = ldb; g.strideB = sB;
    g.C = C; g.ldc = ldc; g.strideC = sC;
    g.D = D; g.ldd = ldd; g.strideD = sD;
    g.alpha = alpha; g.beta = beta; g.M = M; g.N = N; g.K = K; g.batch = batch;
    g.tri = tri; g.out_lower = out_lower; g.epi = BO_EPI_LINEAR;
    return bo_gemm_launch(g, st);
}

bool valid_gp(const bo_gp* gp) {
    return gp && gp->n > 0 && gp->d > 0 && gp->d <= BO_MAX_D && gp->M > 0 && (gp->noise || gp->noise_vec) &&
           (gp->kernel_id == BO_KERNEL_RBF || gp->kernel_id == BO_KERNEL_MATERN52);
}

}  // namespace

extern "C" {

const char* bo_strerror(int code) {
    switch (code) {
        case BO_OK: return "ok";
        case BO_ERR_INVALID: return "invalid argument (shape, alignment or null pointer)";
        case BO_ERR_CUDA: return "CUDA runtime error (see bo_last_cuda_error)";
        case BO_ERR_WORKSPACE: return "workspace too small";
        case BO_ERR_UNSUPPORTED: return "unsupported size (q > 32, d > 128 or grid limit)";
        case BO_ERR_STATIC_SCRATCH:
            return "problem exceeds the built-in scratch of the *_static entry point (M * kp(d) * pad8(Nc) > "
                   "BO_STATIC_SCRATCH_DOUBLES = 65536 doubles); use the workspace-taking entry point";
        default: return "unknown error";
    }
}
const char* bo_last_cuda_error(void) { return g_last_cuda_err; }
#ifndef BO_SOURCE_DIGEST
#define BO_SOURCE_DIGEST "unknown"
#endif
const char* bo_source_digest(void) { return BO_SOURCE_DIGEST; }

int bo_ctx_create(int device, bo_ctx** out) {
    if (!out) return BO_ERR_INVALID;
    int count = 0;
    BO_CUDA_TRY(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) return BO_ERR_INVALID;
    cudaDeviceProp p;
    BO_CUDA_TRY(cudaGetDeviceProperties(&p, device));
    if (p.major != 10) return BO_ERR_UNSUPPORTED;  // built for sm_100a only
    bo_ctx* c = new bo_ctx;
    c->device = device;
    c->sms = p.multiProcessorCount;
    *out = c;
    return BO_OK;
}
void bo_ctx_destroy(bo_ctx* ctx) { delete ctx; }
long bo_launch_count(void) { return g_bo_launch_count; }
void bo_launch_count_reset(void) { g_bo_launch_count = 0; }
long bo_pad8(long x) { return bo_round_up(x, 8); }
long bo_kp(int d) { return bo_round_up(d + 2, 4); }

// ------------------------------------------------------------------------------------------------
// setup
// ------------------------------------------------------------------------------------------------
static size_t gp_setup_layout(const bo_gp* gp, void* work, size_t cap, double** baug, double** scratch, double** rv,
                              double** t1, double** av, double** nz, bool* ok) {
    const long n = gp->n, M = gp->M, np = bo_pad8(n), Kp = bo_kp(gp->d);
    Bump b(work, cap);
    *baug = b.doubles(M * Kp * np);
    *scratch = b.doubles(bo_linalg_scratch_doubles((int)n, (int)M));
    *rv = b.doubles(M * n * 2);
    *t1 = b.doubles(M * n * 2);
    *av = b.doubles(M * n * 2);
    *nz = b.doubles(M);
    *ok = b.ok;
    return b.off;
}

size_t bo_gp_setup_workspace_bytes(const bo_gp* gp) {
    double *a, *b, *c, *d, *e, *f;
    bool ok;
    return gp_setup_layout(gp, nullptr, 0, &a, &b, &c, &d, &e, &f, &ok);
}

int bo_gp_setup(bo_ctx* ctx, const bo_gp* gp, void* work, size_t work_bytes, int32_t* info, void* stream) {
    if (!ctx || !valid_gp(gp) || !work || !info) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const long n = gp->n, M = gp->M, np = bo_pad8(n), n1p = bo_pad8(n + 1), Kp = bo_kp(gp->d);
    double *baug, *scratch, *rv, *t1, *av, *nz;
    bool ok;
    gp_setup_layout(gp, work, work_bytes, &baug, &scratch, &rv, &t1, &av, &nz, &ok);
    if (!ok) return BO_ERR_WORKSPACE;
    const BoKernelHyper h = hyper_of(gp);
    BO_CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int32_t) * M, st));
    BO_TRY(bo_aug_rows(h, gp->train_X, n, gp->aug_train, Kp, n * Kp, st));
    BO_TRY(bo_aug_cols(h, gp->train_X, n, baug, Kp, np, Kp * np, st));
    if (gp->xt_aug) {  // [train_X^T; 1]: operand of the gradient contraction (bo_qlognei_bwd step 6)
        BO_TRY(bo_fill(gp->xt_aug, (long)(gp->d + 1) * np, 0.0, st));
        BO_TRY(bo_transpose(gp->train_X, gp->d, 0, gp->xt_aug, np, 0, (int)n, gp->d, 1, st));
        BO_TRY(bo_fill(gp->xt_aug + (long)gp->d * np, n, 1.0, st));
    }
    vec_add_kernel<<<(unsigned)((M + 127) / 128), 128, 0, st>>>(gp->noise_vec ? nullptr : gp->noise, gp->jitter, nz, (int)M);
    g_bo_launch_count++;
    // K + (s2 + jitter) I, unit diagonal of the kernel enforced exactly (x1_eq_x2 branch of gpytorch's sq_dist);
    // with fixed per-observation noise the diagonal gets noise_vec[i] + jitter instead
    BO_TRY(kmat(gp, gp->aug_train, (int)n, baug, np, n, gp->chol, np, n * np, 1, nz, st));
    BO_TRY(bo_add_diag_vec(gp->chol, np, n * np, (int)n, (int)M, gp->noise_vec, n, st));
    BO_TRY(bo_potrf_blocked(gp->chol, np, n * np, (int)n, (int)M, gp->linv_aug, np, (n + 1) * np, scratch, info, st));
    BO_TRY(bo_trtri_blocked(gp->chol, np, n * np, (int)n, (int)M, gp->linv_aug, np, (n + 1) * np, scratch, st));
    BO_TRY(bo_fill(gp->linv_aug_t, M * n * n1p, 0.0, st));
    BO_TRY(bo_transpose(gp->linv_aug, np, (n + 1) * np, gp->linv_aug_t, n1p, n * n1p, (int)n, (int)n, (int)M, st));
    // alpha = inv(L)^T (inv(L) (y - m))
    dim3 vgrid((unsigned)((n + 127) / 128), (unsigned)M);
    resid_kernel<<<vgrid, 128, 0, st>>>(gp->targets, gp->mean, rv, (int)n);
    g_bo_launch_count++;
    BO_TRY(gemm((int)n, 1, (int)n, (int)M, 1.0, gp->linv_aug, np, (n + 1) * np, rv, 2, n * 2, 0.0, nullptr, 0, 0, t1, 2,
                n * 2, BO_TRI_A_LOWER, 0, st));
    BO_TRY(gemm((int)n, 1, (int)n, (int)M, 1.0, gp->linv_aug_t, n1p, n * n1p, t1, 2, n * 2, 0.0, nullptr, 0, 0, av, 2,
                n * 2, BO_TRI_A_UPPER, 0, st));
    scatter_alpha_kernel<<<vgrid, 128, 0, st>>>(av, gp->linv_aug, np, gp->linv_aug_t, n1p, (int)n);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

// lw holds W = A^T A (lower block-triangle, pitch ldk) for the stacked root A = [ys inv(L); Q] (Q [nb x n] may be null);
// on return lw = L_W with L_W L_W^T = W to the accuracy of a QR of A.
// W has the SQUARED condition number of A, and so does the error of its Cholesky factor (measured: 5e-7 relative on the
// Schur block of a densely sampled low-d problem).  Second pass of Cholesky-QR ("CholeskyQR2"): A1 = A inv(R1) is nearly
// orthonormal, R2 = chol(A1^T A1)^T, R = R2 R1, which is as accurate as a Householder QR of A for cond(A) < 1e8.
// L_W = R^T = L_1 L_G.  qr: six n x n regions per model.
static int cholqr2_root(const bo_gp* gp, const double* qmat, long nb, double* lw, long ldk, double* qr, double* scratch,
                        int32_t* info, cudaStream_t st) {
    const long n = gp->n, M = gp->M, np = bo_pad8(n);
    const long sq = n * np;  // one n x n region per model
    double *x1 = qr, *x1t = qr + M * sq, *a1 = qr + 2 * M * sq, *a2 = qr + 3 * M * sq, *at = qr + 4 * M * sq,
           *gm = qr + 5 * M * sq;
    BO_TRY(bo_fill(qr, 6 * M * sq, 0.0, st));
    BO_TRY(bo_potrf_blocked(lw, ldk, n * ldk, (int)n, (int)M, x1, np, sq, scratch, info, st));
    BO_TRY(bo_trtri_blocked(lw, ldk, n * ldk, (int)n, (int)M, x1, np, sq, scratch, st));  // x1 = inv(L_1)
    BO_TRY(bo_transpose(x1, np, sq, x1t, np, sq, (int)n, (int)n, (int)M, st));                 // inv(R1)
    BO_TRY(gemm((int)n, (int)n, (int)n, (int)M, gp->y_std, gp->linv_aug, np, (n + 1) * np, x1t, np, sq, 0.0, nullptr, 0, 0,
                a1, np, sq, BO_TRI_A_LOWER, 0, st));                                            // ys inv(L) inv(R1)
    BO_TRY(bo_transpose(a1, np, sq, at, np, sq, (int)n, (int)n, (int)M, st));
    BO_TRY(gemm((int)n, (int)n, (int)n, (int)M, 1.0, at, np, sq, a1, np, sq, 0.0, nullptr, 0, 0, gm, np, sq, BO_TRI_NONE, 1,
                st));
    if (qmat) {
        BO_TRY(gemm((int)nb, (int)n, (int)n, (int)M, 1.0, qmat, np, nb * np, x1t, np, sq, 0.0, nullptr, 0, 0, a2, np, sq,
                    BO_TRI_NONE, 0, st));                                                       // Q inv(R1)
        BO_TRY(bo_transpose(a2, np, sq, at, np, sq, (int)nb, (int)n, (int)M, st));
        BO_TRY(gemm((int)n, (int)n, (int)nb, (int)M, 1.0, at, np, sq, a2, np, sq, 1.0, gm, np, sq, gm, np, sq, BO_TRI_NONE, 1,
                    st));                                                                       // G = A1^T A1
    }
    BO_TRY(bo_potrf_blocked(gm, np, sq, (int)n, (int)M, nullptr, np, sq, scratch, info, st));      // L_G
    // x1 (inv(L_1), no longer needed) has an all-zero strict upper part: the lower-tile product lands there
    BO_TRY(gemm((int)n, (int)n, (int)n, (int)M, 1.0, lw, ldk, n * ldk, gm, np, sq, 0.0, nullptr, 0, 0, x1, np, sq,
                BO_TRI_A_LOWER, 1, st));                                                        // L_W = L_1 L_G
    return bo_copy2d(x1, np, sq, lw, ldk, n * ldk, (int)n, (int)n, (int)M, st);
}

static size_t nei_setup_layout(const bo_gp* gp, const bo_nei* nei, void* work, size_t cap, double** baug_b,
                               double** kb, double** fb, double** scratch, double** nz, double** khinv, double** khs,
                               double** qr, bool* ok) {
    const long n = gp->n, M = gp->M, nb = nei->n_b, nbp = bo_pad8(nb), Sp = bo_pad8(nei->S), Kp = bo_kp(gp->d);
    Bump b(work, cap);
    *baug_b = b.doubles(M * Kp * nbp);
    *kb = b.doubles(M * n * nbp);
    *fb = b.doubles(M * nb * Sp);
    *scratch = b.doubles(bo_linalg_scratch_doubles((int)nb, (int)M));
    *nz = b.doubles(M);
    *khinv = nei->baseline_is_train ? b.doubles(M * n * bo_pad8(n)) : nullptr;  // (K + s2 I)^-1
    *khs = (nei->baseline_is_train && gp->noise_vec) ? b.doubles(M * n * bo_pad8(n)) : nullptr;  // D (K + D)^-1
    // six n x n temporaries of the second Cholesky-QR pass over the stacked root [ys inv(L); Q]
    *qr = (nei->baseline_is_train && nei->fwd_cat && nei->bwd_cat) ? b.doubles(6 * M * n * bo_pad8(n)) : nullptr;
    *ok = b.ok;
    return b.off;
}

size_t bo_nei_setup_workspace_bytes(const bo_gp* gp, const bo_nei* nei) {
    double *a, *b, *c, *d, *e, *f, *g, *h;
    bool ok;
    return nei_setup_layout(gp, nei, nullptr, 0, &a, &b, &c, &d, &e, &f, &g, &h, &ok);
}

int bo_nei_setup(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, void* work, size_t work_bytes, int32_t* info,
                 void* stream) {
    if (!ctx || !valid_gp(gp) || !nei || nei->n_b <= 0 || nei->S <= 0 || !work || !info) return BO_ERR_INVALID;
    if (nei->baseline_is_train && nei->n_b != gp->n) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const long n = gp->n, M = gp->M, np = bo_pad8(n), nb = nei->n_b, nbp = bo_pad8(nb), S = nei->S, Sp = bo_pad8(S);
    const long Kp = bo_kp(gp->d);
    const double ys2 = gp->y_std * gp->y_std;
    double *baug_b, *kb, *fb, *scratch, *nz, *khinv, *khs, *qr;
    bool ok;
    nei_setup_layout(gp, nei, work, work_bytes, &baug_b, &kb, &fb, &scratch, &nz, &khinv, &khs, &qr, &ok);
    if (!ok) return BO_ERR_WORKSPACE;
    const BoKernelHyper h = hyper_of(gp);
    // bo_nei::setup_phase: 1 = the factors that do not read Z_b, 2 = only what reads Z_b, 0 = both (same launches, same order
    // within each part, so 1 followed by 2 leaves the same bits as 0)
    const bool do_factors = nei->setup_phase != 2, do_samples = nei->setup_phase != 1;
    if (do_samples && !nei->Z_b) return BO_ERR_INVALID;
    if (do_factors) BO_CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int32_t) * M, st));
    if (do_samples) {  // base samples with padded pitch, and transposed
        BO_TRY(bo_fill(nei->zb_pad, S * nbp, 0.0, st));
        BO_TRY(bo_copy2d(nei->Z_b, nb, 0, nei->zb_pad, nbp, 0, (int)S, (int)nb, 1, st));
        BO_TRY(bo_fill(nei->zb_t, nb * Sp, 0.0, st));
        BO_TRY(bo_transpose(nei->zb_pad, nbp, 0, nei->zb_t, Sp, 0, (int)S, (int)nb, 1, st));
    }
    if (do_factors) {
        // cross covariance K(X_tr, X_b) and V_b = inv(L) K, mean row
        const double* rows_b = gp->aug_train;
        if (!nei->baseline_is_train) {
            BO_TRY(bo_aug_rows(h, nei->X_b, nb, nei->aug_b, Kp, nb * Kp, st));
            rows_b = nei->aug_b;
        }
        BO_TRY(bo_aug_cols(h, nei->X_b, nb, baug_b, Kp, nbp, Kp * nbp, st));
        BO_TRY(kmat(gp, gp->aug_train, (int)n, baug_b, nbp, nb, kb, nbp, n * nbp, 0, nullptr, st));
        BO_TRY(gemm((int)n + 1, (int)nb, (int)n, (int)M, 1.0, gp->linv_aug, np, (n + 1) * np, kb, nbp, n * nbp, 0.0, nullptr,
                    0, 0, nei->vb, nbp, (n + 1) * nbp, BO_TRI_A_LOWER, 0, st));
        BO_TRY(bo_fill(nei->vb_t, M * nb * np, 0.0, st));
        BO_TRY(bo_transpose(nei->vb, nbp, (n + 1) * nbp, nei->vb_t, np, nb * np, (int)n, (int)nb, (int)M, st));
        dim3 mgrid((unsigned)((nb + 127) / 128), (unsigned)M);
        mean_b_kernel<<<mgrid, 128, 0, st>>>(nei->vb, (n + 1) * nbp, nbp, (int)n, (int)nb, gp->y_mean, gp->y_std, gp->mean,
                                             nei->mean_b);
        g_bo_launch_count++;
        // Sigma_bb = ys^2 (K(X_b,X_b) - V_b^T V_b), lower block-triangle, then its Cholesky root and inverse
        BO_TRY(kmat(gp, rows_b, (int)nb, baug_b, nbp, nb, nei->lb, nbp, nb * nbp, 1, nullptr, st));
        BO_TRY(gemm((int)nb, (int)nb, (int)n, (int)M, -ys2, nei->vb_t, np, nb * np, nei->vb, nbp, (n + 1) * nbp, ys2, nei->lb,
                    nbp, nb * nbp, nei->lb, nbp, nb * nbp, BO_TRI_NONE, 1, st));
        BO_TRY(bo_add_diag(nei->lb, nbp, nb * nbp, (int)nb, (int)M, nei->jitter, st));
        BO_TRY(bo_potrf_blocked(nei->lb, nbp, nb * nbp, (int)nb, (int)M, nei->lb_inv, nbp, nb * nbp, scratch, info, st));
        BO_TRY(bo_trtri_blocked(nei->lb, nbp, nb * nbp, (int)nb, (int)M, nei->lb_inv, nbp, nb * nbp, scratch, st));
        BO_TRY(bo_fill(nei->lb_inv_t, M * nb * nbp, 0.0, st));
        BO_TRY(bo_transpose(nei->lb_inv, nbp, nb * nbp, nei->lb_inv_t, nbp, nb * nbp, (int)nb, (int)nb, (int)M, st));
    }
    if (do_samples) {  // baseline samples f_b[s] = mean_b + L_b z_s and best_f[s] = max_i f_b[s][i]
        BO_TRY(gemm((int)nb, (int)S, (int)nb, (int)M, 1.0, nei->lb, nbp, nb * nbp, nei->zb_t, Sp, 0, 0.0, nullptr, 0, 0, fb, Sp,
                    nb * Sp, BO_TRI_A_LOWER, 0, st));
        BO_TRY(bo_col_max(fb, nb * Sp, Sp, nei->mean_b, nbp, (int)nb, (int)S, (int)M, nei->best_f, Sp,
                          nei->obj_weight != 0.0 ? nei->obj_weight : 1.0, st));
    }
    if (nei->baseline_is_train && nei->qmat && nei->qmat_t) {
        // X_baseline == train_X: Sigma(X_b, X) = ys^2 (K - K Khat^-1 K_x) = ys^2 s2 Khat^-1 K_x with Khat = K + s2 I,
        // so bl_chol^T = inv(L_b) Sigma(X_b, X) = Q K(X_tr, X),  Q = ys^2 s2 inv(L_b) Khat^-1  (no cancellation).
        if (do_factors) {
            vec_add_kernel<<<(unsigned)((M + 127) / 128), 128, 0, st>>>(gp->noise, gp->jitter, nz, (int)M);
            g_bo_launch_count++;
            const long n1p = bo_pad8(n + 1);
            BO_TRY(gemm((int)n, (int)n, (int)n, (int)M, 1.0, gp->linv_aug_t, n1p, n * n1p, gp->linv_aug, np, (n + 1) * np, 0.0,
                        nullptr, 0, 0, khinv, np, n * np, BO_TRI_A_UPPER, 0, st));
            // Khat - K = D (s2 I, or diag(noise_vec) for fixed per-observation noise): Sigma(X_b, X) = ys^2 D Khat^-1 K_x
            const double* qrhs = khinv;
            if (gp->noise_vec) {
                BO_TRY(bo_scale_rows(khinv, khs, np, n * np, (int)n, (int)n, (int)M, gp->noise_vec, n, gp->jitter, st));
                qrhs = khs;
            }
            BoGemm g{};
            g.A = nei->lb_inv; g.lda = nbp; g.strideA = nb * nbp;
            g.B = qrhs; g.ldb = np; g.strideB = n * np;
            g.C = nei->qmat; g.ldc = np; g.strideC = nb * np;
            g.CT = nei->qmat_t; g.ldct = nbp; g.strideCT = n * nbp;
            g.alpha = ys2; g.alpha_vec = gp->noise_vec ? nullptr : nz; g.M = (int)nb; g.N = (int)n; g.K = (int)nb; g.batch = (int)M;
            g.tri = BO_TRI_A_LOWER; g.epi = BO_EPI_LINEAR;
            BO_TRY(bo_fill(nei->qmat_t, M * n * nbp, 0.0, st));
            BO_TRY(bo_gemm_launch(g, st));
        }
        if (nei->fwd_cat && nei->bwd_cat) {
            // W = ys^2 Khat^-1 + Q^T Q = L_W L_W^T, factored in place inside the backward operand [ . | L_W ]
            const long S1 = S + 1, S1p = bo_pad8(S1), ldk = S1p + np, rowsF = S1p + n;
            double* lw = nei->bwd_cat + S1p;
            if (do_factors) {
                BO_TRY(bo_fill(nei->bwd_cat, M * n * ldk, 0.0, st));
                BO_TRY(bo_fill(nei->fwd_cat, M * rowsF * np, 0.0, st));
                BO_TRY(gemm((int)n, (int)n, (int)nb, (int)M, 1.0, nei->qmat_t, nbp, n * nbp, nei->qmat, np, nb * np, ys2, khinv, np,
                            n * np, lw, ldk, n * ldk, BO_TRI_NONE, 1, st));
                BO_TRY(cholqr2_root(gp, nei->qmat, nb, lw, ldk, qr, scratch, info, st));
                BO_TRY(bo_transpose(lw, ldk, n * ldk, nei->fwd_cat + S1p * np, np, rowsF * np, (int)n, (int)n, (int)M, st));
                // row S: alpha
                BO_TRY(bo_copy2d(gp->linv_aug + n * np, np, (n + 1) * np, nei->fwd_cat + S * np, np, rowsF * np, 1, (int)n,
                                 (int)M, st));
            }
            if (do_samples) {
                // rows 0..S-1: Z_b Q; then rows 0..S transposed into the leading columns of the backward operand
                BO_TRY(gemm((int)S, (int)n, (int)nb, (int)M, 1.0, nei->zb_pad, nbp, 0, nei->qmat, np, nb * np, 0.0, nullptr, 0, 0,
                            nei->fwd_cat, np, rowsF * np, BO_TRI_NONE, 0, st));
                BO_TRY(bo_transpose(nei->fwd_cat, np, rowsF * np, nei->bwd_cat, ldk, n * ldk, (int)S1, (int)n, (int)M, st));
            }
        }
    }
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

long bo_nei_dense_rows(const bo_nei* nei) { return bo_pad8((nei->no_baseline ? 0 : (long)nei->S) + 1); }

static size_t ei_setup_layout(const bo_gp* gp, void* work, size_t cap, double** scratch, double** qr, bool* ok) {
    Bump b(work, cap);
    *scratch = b.doubles(bo_linalg_scratch_doubles(gp->n, gp->M));
    *qr = b.doubles(6 * (long)gp->M * gp->n * bo_pad8(gp->n));
    *ok = b.ok;
    return b.off;
}

size_t bo_ei_setup_workspace_bytes(const bo_gp* gp, const bo_nei* nei) {
    (void)nei;
    double *a, *b;
    bool ok;
    return ei_setup_layout(gp, nullptr, 0, &a, &b, &ok);
}

int bo_ei_setup(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, void* work, size_t work_bytes, int32_t* info,
                void* stream) {
    if (!ctx || !valid_gp(gp) || !nei || !nei->no_baseline || nei->S <= 0 || !nei->fwd_cat || !nei->bwd_cat ||
        !nei->best_f || !work || !info)
        return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const long n = gp->n, M = gp->M, np = bo_pad8(n), n1p = bo_pad8(n + 1);
    const long S1p = bo_nei_dense_rows(nei), ldk = S1p + np, rowsF = S1p + n;
    const double ys2 = gp->y_std * gp->y_std;
    double *scratch, *qr;
    bool ok;
    ei_setup_layout(gp, work, work_bytes, &scratch, &qr, &ok);
    if (!ok) return BO_ERR_WORKSPACE;
    BO_CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int32_t) * M, st));
    double* lw = nei->bwd_cat + S1p;
    BO_TRY(bo_fill(nei->bwd_cat, M * n * ldk, 0.0, st));
    BO_TRY(bo_fill(nei->fwd_cat, M * rowsF * np, 0.0, st));
    // W = ys^2 (K + s2 I)^-1 = ys^2 L^-T L^-1 (lower block-triangle), factored in place inside the backward operand
    BO_TRY(gemm((int)n, (int)n, (int)n, (int)M, ys2, gp->linv_aug_t, n1p, n * n1p, gp->linv_aug, np, (n + 1) * np, 0.0,
                nullptr, 0, 0, lw, ldk, n * ldk, BO_TRI_A_UPPER, 1, st));
    BO_TRY(cholqr2_root(gp, nullptr, 0, lw, ldk, qr, scratch, info, st));
    BO_TRY(bo_transpose(lw, ldk, n * ldk, nei->fwd_cat + S1p * np, np, rowsF * np, (int)n, (int)n, (int)M, st));
    // row 0: alpha, and its transpose as column 0 of the backward operand
    BO_TRY(bo_copy2d(gp->linv_aug + n * np, np, (n + 1) * np, nei->fwd_cat, np, rowsF * np, 1, (int)n, (int)M, st));
    BO_TRY(bo_transpose(nei->fwd_cat, np, rowsF * np, nei->bwd_cat, ldk, n * ldk, 1, (int)n, (int)M, st));
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// the hot path
// ------------------------------------------------------------------------------------------------
namespace {

struct QWork {
    long ldn, R, regionStride;
    double *baug, *W0, *W1, *W2, *W3, *F, *W5, *W6, *Lq, *lse, *Sbar, *Wd, *H;
    double *gram, *post;           // fused combined-root pieces (BoGemm::gram / BoGemm::post), 8 % q == 0 only
    double* msave;                 // per-sample maxima of the forward [M][B][Sp] (bo_nei::local_derivs), with_grad only
    long strideGram, stridePost;
    int gram_tiles;
    long strideH;
    int ksplit;  // row slices of the [train_X^T; 1] G contraction (split-K), summed in order by the finalize kernel
    bool qpath;  // X_baseline == train_X with the combined root L_W (bo_nei.lw): Schur = ys^2 k - U^T U, U = L_W^T K(X_tr, X)
    int32_t* info;
    bool ok;
    size_t bytes;
};

// The (d+1) x N x n contraction H = [train_X^T; 1] G has one row tile only: split its n-long sum into slices that run as
// separate batch members (single-model case), so the launch fills the SMs; slices must be equal and 16-byte aligned.
int hsplit(const bo_gp* gp) {
    if (gp->M != 1 || gp->n < 512) return 1;
    for (int ks = 8; ks >= 2; ks--)
        if (gp->n % ks == 0 && (gp->n / ks) % 2 == 0) return ks;
    return 1;
}

// H slices <- [train_X^T; 1] G, then gradX += (H_t - x_t H_d) / l_t^2
int kernel_gradient(const bo_gp* gp, const BoKernelHyper& h, const QWork& w, const double* G, const double* X, long N,
                    double* grad_X, cudaStream_t st);

// BO_FUSE=0 disables the fused Gram / adjoint pieces of the combined-root path (A/B measurements)
static bool fuse_enabled() {
    static const bool on = [] { const char* e = getenv("BO_FUSE"); return !e || atoi(e) != 0; }();
    return on;
}

static bool fuse_enabled_h() {  // BO_FUSE_H=0: keep the separate (d+1)-row gradient GEMM (A/B measurements)
    static const bool on = [] { const char* e = getenv("BO_FUSE_H"); return !e || atoi(e) != 0; }();
    return on;
}

QWork qlayout(const bo_gp* gp, const bo_nei* nei, int B, int q, int with_grad, void* work, size_t cap) {
    QWork w{};
    const long n = gp->n, nb = nei->n_b, M = gp->M, Kp = bo_kp(gp->d);
    const long N = (long)B * q;
    w.ldn = bo_pad8(N);
    w.qpath = (nei->baseline_is_train || nei->no_baseline) && nei->fwd_cat && nei->bwd_cat && gp->xt_aug;
    const long S1p = bo_nei_dense_rows(nei);
    // rows of every [rows x N] region: the stacked forward output [samples; pad; U] on the shared-baseline path
    w.R = w.qpath ? S1p + n : (n > nb ? n : nb) + 1;
    w.regionStride = w.R * w.ldn;
    Bump b(work, cap);
    w.baug = b.doubles(M * Kp * w.ldn);
    w.W0 = b.doubles(M * w.regionStride);
    w.W1 = b.doubles(M * w.regionStride);
    w.Lq = b.doubles(M * (long)B * q * q);
    w.lse = b.doubles(M * (long)B);
    w.info = reinterpret_cast<int32_t*>(b.doubles((M * (long)B + 1) / 2));
    if (!w.qpath) {
        w.W2 = b.doubles(M * w.regionStride);
        w.W3 = b.doubles(M * w.regionStride);
        w.F = b.doubles(M * (long)(nei->S + 1) * w.ldn);
        w.W6 = nei->baseline_is_train ? nullptr : b.doubles(M * w.regionStride);
    } else if (nei->no_baseline) {
        w.F = b.doubles(M * (long)nei->S * w.ldn);  // the samples (no baseline rows in the stacked output)
    }
    if (with_grad) {
        w.Sbar = b.doubles(M * (long)B * q * q);
        w.strideH = (long)(gp->d + 1) * w.ldn;
        w.ksplit = hsplit(gp);
        w.H = gp->xt_aug ? b.doubles(M * w.strideH * w.ksplit) : nullptr;
        if (gp->kernel_id != BO_KERNEL_RBF) w.Wd = b.doubles(M * w.regionStride);  // Matern derivative factor
        if (w.qpath) w.W2 = b.doubles(M * w.regionStride);
        else w.W5 = b.doubles(M * w.regionStride);
    }
    if (w.qpath && q <= 8 && 8 % q == 0 && fuse_enabled()) {  // appended last: the offsets of every other region stay what they were
        const long groups = (N + 7) / 8;
        w.gram_tiles = (int)((S1p + n + BO_GRAM_BM - 1) / BO_GRAM_BM);
        w.strideGram = (long)w.gram_tiles * groups * 64;
        w.stridePost = groups * 64;
        w.gram = b.doubles(M * w.strideGram);
        if (with_grad) w.post = b.doubles(M * w.stridePost);
    }
    if (with_grad) w.msave = b.doubles(M * (long)B * bo_pad8(nei->S));  // appended last as well
    w.ok = b.ok;
    w.bytes = b.off;
    return w;
}

int kernel_gradient(const bo_gp* gp, const BoKernelHyper& h, const QWork& w, const double* G, const double* X, long N,
                    double* grad_X, cudaStream_t st) {
    const long n = gp->n, M = gp->M, np = bo_pad8(n);
    if (w.ksplit > 1) {  // M == 1: batch index = slice
        const long kc = n / w.ksplit;
        BO_TRY(gemm(gp->d + 1, (int)N, (int)kc, w.ksplit, 1.0, gp->xt_aug, np, kc, G, w.ldn, kc * w.ldn, 0.0, nullptr, 0, 0,
                    w.H, w.ldn, w.strideH, BO_TRI_NONE, 0, st));
        return bo_kgrad_finalize(h, w.H, 0, w.ldn, X, N, 1.0, grad_X, w.ksplit, w.strideH, st);
    }
    BO_TRY(gemm(gp->d + 1, (int)N, (int)n, (int)M, 1.0, gp->xt_aug, np, 0, G, w.ldn, w.regionStride, 0.0, nullptr, 0, 0, w.H,
                w.ldn, w.strideH, BO_TRI_NONE, 0, st));
    return bo_kgrad_finalize(h, w.H, w.strideH, w.ldn, X, N, 1.0, grad_X, 1, 0, st);
}

BoCandArgs cand_args(const bo_gp* gp, const bo_nei* nei, int B, int q, const double* X, const double* Zq,
                     double tau_relu, double tau_max, const QWork& w) {
    BoCandArgs a{};
    a.h = hyper_of(gp);
    a.n = gp->n; a.n_b = nei ? nei->n_b : 0; a.S = nei ? nei->S : 0; a.B = B; a.q = q;
    a.y_mean = gp->y_mean; a.y_std = gp->y_std;
    a.mean_const = gp->mean;
    a.X = X; a.Zq = Zq;
    a.best_f = nei ? nei->best_f : nullptr;
    a.Sp = nei ? bo_pad8(nei->S) : 0;
    a.tau_relu = tau_relu; a.tau_max = tau_max;
    a.reduction = nei ? nei->reduction : BO_REDUCE_LOG;
    a.obj_w = (nei && nei->obj_weight != 0.0) ? nei->obj_weight : 1.0;
    a.schur_jitter = nei ? nei->schur_jitter : 0.0;
    a.schur_fixed = nei ? nei->schur_fixed : 0;
    // local derivatives only where the reductions are the log ones and nobody else writes the sample rows
    a.local_derivs = (nei && nei->local_derivs && nei->reduction != BO_REDUCE_PLAIN && !nei->external_df && w.msave && q <= 8) ? 1 : 0;
    a.msave = w.msave;
    a.ldn = w.ldn;
    a.V = w.W1; a.strideV = w.regionStride;
    a.BL = w.W3; a.strideBL = w.regionStride;
    a.F = w.F; a.strideF = nei ? (long)(nei->S + 1) * w.ldn : 0;
    a.Lq = w.Lq; a.lse = w.lse;
    return a;
}

}  // namespace

extern "C" {

size_t bo_qlognei_workspace_bytes(const bo_gp* gp, const bo_nei* nei, int B, int q, int with_grad) {
    return qlayout(gp, nei, B, q, with_grad, nullptr, 0).bytes;
}

int bo_qlognei_fwd(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, int B, int q, const double* X,
                   const double* Z_q, double tau_relu, double tau_max, double* acq, int32_t* info, void* work,
                   size_t work_bytes, int with_grad, void* stream) {
    if (!ctx || !valid_gp(gp) || !nei || B <= 0 || q <= 0 || !X || !Z_q || !acq || !info || !work)
        return BO_ERR_INVALID;
    if (q > BO_MAX_Q) return BO_ERR_UNSUPPORTED;
    if (!(tau_relu > 0.0) || !(tau_max > 0.0)) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const QWork w = qlayout(gp, nei, B, q, with_grad, work, work_bytes);
    if (!w.ok) return BO_ERR_WORKSPACE;
    if (nei->no_baseline && !w.qpath) return BO_ERR_INVALID;
    const long n = gp->n, nb = nei->n_b, M = gp->M, np = bo_pad8(n), nbp = bo_pad8(nb), Kp = bo_kp(gp->d);
    const long S = nei->S, N = (long)B * q;
    const double ys2 = gp->y_std * gp->y_std;
    const BoKernelHyper h = hyper_of(gp);
    // 1. K(X_tr, X) (and K(X_b, X)) through the augmented squared-distance contraction
    BO_TRY(bo_aug_cols(h, X, N, w.baug, Kp, w.ldn, Kp * w.ldn, st));
    BO_TRY(kmat(gp, gp->aug_train, (int)n, w.baug, w.ldn, N, w.W0, w.ldn, w.regionStride, 0, nullptr, st,
                with_grad ? w.Wd : nullptr));
    const double* KB = w.W0;
    if (!nei->baseline_is_train) {
        BO_TRY(kmat(gp, nei->aug_b, (int)nb, w.baug, w.ldn, N, w.W6, w.ldn, w.regionStride, 0, nullptr, st));
        KB = w.W6;
    }
    if (w.qpath) {
        // 2r. [Z_b Q; alpha^T; 0; L_W^T] K(X_tr, X) in ONE launch: rows 0..S-1 baseline part of every sample, row S the mean,
        //     rows S1p.. U = L_W^T K(X_tr, X) (upper-triangular part of the operand: k-range skipped per row tile)
        const long S1p = bo_nei_dense_rows(nei), Sd = nei->no_baseline ? 0 : S;  // Sd sample rows, then the mean row
        BoGemm g{};
        g.A = nei->fwd_cat; g.lda = np; g.strideA = (S1p + n) * np;
        g.B = w.W0; g.ldb = w.ldn; g.strideB = w.regionStride;
        g.C = w.W1; g.ldc = w.ldn; g.strideC = w.regionStride;
        g.alpha = 1.0; g.M = (int)(S1p + n); g.N = (int)N; g.K = (int)n; g.batch = (int)M;
        g.tri = BO_TRI_A_UPPER; g.tri_off = (int)S1p; g.epi = BO_EPI_LINEAR;
        // q x q Gram U^T U of every t-batch from the epilogue of this launch (no second pass over the n x N block U)
        const bool fuse_gram = w.gram != nullptr && bo_gemm_fused_ok(g);
        if (fuse_gram) { g.gram = w.gram; g.strideGram = w.strideGram; g.gram_row0 = (int)S1p; }
        BO_TRY(bo_gemm_launch(g, st));
        BoCandArgs a = cand_args(gp, nei, B, q, X, Z_q, tau_relu, tau_max, w);
        if (fuse_gram) { a.gram = w.gram; a.strideGram = w.strideGram; a.gram_tiles = w.gram_tiles; }
        a.V = nullptr;
        if (nei->no_baseline) { a.F = w.F; a.strideF = S * w.ldn; a.f_zero = 1; }
        else { a.F = w.W1; a.strideF = w.regionStride; }
        a.BL = w.W1 + S1p * w.ldn; a.strideBL = w.regionStride; a.n_b = (int)n;
        a.mu_row = w.W1 + Sd * w.ldn; a.strideMu = w.regionStride;
        a.acq = acq;
        a.info = w.info;
        BO_TRY(bo_cand_fwd(a, st));
        BO_CUDA_TRY(cudaMemcpyAsync(info, w.info, sizeof(int32_t) * M * B, cudaMemcpyDeviceToDevice, st));
        return BO_OK;
    }
    // 2. V = [inv(L); alpha^T] K(X_tr, X)           (n+1) x N
    BO_TRY(gemm((int)n + 1, (int)N, (int)n, (int)M, 1.0, gp->linv_aug, np, (n + 1) * np, w.W0, w.ldn, w.regionStride, 0.0,
                nullptr, 0, 0, w.W1, w.ldn, w.regionStride, BO_TRI_A_LOWER, 0, st));
    // 3. Sigma(X_b, X) = ys^2 (K(X_b, X) - V_b^T V)   n_b x N
    BO_TRY(gemm((int)nb, (int)N, (int)n, (int)M, -ys2, nei->vb_t, np, nb * np, w.W1, w.ldn, w.regionStride, ys2, KB, w.ldn,
                w.regionStride, w.W2, w.ldn, w.regionStride, BO_TRI_NONE, 0, st));
    // 4. bl_chol^T = inv(L_b) Sigma(X_b, X)
    BO_TRY(gemm((int)nb, (int)N, (int)nb, (int)M, 1.0, nei->lb_inv, nbp, nb * nbp, w.W2, w.ldn, w.regionStride, 0.0,
                nullptr, 0, 0, w.W3, w.ldn, w.regionStride, BO_TRI_A_LOWER, 0, st));
    // 5. baseline part of the samples: Z_b bl_chol^T    S x N
    BO_TRY(gemm((int)S, (int)N, (int)nb, (int)M, 1.0, nei->zb_pad, nbp, 0, w.W3, w.ldn, w.regionStride, 0.0, nullptr, 0, 0,
                w.F, w.ldn, (S + 1) * w.ldn, BO_TRI_NONE, 0, st));
    // 6. Schur block, q x q Cholesky, samples, log_fatplus / fatmax / logmeanexp
    BoCandArgs a = cand_args(gp, nei, B, q, X, Z_q, tau_relu, tau_max, w);
    a.acq = acq;
    a.info = w.info;  // kept in the workspace for the backward pass, copied out for the caller
    BO_TRY(bo_cand_fwd(a, st));
    BO_CUDA_TRY(cudaMemcpyAsync(info, w.info, sizeof(int32_t) * M * B, cudaMemcpyDeviceToDevice, st));
    return BO_OK;
}

int bo_qlognei_bwd(bo_ctx* ctx, const bo_gp* gp, const bo_nei* nei, int B, int q, const double* X,
                   const double* Z_q, double tau_relu, double tau_max, const double* grad_acq, double* grad_X,
                   void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || !nei || B <= 0 || q <= 0 || !X || !Z_q || (!grad_acq && !nei->external_df) || !grad_X || !work)
        return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const QWork w = qlayout(gp, nei, B, q, 1, work, work_bytes);
    if (!w.ok) return BO_ERR_WORKSPACE;
    if (nei->no_baseline && !w.qpath) return BO_ERR_INVALID;
    const long n = gp->n, nb = nei->n_b, M = gp->M, n1p = bo_pad8(n + 1), nbp = bo_pad8(nb), Sp = bo_pad8(nei->S);
    const long S = nei->S, N = (long)B * q;
    const double ys2 = gp->y_std * gp->y_std;
    const BoKernelHyper h = hyper_of(gp);
    // 1. reductions, samples and q x q Cholesky in reverse: F <- dF, Sbar, y_std * mubar, gradX <- k(X,X) part
    BoCandArgs a = cand_args(gp, nei, B, q, X, Z_q, tau_relu, tau_max, w);
    a.info = w.info;
    a.grad_acq = grad_acq;
    a.df_external = nei->external_df;
    a.Sbar = w.Sbar;
    a.gradX = grad_X;
    if (w.qpath) {
        const long np = bo_pad8(n), S1p = bo_nei_dense_rows(nei), ldk = S1p + np, Sd = nei->no_baseline ? 0 : S;
        if (nei->no_baseline) { a.F = w.F; a.strideF = S * w.ldn; }  // samples / dF in their own scratch
        else { a.F = w.W1; a.strideF = w.regionStride; }  // samples live in rows 0..S-1 of the stacked forward output
        a.Vbar_mu = w.W1 + Sd * w.ldn;            // row Sd <- y_std * mubar: rows 0..Sd become [dF; y_std mubar]
        a.strideVbar = w.regionStride;
        // 2r. Ubar = -2 U Sbar, in place on rows S1p.. of the same region: inside the reverse kernel for q <= 8 (the
        //     q x q block lives in registers there), as a separate pass otherwise
        // 3r. G = ([Z_b Q; alpha^T]^T [dF; y_std mubar] + L_W Ubar) .* g(r) in ONE launch over the stacked operand;
        //     g = K for RBF, the stored derivative factor for Matern                                          (W2)
        BoGemm g{};
        g.A = nei->bwd_cat; g.lda = ldk; g.strideA = n * ldk;
        g.B = w.W1; g.ldb = w.ldn; g.strideB = w.regionStride;
        g.C = w.W2; g.ldc = w.ldn; g.strideC = w.regionStride;
        g.E = w.Wd ? w.Wd : w.W0; g.lde = w.ldn; g.strideE = w.regionStride;
        g.alpha = 1.0; g.M = (int)n; g.N = (int)N; g.K = (int)(S1p + n); g.batch = (int)M;
        g.tri = BO_TRI_A_LOWER; g.tri_off = (int)S1p; g.epi = BO_EPI_LINEAR;
        // L_W (U (-2 Sbar)) = (L_W U) (-2 Sbar): the block-diagonal factor is applied to the accumulators inside the
        // GEMM (BoGemm::post), so U stays as the forward left it and nobody rewrites the n x N block
        const bool fuse_post = w.post != nullptr && bo_gemm_fused_ok(g);
        const bool fuse_u = !fuse_post && q <= 8;
        if (fuse_post) { a.post = w.post; a.stridePost = w.stridePost; a.u_coef = -2.0; }
        if (fuse_u) { a.Uinplace = w.W1 + S1p * w.ldn; a.strideU = w.regionStride; a.u_coef = -2.0; a.n_b = (int)n; }
        BO_TRY(bo_cand_bwd(a, st));
        if (!fuse_u && !fuse_post)
            BO_TRY(bo_colblock_inplace(w.W1 + S1p * w.ldn, w.regionStride, w.Sbar, (int)n, B, q, (int)M, w.ldn, -2.0, st));
        if (fuse_post) { g.post = w.post; g.stridePost = w.stridePost; g.post_k0 = (int)S1p; }
        // 4r. H = [train_X^T; 1] G  ((d+1) x N), then gradX += (H_t - x_t H_d) / l_t^2.  With the fused adjoint the
        //     contraction rides on the GEMM's accumulators (BoGemm::hfuse): per-row-tile partials of H land in the region
        //     that used to hold G, G itself is never written, and the separate (d+1)-row GEMM disappears.
        const int tiles_m = (int)((n + BO_GRAM_BM - 1) / BO_GRAM_BM);
        const bool fuse_h = fuse_post && gp->xt_aug != nullptr && gp->d + 1 <= BO_HFUSE_ROWS &&
                            (long)tiles_m * BO_HFUSE_ROWS <= n && fuse_enabled_h();
        if (fuse_h) {
            g.hA = gp->xt_aug; g.ldhA = np; g.hrows = gp->d + 1;
            g.hpart = w.W2; g.strideHpart = w.regionStride;
            g.C = nullptr;
            BO_TRY(bo_gemm_launch(g, st));
            return bo_kgrad_finalize(h, w.W2, w.regionStride, w.ldn, X, N, 1.0, grad_X, tiles_m,
                                     (long)BO_HFUSE_ROWS * w.ldn, st);
        }
        BO_TRY(bo_gemm_launch(g, st));
        return kernel_gradient(gp, h, w, w.W2, X, N, grad_X, st);
    }
    a.Vbar_mu = w.W5 + n * w.ldn;
    a.strideVbar = w.regionStride;
    BO_TRY(bo_cand_bwd(a, st));
    // 2. BLbar = Z_b^T dF - 2 BL Sbar                       (W2)
    BO_TRY(gemm((int)nb, (int)N, (int)S, (int)M, 1.0, nei->zb_t, Sp, 0, w.F, w.ldn, (S + 1) * w.ldn, 0.0, nullptr, 0, 0, w.W2,
                w.ldn, w.regionStride, BO_TRI_NONE, 0, st));
    BO_TRY(bo_colblock_update(w.W2, w.regionStride, w.W3, w.regionStride, w.Sbar, (int)nb, B, q, (int)M, w.ldn, -2.0, 1, st));
    // 3. Cbar = inv(L_b)^T BLbar                            (W3: bl_chol^T is dead after step 2; W0 keeps K(X_tr, X))
    BO_TRY(gemm((int)nb, (int)N, (int)nb, (int)M, 1.0, nei->lb_inv_t, nbp, nb * nbp, w.W2, w.ldn, w.regionStride, 0.0,
                nullptr, 0, 0, w.W3, w.ldn, w.regionStride, BO_TRI_A_UPPER, 0, st));
    // 4. Vbar = -ys^2 V_b Cbar - 2 ys^2 V Sbar              (W5 rows 0..n-1; row n holds y_std * mubar)
    BO_TRY(gemm((int)n, (int)N, (int)nb, (int)M, -ys2, nei->vb, nbp, (n + 1) * nbp, w.W3, w.ldn, w.regionStride, 0.0,
                nullptr, 0, 0, w.W5, w.ldn, w.regionStride, BO_TRI_NONE, 0, st));
    BO_TRY(bo_colblock_update(w.W5, w.regionStride, w.W1, w.regionStride, w.Sbar, (int)n, B, q, (int)M, w.ldn, -2.0 * ys2,
                              1, st));
    // 5. KXbar = [inv(L)^T | alpha] Vbar (+ ys^2 Cbar when X_baseline is train_X), and with the gradient contraction
    //    operand available G = KXbar .* g(r) straight from the epilogue                                    (W2)
    const bool shared = nei->baseline_is_train != 0;
    const bool gemm_grad = w.H != nullptr;
    {
        BoGemm g{};
        g.A = gp->linv_aug_t; g.lda = n1p; g.strideA = n * n1p;
        g.B = w.W5; g.ldb = w.ldn; g.strideB = w.regionStride;
        g.C = w.W2; g.ldc = w.ldn; g.strideC = w.regionStride;
        if (shared) { g.D = w.W3; g.ldd = w.ldn; g.strideD = w.regionStride; g.beta = ys2; }
        if (gemm_grad) { g.E = w.Wd ? w.Wd : w.W0; g.lde = w.ldn; g.strideE = w.regionStride; }
        g.alpha = 1.0; g.M = (int)n; g.N = (int)N; g.K = (int)n + 1; g.batch = (int)M;
        g.tri = BO_TRI_A_UPPER; g.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(g, st));
    }
    // 6. chain through the covariance function to X: H = [train_X^T; 1] G, gradX += (H_t - x_t H_d) / l_t^2
    if (gemm_grad) {
        BO_TRY(kernel_gradient(gp, h, w, w.W2, X, N, grad_X, st));
    } else {
        BO_TRY(bo_kgrad(h, gp->train_X, (int)n, w.W2, w.regionStride, w.ldn, X, B, q, 1.0, grad_X, st));
    }
    if (!shared) BO_TRY(bo_kgrad(h, nei->X_b, (int)nb, w.W3, w.regionStride, w.ldn, X, B, q, ys2, grad_X, st));
    return BO_OK;
}

// ------------------------------------------------------------------------------------------------
// outcome constraints across several models
// ------------------------------------------------------------------------------------------------
int bo_qlognei_samples_view(const bo_gp* gp, const bo_nei* nei, int B, int q, long* offset, long* pitch) {
    if (!valid_gp(gp) || !nei || B <= 0 || q <= 0 || !offset || !pitch || gp->M != 1) return BO_ERR_INVALID;
    char base[1];
    const QWork w = qlayout(gp, nei, B, q, 1, base, (size_t)1 << 62);
    // combined-root path with a baseline: the samples are rows 0..S-1 of the stacked forward output W1; otherwise F
    const double* rows = (w.qpath && !nei->no_baseline) ? w.W1 : w.F;
    *offset = (long)(rows - reinterpret_cast<const double*>(base));
    *pitch = w.ldn;
    return BO_OK;
}

static BoReduceArgs reduce_args(int S, int B, int q, long ldn, const double* F, const double* lw, const double* best_f,
                                double obj_weight, double tau_relu, double tau_max, int reduction) {
    BoReduceArgs a{};
    a.S = S; a.B = B; a.q = q; a.reduction = reduction; a.ldn = ldn;
    a.F = F; a.lw = lw; a.best_f = best_f;
    a.obj_w = obj_weight != 0.0 ? obj_weight : 1.0; a.tau_relu = tau_relu; a.tau_max = tau_max;
    return a;
}

int bo_mc_reduce_fwd(bo_ctx* ctx, int S, int B, int q, long ldn, const double* F, const double* lw, const double* best_f,
                     double obj_weight, double tau_relu, double tau_max, int reduction, double* acq, double* lse,
                     void* stream) {
    if (!ctx || S <= 0 || B <= 0 || q <= 0 || !F || !best_f || !acq || !lse || !(tau_relu > 0.0) || !(tau_max > 0.0))
        return BO_ERR_INVALID;
    BoReduceArgs a = reduce_args(S, B, q, ldn, F, lw, best_f, obj_weight, tau_relu, tau_max, reduction);
    a.acq = acq; a.lse = lse;
    return bo_mc_reduce_fwd_launch(a, (cudaStream_t)stream);
}

int bo_mc_reduce_bwd(bo_ctx* ctx, int S, int B, int q, long ldn, const double* F, const double* lw, const double* best_f,
                     double obj_weight, double tau_relu, double tau_max, int reduction, const double* lse,
                     const double* grad_acq, double* dF, double* lw_bar, void* stream) {
    if (!ctx || S <= 0 || B <= 0 || q <= 0 || !F || !best_f || !lse || !grad_acq || !dF) return BO_ERR_INVALID;
    BoReduceArgs a = reduce_args(S, B, q, ldn, F, lw, best_f, obj_weight, tau_relu, tau_max, reduction);
    a.lse = const_cast<double*>(lse); a.grad_acq = grad_acq; a.dF = dF; a.lw_bar = lw_bar;
    return bo_mc_reduce_bwd_launch(a, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// posterior API
// ------------------------------------------------------------------------------------------------
static size_t post_layout(const bo_gp* gp, long Nc, void* work, size_t cap, double** baug, double** kx, double** v,
                          long* ldn, bool* ok) {
    const long n = gp->n, M = gp->M, Kp = bo_kp(gp->d);
    *ldn = bo_pad8(Nc);
    Bump b(work, cap);
    *baug = b.doubles(M * Kp * *ldn);
    *kx = b.doubles(M * n * *ldn);
    *v = b.doubles(M * (n + 1) * *ldn);
    *ok = b.ok;
    return b.off;
}

size_t bo_posterior_workspace_bytes(const bo_gp* gp, long Nc) {
    double *a, *b, *c;
    long ldn;
    bool ok;
    return post_layout(gp, Nc, nullptr, 0, &a, &b, &c, &ldn, &ok);
}

static int posterior_v(const bo_gp* gp, long Nc, const double* Xc, void* work, size_t work_bytes, double** v_out,
                       double** kx_out, long* ldn_out, cudaStream_t st) {
    double *baug, *kx, *v;
    long ldn;
    bool ok;
    post_layout(gp, Nc, work, work_bytes, &baug, &kx, &v, &ldn, &ok);
    if (!ok) return BO_ERR_WORKSPACE;
    const long n = gp->n, M = gp->M, np = bo_pad8(n), Kp = bo_kp(gp->d);
    if (Nc > 0x7fffffffL) return BO_ERR_UNSUPPORTED;
    const BoKernelHyper h = hyper_of(gp);
    BO_TRY(bo_aug_cols(h, Xc, Nc, baug, Kp, ldn, Kp * ldn, st));
    BO_TRY(kmat(gp, gp->aug_train, (int)n, baug, ldn, Nc, kx, ldn, n * ldn, 0, nullptr, st));
    BO_TRY(gemm((int)n + 1, (int)Nc, (int)n, (int)M, 1.0, gp->linv_aug, np, (n + 1) * np, kx, ldn, n * ldn, 0.0, nullptr, 0,
                0, v, ldn, (n + 1) * ldn, BO_TRI_A_LOWER, 0, st));
    *v_out = v;
    *kx_out = kx;
    *ldn_out = ldn;
    return BO_OK;
}

int bo_posterior_joint(bo_ctx* ctx, const bo_gp* gp, int B, int q, const double* X, double* mean, double* cov,
                       void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || B <= 0 || q <= 0 || !X || !mean || !cov || !work) return BO_ERR_INVALID;
    if (q > BO_MAX_Q) return BO_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    double *v, *kx;
    long ldn;
    BO_TRY(posterior_v(gp, (long)B * q, X, work, work_bytes, &v, &kx, &ldn, st));
    BoCandArgs a{};
    a.h = hyper_of(gp);
    a.n = gp->n; a.B = B; a.q = q;
    a.y_mean = gp->y_mean; a.y_std = gp->y_std;
    a.mean_const = gp->mean;
    a.X = X;
    a.ldn = ldn;
    a.V = v; a.strideV = (long)(gp->n + 1) * ldn;
    return bo_cand_posterior(a, mean, cov, st);
}

// ---- joint posterior of ONE large point set (q up to n and beyond): all GEMM-shaped ------------------------------
static size_t post_large_layout(const bo_gp* gp, long q, void* work, size_t cap, double** arows, double** vt,
                                double** var, size_t* base_bytes, bool* ok) {
    const long M = gp->M, Kp = bo_kp(gp->d), np = bo_pad8(gp->n), ldq = bo_pad8(q);
    *base_bytes = bo_posterior_workspace_bytes(gp, q);
    Bump b(work ? (char*)work + *base_bytes : nullptr, cap > *base_bytes ? cap - *base_bytes : 0);
    *arows = b.doubles(M * q * Kp);
    *vt = b.doubles(M * q * np);
    *var = b.doubles(M * ldq);
    *ok = b.ok && (!work || cap >= *base_bytes);
    return *base_bytes + b.off;
}

size_t bo_posterior_joint_large_workspace_bytes(const bo_gp* gp, long q) {
    double *a, *b, *c;
    size_t base;
    bool ok;
    return post_large_layout(gp, q, nullptr, 0, &a, &b, &c, &base, &ok);
}

int bo_posterior_joint_large(bo_ctx* ctx, const bo_gp* gp, long q, const double* X, double* mean, double* cov,
                             void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || q <= 0 || !X || !mean || !cov || !work) return BO_ERR_INVALID;
    if (q > 0x7fffffffL) return BO_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    double *arows, *vt, *var, *v, *kx;
    size_t base;
    bool ok;
    post_large_layout(gp, q, work, work_bytes, &arows, &vt, &var, &base, &ok);
    if (!ok) return BO_ERR_WORKSPACE;
    long ldn;
    BO_TRY(posterior_v(gp, q, X, work, base, &v, &kx, &ldn, st));  // V = [inv(L); alpha^T] K(X_tr, X)
    const long n = gp->n, M = gp->M, np = bo_pad8(n), Kp = bo_kp(gp->d);
    const double ys2 = gp->y_std * gp->y_std;
    const BoKernelHyper h = hyper_of(gp);
    // mean (and a scratch marginal variance) from row n and the column norms of V
    BO_TRY(bo_col_mean_var(v, (n + 1) * ldn, ldn, (int)n, q, (int)M, gp->y_mean, gp->y_std, gp->mean, gp->outputscale, mean,
                           var, st));
    // cov = ys^2 (k(X, X) - V^T V): the augmented operands of X on both sides, then one q x q x n GEMM
    BO_TRY(bo_aug_rows(h, X, q, arows, Kp, q * Kp, st));
    double* baug = reinterpret_cast<double*>(work);  // first region of the base layout: augmented columns of X
    const long ldq = bo_pad8(q);  // pitch of cov: [M][q x ldq]
    BO_TRY(kmat(gp, arows, (int)q, baug, ldn, q, cov, ldq, q * ldq, 1, nullptr, st));
    BO_TRY(bo_fill(vt, M * q * np, 0.0, st));
    BO_TRY(bo_transpose(v, ldn, (n + 1) * ldn, vt, np, q * np, (int)n, (int)q, (int)M, st));
    BO_TRY(gemm((int)q, (int)q, (int)n, (int)M, -ys2, vt, np, q * np, v, ldn, (n + 1) * ldn, ys2, cov, ldq, q * ldq, cov, ldq,
                q * ldq, BO_TRI_NONE, 0, st));
    return BO_OK;
}

size_t bo_posterior_bwd_workspace_bytes(const bo_gp* gp, int B, int q) {
    const long Nc = (long)B * q, ldn = bo_pad8(Nc);
    Bump b(nullptr, 0);
    b.doubles((long)gp->M * (gp->n + 1) * ldn);
    b.doubles((long)gp->M * B * q * q);
    return bo_posterior_workspace_bytes(gp, Nc) + b.off;
}

int bo_posterior_joint_bwd(bo_ctx* ctx, const bo_gp* gp, int B, int q, const double* X, const double* mean_bar,
                           const double* cov_bar, double* grad_X, void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || B <= 0 || q <= 0 || !X || !mean_bar || !cov_bar || !grad_X || !work)
        return BO_ERR_INVALID;
    if (q > BO_MAX_Q) return BO_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const long n = gp->n, M = gp->M, n1p = bo_pad8(n + 1), Nc = (long)B * q;
    const size_t fwd_bytes = bo_posterior_workspace_bytes(gp, Nc);
    if (work_bytes < bo_posterior_bwd_workspace_bytes(gp, B, q)) return BO_ERR_WORKSPACE;
    double *v, *kx;
    long ldn;
    BO_TRY(posterior_v(gp, Nc, X, work, fwd_bytes, &v, &kx, &ldn, st));  // recompute V (cheaper than keeping it)
    Bump b((char*)work + fwd_bytes, work_bytes - fwd_bytes);
    const long regionV = (n + 1) * ldn;
    double* vbar = b.doubles(M * regionV);
    double* tbuf = b.doubles(M * (long)B * q * q);
    const double ys2 = gp->y_std * gp->y_std;
    BoCandArgs a{};
    a.h = hyper_of(gp);
    a.n = gp->n; a.B = B; a.q = q;
    a.y_mean = gp->y_mean; a.y_std = gp->y_std;
    a.X = X; a.ldn = ldn;
    a.Vbar_mu = vbar + n * ldn; a.strideVbar = regionV;
    a.gradX = grad_X;
    BO_TRY(bo_fill(vbar, M * regionV, 0.0, st));
    BO_TRY(bo_cand_posterior_bwd(a, mean_bar, cov_bar, tbuf, st));
    // Vbar = -ys^2 V (cov_bar + cov_bar^T)
    BO_TRY(bo_colblock_update(vbar, regionV, v, regionV, tbuf, (int)n, B, q, (int)M, ldn, -ys2, 1, st));
    // KXbar = [inv(L)^T | alpha] Vbar     (into the kx region)
    BO_TRY(gemm((int)n, (int)Nc, (int)n + 1, (int)M, 1.0, gp->linv_aug_t, n1p, n * n1p, vbar, ldn, regionV, 0.0, nullptr, 0,
                0, kx, ldn, n * ldn, BO_TRI_A_UPPER, 0, st));
    return bo_kgrad(hyper_of(gp), gp->train_X, (int)n, kx, n * ldn, ldn, X, B, q, 1.0, grad_X, st);
}

int bo_posterior_mean_var(bo_ctx* ctx, const bo_gp* gp, long Nc, const double* Xc, double* mean, double* var,
                          void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || Nc <= 0 || !Xc || !mean || !var || !work) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    double *v, *kx;
    long ldn;
    BO_TRY(posterior_v(gp, Nc, Xc, work, work_bytes, &v, &kx, &ldn, st));
    return bo_col_mean_var(v, (long)(gp->n + 1) * ldn, ldn, gp->n, Nc, gp->M, gp->y_mean, gp->y_std, gp->mean,
                           gp->outputscale, mean, var, st);
}

int bo_ts_argmax(bo_ctx* ctx, int M, long Nc, const double* mean, const double* var, const double* z,
                 double* out_val, int64_t* out_idx, void* stream) {
    if (!ctx || !mean || !out_val || !out_idx || (z && !var)) return BO_ERR_INVALID;
    return bo_ts_argmax_launch(M, Nc, mean, var, z, out_val, reinterpret_cast<long long*>(out_idx), (cudaStream_t)stream);
}

int bo_argmax(bo_ctx* ctx, long n, const double* v, double* out_val, int64_t* out_idx, void* stream) {
    if (!ctx || !v || !out_val || !out_idx) return BO_ERR_INVALID;
    return bo_ts_argmax_launch(1, n, v, nullptr, nullptr, out_val, reinterpret_cast<long long*>(out_idx),
                               (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// exact marginal log likelihood, value + gradient w.r.t. the hyperparameters
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace {

constexpr int MLL_CHUNKS = 16;  // row chunks of the column sum-of-squares of inv(L) (summed in order)

// Bn[z][k][j] = -inv(L)[z][k][j] (k < n), alpha[z][j] (k == n): [inv(L)^T | alpha] Bn = alpha alpha^T - (K + s2 I)^-1
__global__ void mll_negate_kernel(const double* linv_aug, double* Bn, long np, int n) {
    const long z = blockIdx.z, k = blockIdx.y;
    const long base = z * (long)(n + 1) * np + k * np;
    for (long j = blockIdx.x * (long)blockDim.x + threadIdx.x; j < np; j += (long)gridDim.x * blockDim.x)
        Bn[base + j] = (k < n) ? -linv_aug[base + j] : linv_aug[base + j];
}
// xct[t][j] = train_X[j][t] - center[t] (t < d), 1 (t == d)
__global__ void mll_xct_kernel(const double* X, const double* center, double* xct, long np, int n, int d) {
    const int t = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < np) xct[(long)t * np + j] = (j < n) ? (t < d ? X[(long)j * d + t] - center[t] : 1.0) : 0.0;
}
// part[z][c][j] = sum over the c-th row chunk of inv(L)[z][k][j]^2   (diag of (K + s2 I)^-1 = column norms of inv(L))
__global__ void mll_colsq_kernel(const double* linv_aug, double* part, long np, int n) {
    const long z = blockIdx.z;
    const int c = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int rows = (n + MLL_CHUNKS - 1) / MLL_CHUNKS;
    int k0 = c * rows, k1 = k0 + rows < n ? k0 + rows : n;
    if (k0 < j) k0 = j;  // inv(L) is lower triangular
    const double* L = linv_aug + z * (long)(n + 1) * np;
    double s = 0.0;
    for (int k = k0; k < k1; k++) { const double v = L[(long)k * np + j]; s += v * v; }
    part[(z * MLL_CHUNKS + c) * np + j] = s;
}

__device__ double mll_block_sum(double v, double* sm) {  // fixed-order tree: deterministic
    const int tid = threadIdx.x;
    sm[tid] = v;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
        if (tid < s) sm[tid] += sm[tid + s];
        __syncthreads();
    }
    const double r = sm[0];
    __syncthreads();
    return r;
}

struct MllFin {
    int n, d, ks;
    long np, strideH, strideKs;
    const double *H, *xct, *inv_ls, *outputscale, *noise, *noise_vec, *jitter, *mean, *targets, *linv_aug, *chol, *part;
    double *value, *grad;
};
// grid (d+1, M): blocks 0..d-1 the lengthscale gradients, block d value + noise / mean / outputscale gradients
__global__ void __launch_bounds__(256) mll_finalize_kernel(const MllFin a) {
    __shared__ double sm[256];
    const int z = blockIdx.y, t = blockIdx.x, tid = threadIdx.x, n = a.n, d = a.d;
    double* g = a.grad + (long)z * (d + 3);
    if (t < d) {
        double s = 0.0;
        for (int j = tid; j < n; j += blockDim.x) {
            double hd = 0.0, ht = 0.0;
            for (int k = 0; k < a.ks; k++) {
                const double* Hk = a.H + (long)z * a.strideH + (long)k * a.strideKs;
                hd += Hk[(long)d * a.np + j];
                ht += Hk[(long)t * a.np + j];
            }
            const double x = a.xct[(long)t * a.np + j];
            s += x * (x * hd - ht);
        }
        s = mll_block_sum(s, sm);
        if (tid == 0) { const double il = a.inv_ls[(long)z * d + t]; g[t] = s * il * il * il; }
        return;
    }
    const double* alpha = a.linv_aug + (long)z * (n + 1) * a.np + (long)n * a.np;
    const double* L = a.chol + (long)z * n * a.np;
    const double jit = a.jitter ? a.jitter[z] : 0.0;
    double ra = 0.0, sa = 0.0, ld = 0.0, trw = 0.0, trwd = 0.0;
    for (int i = tid; i < n; i += blockDim.x) {
        const double al = alpha[i];
        ra += (a.targets[i] - a.mean[z]) * al;
        sa += al;
        ld += log(L[(long)i * a.np + i]);
        double kd = 0.0;
        for (int c = 0; c < MLL_CHUNKS; c++) kd += a.part[((long)z * MLL_CHUNKS + c) * a.np + i];
        const double w = al * al - kd;
        trw += w;
        trwd += w * ((a.noise_vec ? a.noise_vec[(long)z * n + i] : a.noise[z]) + jit);
    }
    ra = mll_block_sum(ra, sm); sa = mll_block_sum(sa, sm); ld = mll_block_sum(ld, sm);
    trw = mll_block_sum(trw, sm); trwd = mll_block_sum(trwd, sm);
    if (tid == 0) {
        a.value[z] = -0.5 * ra - ld - 0.5 * n * 1.8378770664093454835606594728112;  // log(2 pi)
        const double os = a.outputscale ? a.outputscale[z] : 1.0;
        g[d] = (ra - n - trwd) / (2.0 * os);  // sum_ij W_ij K_ij = tr(W (Khat - D)) = r.alpha - n - tr(W D)
        g[d + 1] = 0.5 * trw;
        g[d + 2] = sa;
    }
}

struct MllWork {
    void* setup; size_t setup_bytes;
    double *baug, *Kg, *Bn, *G, *xct, *H, *part;
    int ks; bool ok; size_t bytes;
};
MllWork mll_layout(const bo_gp* gp, void* work, size_t cap) {
    MllWork w{};
    const long n = gp->n, M = gp->M, np = bo_pad8(n), Kp = bo_kp(gp->d);
    w.setup_bytes = (size_t)bo_round_up((long)bo_gp_setup_workspace_bytes(gp), 256);
    Bump b(work, cap);
    w.setup = b.doubles((long)(w.setup_bytes / sizeof(double)));
    w.baug = b.doubles(M * Kp * np);
    w.Kg = b.doubles(M * n * np);
    w.Bn = b.doubles(M * (n + 1) * np);
    w.G = b.doubles(M * n * np);
    w.xct = b.doubles((long)(gp->d + 1) * np);
    w.ks = hsplit(gp);
    w.H = b.doubles(M * w.ks * (long)(gp->d + 1) * np);
    w.part = b.doubles(M * MLL_CHUNKS * np);
    w.ok = b.ok;
    w.bytes = b.off;
    return w;
}

}  // namespace

extern "C" {

size_t bo_mll_workspace_bytes(const bo_gp* gp) { return mll_layout(gp, nullptr, 0).bytes; }

int bo_mll_value_grad(bo_ctx* ctx, const bo_gp* gp, void* work, size_t work_bytes, double* value, double* grad,
                      int32_t* info, void* stream) {
    if (!ctx || !valid_gp(gp) || !work || !value || !grad || !info || !gp->xt_aug) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const MllWork w = mll_layout(gp, work, work_bytes);
    if (!w.ok) return BO_ERR_WORKSPACE;
    const long n = gp->n, M = gp->M, np = bo_pad8(n), n1p = bo_pad8(n + 1), Kp = bo_kp(gp->d);
    const int d = gp->d;
    if (n + 1 > 65535 || M > 65535) return BO_ERR_UNSUPPORTED;
    // 1. K + D = L L^T, inv(L), alpha: the prediction caches of these hyperparameters stay valid afterwards
    BO_TRY(bo_gp_setup(ctx, gp, w.setup, w.setup_bytes, info, stream));
    // 2. derivative factor g(r) of the covariance function (g = K for RBF)
    const BoKernelHyper h = hyper_of(gp);
    BO_TRY(bo_aug_cols(h, gp->train_X, n, w.baug, Kp, np, Kp * np, st));
    if (gp->kernel_id == BO_KERNEL_RBF)
        BO_TRY(kmat(gp, gp->aug_train, (int)n, w.baug, np, n, w.Kg, np, n * np, 1, nullptr, st));
    else
        BO_TRY(kmat(gp, gp->aug_train, (int)n, w.baug, np, n, w.G, np, n * np, 1, nullptr, st, w.Kg));
    // 3. G = (alpha alpha^T - (K + D)^-1) .* g   in one GEMM over the augmented factors
    {
        dim3 grid((unsigned)((np + 255) / 256 > 8 ? 8 : (np + 255) / 256), (unsigned)(n + 1), (unsigned)M);
        mll_negate_kernel<<<grid, 256, 0, st>>>(gp->linv_aug, w.Bn, np, (int)n);
        g_bo_launch_count++;
        BoGemm g{};
        g.A = gp->linv_aug_t; g.lda = n1p; g.strideA = n * n1p;
        g.B = w.Bn; g.ldb = np; g.strideB = (n + 1) * np;
        g.C = w.G; g.ldc = np; g.strideC = n * np;
        g.E = w.Kg; g.lde = np; g.strideE = n * np;
        g.alpha = 1.0; g.M = (int)n; g.N = (int)n; g.K = (int)n + 1; g.batch = (int)M;
        g.tri = BO_TRI_A_UPPER; g.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(g, st));
    }
    // 4. H = [x~^T; 1] G   ((d+1) x n), x~ = train_X - center
    {
        dim3 grid((unsigned)((np + 127) / 128), (unsigned)(d + 1));
        mll_xct_kernel<<<grid, 128, 0, st>>>(gp->train_X, gp->center, w.xct, np, (int)n, d);
        g_bo_launch_count++;
    }
    const long strideH = (long)(d + 1) * np;
    if (w.ks > 1) {
        const long kc = n / w.ks;
        BO_TRY(gemm(d + 1, (int)n, (int)kc, w.ks, 1.0, w.xct, np, kc, w.G, np, kc * np, 0.0, nullptr, 0, 0, w.H, np, strideH,
                    BO_TRI_NONE, 0, st));
    } else {
        BO_TRY(gemm(d + 1, (int)n, (int)n, (int)M, 1.0, w.xct, np, 0, w.G, np, n * np, 0.0, nullptr, 0, 0, w.H, np, strideH,
                    BO_TRI_NONE, 0, st));
    }
    // 5. diag of (K + D)^-1 and the reductions
    {
        dim3 grid((unsigned)((n + 127) / 128), MLL_CHUNKS, (unsigned)M);
        mll_colsq_kernel<<<grid, 128, 0, st>>>(gp->linv_aug, w.part, np, (int)n);
        g_bo_launch_count++;
    }
    MllFin f{};
    f.n = (int)n; f.d = d; f.ks = w.ks; f.np = np;
    f.strideH = w.ks > 1 ? 0 : strideH; f.strideKs = w.ks > 1 ? strideH : 0;
    f.H = w.H; f.xct = w.xct; f.inv_ls = gp->inv_ls; f.outputscale = gp->outputscale; f.noise = gp->noise;
    f.noise_vec = gp->noise_vec; f.jitter = gp->jitter; f.mean = gp->mean; f.targets = gp->targets;
    f.linv_aug = gp->linv_aug; f.chol = gp->chol; f.part = w.part; f.value = value; f.grad = grad;
    mll_finalize_kernel<<<dim3((unsigned)(d + 1), (unsigned)M), 256, 0, st>>>(f);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

// ------------------------------------------------------------------------------------------------
// SAAS potential energy: what pyro's NUTS evaluates per leapfrog step inside
// fit_fully_bayesian_model_nuts (fit.py:335-396 over SaasPyroModel.sample, models/fully_bayesian.py:167-246)
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace {

constexpr double SAAS_MIN_NOISE = 1e-4;  // MIN_INFERRED_NOISE_LEVEL (models/utils/gpytorch_modules.py:29)
constexpr double SAAS_U_MAX = 80.0;      // |u| beyond this: exp() leaves the useful float64 range -> divergent

// u [C x (d+4)] = [log tausq, log inv_length_sq (d), mean, log noise_raw, log outputscale] -> the hyperparameter buffers of gp
__global__ void saas_transform_kernel(const double* u, int d, double* inv_ls, double* outputscale, double* noise,
                                      double* mean, int* wild) {
    const int c = blockIdx.x, D = d + 4;
    const double* uc = u + (long)c * D;
    __shared__ int bad;
    if (threadIdx.x == 0) bad = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < D; i += blockDim.x)
        if (!(fabs(uc[i]) < SAAS_U_MAX)) bad = 1;  // also catches NaN
    __syncthreads();
    const int b = bad;
    for (int i = threadIdx.x; i < d; i += blockDim.x)
        inv_ls[(long)c * d + i] = b ? 1.0 : exp(0.5 * (uc[0] + uc[1 + i]));  // 1 / lengthscale = sqrt(tausq * inv_length_sq)
    if (threadIdx.x == 0) {
        mean[c] = b ? 0.0 : uc[1 + d];
        noise[c] = SAAS_MIN_NOISE + (b ? 1.0 : exp(uc[2 + d]));
        outputscale[c] = b ? 1.0 : exp(uc[3 + d]);
        wild[c] = b;
    }
}

// energy = -(log lik + log priors + log|dx/du|), gradient by the closed-form chain rule (x = exp(u): d/du = x d/dx;
// lengthscale_i = exp(-(u_0 + u_{1+i}) / 2)  =>  d lengthscale_i / d u_0 = d lengthscale_i / d u_{1+i} = -lengthscale_i / 2)
__global__ void __launch_bounds__(128) saas_finalize_kernel(const double* u, int d, const double* value, const double* grad,
                                                            const int* info, const int* wild, double* energy,
                                                            double* grad_u) {
    __shared__ double sm[128];
    const int c = blockIdx.x, D = d + 4, tid = threadIdx.x;
    const double* uc = u + (long)c * D;
    const double* g = grad + (long)c * (d + 3);  // [d lengthscale (d), outputscale, noise, mean]
    double* gu = grad_u + (long)c * D;
    const bool bad = wild[c] || info[c] != 0;
    const double u0 = bad ? 0.0 : uc[0];
    double lp = 0.0, gsum = 0.0;  // half-Cauchy(1) sites + their Jacobians; sum_i d ll / d u_{1+i}
    for (int i = tid; i < d; i += blockDim.x) {
        const double ui = bad ? 0.0 : uc[1 + i];
        const double x = exp(ui), r2 = x * x;
        const double hl = -0.5 * exp(-0.5 * (u0 + ui)) * g[i];
        lp += -0.4515827052894548647 - log1p(r2) + ui;  // log(2/pi)
        gsum += hl;
        gu[1 + i] = -(hl + 1.0 - 2.0 * r2 / (1.0 + r2));
    }
    sm[tid] = lp;
    __syncthreads();
    for (int s_ = 64; s_ > 0; s_ >>= 1) { if (tid < s_) sm[tid] += sm[tid + s_]; __syncthreads(); }
    lp = sm[0];
    __syncthreads();
    sm[tid] = gsum;
    __syncthreads();
    for (int s_ = 64; s_ > 0; s_ >>= 1) { if (tid < s_) sm[tid] += sm[tid + s_]; __syncthreads(); }
    gsum = sm[0];
    if (tid == 0) {
        const double mean = bad ? 0.0 : uc[1 + d], unr = bad ? 0.0 : uc[2 + d], uos = bad ? 0.0 : uc[3 + d];
        const double t = exp(u0), nr = exp(unr), os = exp(uos), r2 = 100.0 * t * t;  // tausq ~ HalfCauchy(0.1)
        // Gamma(a, b): a log b + (a-1) log x - b x - lgamma(a); lgamma(2) = 0, lgamma(0.9) = 0.06637623973474297
        const double lp_os = 2.0 * log(0.15) + uos - 0.15 * os;
        const double lp_nr = 0.9 * log(10.0) - 0.1 * unr - 10.0 * nr - 0.066376239734742966;
        const double lp_t = -0.4515827052894548647 - log(0.1) - log1p(r2);
        const double lp_m = -0.5 * mean * mean - 0.91893853320467274178;
        const double logp = value[c] + lp + lp_t + lp_os + lp_nr + lp_m + u0 + unr + uos;
        double e = -logp;
        double g0 = -(gsum + 1.0 - 2.0 * r2 / (1.0 + r2));
        double gm = -(g[d + 2] - mean);
        double gn = -(g[d + 1] * nr - 0.1 - 10.0 * nr + 1.0);
        double go = -(g[d] * os + 1.0 - 0.15 * os + 1.0);
        bool fail = bad || !isfinite(e) || !isfinite(g0) || !isfinite(gm) || !isfinite(gn) || !isfinite(go);
        sm[0] = fail ? 1.0 : 0.0;
        energy[c] = e; gu[0] = g0; gu[1 + d] = gm; gu[2 + d] = gn; gu[3 + d] = go;
    }
    __syncthreads();
    bool fail = sm[0] != 0.0;
    __syncthreads();
    // a non-finite lengthscale gradient entry also marks the chain as failed
    int nf = 0;
    for (int i = tid; i < d; i += blockDim.x) nf |= !isfinite(gu[1 + i]);
    sm[tid] = (double)nf;
    __syncthreads();
    for (int s_ = 64; s_ > 0; s_ >>= 1) { if (tid < s_) sm[tid] += sm[tid + s_]; __syncthreads(); }
    fail = fail || sm[0] != 0.0;
    if (fail) {
        for (int i = tid; i < D; i += blockDim.x) gu[i] = 0.0;
        if (tid == 0) energy[c] = INFINITY;
    }
}

}  // namespace

extern "C" {

// gp: Matern-5/2 with outputscale, M = number of chains, hyperparameter buffers (inv_ls, outputscale, noise, mean) writable.
int bo_saas_potential(bo_ctx* ctx, const bo_gp* gp, const double* u, void* work, size_t work_bytes, double* value,
                      double* grad, int32_t* wild, double* energy, double* grad_u, int32_t* info, void* stream) {
    if (!ctx || !gp || !u || !value || !grad || !wild || !energy || !grad_u || !info) return BO_ERR_INVALID;
    if (!gp->outputscale || gp->noise_vec) return BO_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const int C = (int)gp->M, d = (int)gp->d;
    saas_transform_kernel<<<C, 64, 0, st>>>(u, d, const_cast<double*>(gp->inv_ls), const_cast<double*>(gp->outputscale),
                                            const_cast<double*>(gp->noise), const_cast<double*>(gp->mean), wild);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    BO_TRY(bo_mll_value_grad(ctx, gp, work, work_bytes, value, grad, info, stream));
    saas_finalize_kernel<<<C, 128, 0, st>>>(u, d, value, grad, info, wild, energy, grad_u);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

// ------------------------------------------------------------------------------------------------
// pathwise (Matheron) posterior samples
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace {

// w_aug[f] = [W[f][0..d-1], sum_t W[f][t] center[t] / l[t], 0, ...]: against the column operand [(x-c)/l; 1; |.|^2] of
// bo_aug_cols the contraction is W[f] . x / l  (InverseLengthscaleTransform, sampling/pathwise/utils.py:101-117)
__global__ void paths_waug_kernel(const double* W, const double* inv_ls, const double* center, double* w_aug, int half,
                                  int d, long Kp) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= half) return;
    double off = 0.0;
    for (int t = 0; t < d; t++) {
        const double wv = W[(long)f * d + t];
        w_aug[(long)f * Kp + t] = wv;
        off += wv * center[t] * inv_ls[t];
    }
    w_aug[(long)f * Kp + d] = off;
    for (long k = d + 1; k < Kp; k++) w_aug[(long)f * Kp + k] = 0.0;
}
// r[s][i] = targets[i] - mean - f0[s][i] - sqrt(D_i) eps[s][i]      (update_strategies.py:82-91)
__global__ void paths_resid_kernel(const double* targets, const double* mean, const double* noise, const double* noise_vec,
                                   const double* f0, const double* eps, double* r, int n, long np) {
    const int s = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    double v = 0.0;
    if (i < n) {
        const double dn = noise_vec ? noise_vec[i] : noise[0];
        v = targets[i] - mean[0] - f0[(long)s * np + i] - sqrt(dn) * eps[(long)s * n + i];
    }
    r[(long)s * np + i] = v;
}
// out = y_mean + y_std * (mean + out)
__global__ void paths_affine_kernel(double* out, long count, double y_mean, double y_std, const double* mean) {
    const double b = y_mean + y_std * mean[0];
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < count; i += (long)gridDim.x * blockDim.x)
        out[i] = b + y_std * out[i];
}

bool valid_paths(const bo_gp* gp, const bo_paths* p) {
    return p && gp->M == 1 && p->S > 0 && p->F >= 2 && (p->F % 2) == 0 && p->W && p->w && p->eps && p->w_aug && p->w_pad && p->v;
}

// PhiT[F x ldn] = features of the points whose column operand is baug; then C[S x ldn] = alpha * w_pad PhiT (+ beta D)
int paths_features(const bo_gp* gp, const bo_paths* p, const double* baug, long npts, long ldn, double* PhiT,
                   cudaStream_t st) {
    const long Kp = bo_kp(gp->d), half = p->F / 2;
    BoGemm g{};
    g.A = p->w_aug; g.lda = Kp; g.strideA = 0;
    g.B = baug; g.ldb = ldn; g.strideB = 0;
    g.C = PhiT; g.ldc = ldn; g.strideC = 0;
    g.C2 = PhiT + half * ldn; g.ldc2 = ldn; g.strideC2 = 0;
    g.alpha = sqrt(2.0 / (double)p->F); g.os = gp->outputscale;
    g.M = (int)half; g.N = (int)npts; g.K = (int)Kp; g.batch = 1; g.epi = BO_EPI_SINCOS;
    return bo_gemm_launch(g, st);
}

struct PathsSetupWork { double *baug, *PhiT, *f0, *r, *t; bool ok; size_t bytes; };
PathsSetupWork paths_setup_layout(const bo_gp* gp, const bo_paths* p, void* work, size_t cap) {
    PathsSetupWork w{};
    const long n = gp->n, np = bo_pad8(n), Kp = bo_kp(gp->d), Sp = bo_pad8(p->S);
    Bump b(work, cap);
    w.baug = b.doubles(Kp * np);
    w.PhiT = b.doubles((long)p->F * np);
    w.f0 = b.doubles(Sp * np);
    w.r = b.doubles(Sp * np);
    w.t = b.doubles(Sp * np);
    w.ok = b.ok; w.bytes = b.off;
    return w;
}
struct PathsEvalWork { double *baug, *PhiT, *kx; long ldn; bool ok; size_t bytes; };
PathsEvalWork paths_eval_layout(const bo_gp* gp, const bo_paths* p, long Nc, void* work, size_t cap) {
    PathsEvalWork w{};
    const long Kp = bo_kp(gp->d);
    w.ldn = bo_pad8(Nc);
    Bump b(work, cap);
    w.baug = b.doubles(Kp * w.ldn);
    w.PhiT = b.doubles((long)p->F * w.ldn);
    w.kx = b.doubles((long)gp->n * w.ldn);
    w.ok = b.ok; w.bytes = b.off;
    return w;
}

}  // namespace

extern "C" {

size_t bo_paths_setup_workspace_bytes(const bo_gp* gp, const bo_paths* paths) {
    return paths_setup_layout(gp, paths, nullptr, 0).bytes;
}

int bo_paths_setup(bo_ctx* ctx, const bo_gp* gp, const bo_paths* paths, void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || !valid_paths(gp, paths) || !work) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const PathsSetupWork w = paths_setup_layout(gp, paths, work, work_bytes);
    if (!w.ok) return BO_ERR_WORKSPACE;
    const long n = gp->n, np = bo_pad8(n), n1p = bo_pad8(n + 1), Kp = bo_kp(gp->d), S = paths->S, F = paths->F, Fp = bo_pad8(F);
    const BoKernelHyper h = hyper_of(gp);
    paths_waug_kernel<<<(unsigned)((F / 2 + 127) / 128), 128, 0, st>>>(paths->W, gp->inv_ls, gp->center, paths->w_aug,
                                                                       (int)(F / 2), gp->d, Kp);
    g_bo_launch_count++;
    BO_TRY(bo_fill(paths->w_pad, S * Fp, 0.0, st));
    BO_TRY(bo_copy2d(paths->w, F, 0, paths->w_pad, Fp, 0, (int)S, (int)F, 1, st));
    // prior paths at the training inputs: f0(X) - m = w Phi(X)^T
    BO_TRY(bo_aug_cols(h, gp->train_X, n, w.baug, Kp, np, Kp * np, st));
    BO_TRY(paths_features(gp, paths, w.baug, n, np, w.PhiT, st));
    BO_TRY(gemm((int)S, (int)n, (int)F, 1, 1.0, paths->w_pad, Fp, 0, w.PhiT, np, 0, 0.0, nullptr, 0, 0, w.f0, np, 0,
                BO_TRI_NONE, 0, st));
    dim3 grid((unsigned)((np + 127) / 128), (unsigned)S);
    paths_resid_kernel<<<grid, 128, 0, st>>>(gp->targets, gp->mean, gp->noise, gp->noise_vec, w.f0, paths->eps, w.r, (int)n, np);
    g_bo_launch_count++;
    // v = r (K + D)^-1 = (r inv(L)^T) inv(L), paths as rows
    BO_TRY(gemm((int)S, (int)n, (int)n, 1, 1.0, w.r, np, 0, gp->linv_aug_t, n1p, 0, 0.0, nullptr, 0, 0, w.t, np, 0, BO_TRI_NONE,
                0, st));
    BO_TRY(bo_fill(paths->v, S * np, 0.0, st));
    BO_TRY(gemm((int)S, (int)n, (int)n, 1, 1.0, w.t, np, 0, gp->linv_aug, np, 0, 0.0, nullptr, 0, 0, paths->v, np, 0,
                BO_TRI_B_LOWER, 0, st));
    if (paths->w_t && paths->v_t) {  // transposed copies for the gradient path
        const long Sp = bo_pad8(S);
        BO_TRY(bo_fill(paths->w_t, F * Sp, 0.0, st));
        BO_TRY(bo_transpose(paths->w, F, 0, paths->w_t, Sp, 0, (int)S, (int)F, 1, st));
        BO_TRY(bo_fill(paths->v_t, n * Sp, 0.0, st));
        BO_TRY(bo_transpose(paths->v, np, 0, paths->v_t, Sp, 0, (int)S, (int)n, 1, st));
    }
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

size_t bo_paths_eval_workspace_bytes(const bo_gp* gp, const bo_paths* paths, long Nc) {
    return paths_eval_layout(gp, paths, Nc, nullptr, 0).bytes;
}

int bo_paths_eval(bo_ctx* ctx, const bo_gp* gp, const bo_paths* paths, long Nc, const double* Xc, double* out, void* work,
                  size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || !valid_paths(gp, paths) || Nc <= 0 || !Xc || !out || !work) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const PathsEvalWork w = paths_eval_layout(gp, paths, Nc, work, work_bytes);
    if (!w.ok) return BO_ERR_WORKSPACE;
    const long n = gp->n, np = bo_pad8(n), Kp = bo_kp(gp->d), S = paths->S, F = paths->F, Fp = bo_pad8(F), ldn = w.ldn;
    const BoKernelHyper h = hyper_of(gp);
    BO_TRY(bo_aug_cols(h, Xc, Nc, w.baug, Kp, ldn, Kp * ldn, st));
    BO_TRY(paths_features(gp, paths, w.baug, Nc, ldn, w.PhiT, st));
    BO_TRY(kmat(gp, gp->aug_train, (int)n, w.baug, ldn, Nc, w.kx, ldn, n * ldn, 0, nullptr, st));
    // out = w Phi(x)^T + v k(X, x), then the mean and the outcome transform
    BO_TRY(gemm((int)S, (int)Nc, (int)F, 1, 1.0, paths->w_pad, Fp, 0, w.PhiT, ldn, 0, 0.0, nullptr, 0, 0, out, ldn, 0,
                BO_TRI_NONE, 0, st));
    BO_TRY(gemm((int)S, (int)Nc, (int)n, 1, 1.0, paths->v, np, 0, w.kx, ldn, 0, 1.0, out, ldn, 0, out, ldn, 0, BO_TRI_NONE, 0,
                st));
    long blocks = (S * ldn + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    paths_affine_kernel<<<(unsigned)blocks, 256, 0, st>>>(out, S * ldn, gp->y_mean, gp->y_std, gp->mean);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

}  // extern "C"

namespace {

// wlsT[t][f] = W[f % half][t] / l[t]   ([d x pad8(F)]: the frequency matrix stacked twice, transposed, in input units)
__global__ void paths_wlsT_kernel(const double* W, const double* inv_ls, double* wlsT, int half, int d, long Fp) {
    const int t = blockIdx.y;
    const long f = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (f < Fp) wlsT[(long)t * Fp + f] = f < 2 * half ? W[(f % half) * d + t] * inv_ls[t] : 0.0;
}

struct PathsGradWork { double *baug, *PhiT, *kx, *wd, *R, *H, *gT, *wlsT; long ldn; int ks; bool ok; size_t bytes; };
PathsGradWork paths_grad_layout(const bo_gp* gp, const bo_paths* p, long Nc, void* work, size_t cap) {
    PathsGradWork w{};
    const long Kp = bo_kp(gp->d);
    w.ldn = bo_pad8(Nc);
    Bump b(work, cap);
    w.baug = b.doubles(Kp * w.ldn);
    w.PhiT = b.doubles((long)p->F * w.ldn);
    w.kx = b.doubles((long)gp->n * w.ldn);
    w.wd = gp->kernel_id != BO_KERNEL_RBF ? b.doubles((long)gp->n * w.ldn) : nullptr;
    w.R = b.doubles((long)p->F * w.ldn);
    w.ks = hsplit(gp);
    w.H = b.doubles((long)w.ks * (gp->d + 1) * w.ldn);
    w.gT = b.doubles((long)gp->d * w.ldn);
    w.wlsT = b.doubles((long)gp->d * bo_pad8(p->F));
    w.ok = b.ok; w.bytes = b.off;
    return w;
}

}  // namespace

extern "C" {

size_t bo_paths_grad_workspace_bytes(const bo_gp* gp, const bo_paths* paths, long Nc) {
    return paths_grad_layout(gp, paths, Nc, nullptr, 0).bytes;
}

int bo_paths_grad(bo_ctx* ctx, const bo_gp* gp, const bo_paths* paths, long Nc, const double* Xc, const double* out_bar,
                  double* grad_X, void* work, size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || !valid_paths(gp, paths) || !paths->w_t || !paths->v_t || !gp->xt_aug || Nc <= 0 || !Xc ||
        !out_bar || !grad_X || !work)
        return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const PathsGradWork w = paths_grad_layout(gp, paths, Nc, work, work_bytes);
    if (!w.ok) return BO_ERR_WORKSPACE;
    const long n = gp->n, np = bo_pad8(n), Kp = bo_kp(gp->d), S = paths->S, F = paths->F, half = F / 2, Sp = bo_pad8(S);
    const long ldn = w.ldn, Fp = bo_pad8(F);
    const int d = gp->d;
    const BoKernelHyper h = hyper_of(gp);
    BO_TRY(bo_aug_cols(h, Xc, Nc, w.baug, Kp, ldn, Kp * ldn, st));
    BO_TRY(paths_features(gp, paths, w.baug, Nc, ldn, w.PhiT, st));
    BO_TRY(kmat(gp, gp->aug_train, (int)n, w.baug, ldn, Nc, w.kx, ldn, n * ldn, 0, nullptr, st, w.wd));
    // feature part: R = ys (w^T out_bar) .* d phi / dz = [ (w_lo^T out_bar) .* cos ; -(w_hi^T out_bar) .* sin ]
    for (int part = 0; part < 2; part++) {
        BoGemm g{};
        g.A = paths->w_t + (long)part * half * Sp; g.lda = Sp; g.strideA = 0;
        g.B = out_bar; g.ldb = ldn; g.strideB = 0;
        g.C = w.R + (long)part * half * ldn; g.ldc = ldn; g.strideC = 0;
        g.E = w.PhiT + (long)(1 - part) * half * ldn; g.lde = ldn; g.strideE = 0;
        g.alpha = part == 0 ? gp->y_std : -gp->y_std; g.M = (int)half; g.N = (int)Nc; g.K = (int)S; g.batch = 1;
        g.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(g, st));
    }
    // grad^T [d x N] = [W / l stacked twice]^T R; its transposed copy IS grad_X [N x d]
    {
        dim3 grid((unsigned)((Fp + 127) / 128), (unsigned)d);
        paths_wlsT_kernel<<<grid, 128, 0, st>>>(paths->W, gp->inv_ls, w.wlsT, (int)half, d, Fp);
        g_bo_launch_count++;
        BoGemm g{};
        g.A = w.wlsT; g.lda = Fp; g.strideA = 0;
        g.B = w.R; g.ldb = ldn; g.strideB = 0;
        g.C = w.gT; g.ldc = ldn; g.strideC = 0;
        g.CT = grad_X; g.ldct = d; g.strideCT = 0;
        g.alpha = 1.0; g.M = d; g.N = (int)Nc; g.K = (int)F; g.batch = 1; g.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(g, st));
    }
    // data part: G = ys (v^T out_bar) .* g(r) (in place over the derivative factor), H = [X^T; 1] G, grad += (H_t - x_t H_d) / l_t^2
    double* G = w.wd ? w.wd : w.kx;
    {
        BoGemm g{};
        g.A = paths->v_t; g.lda = Sp; g.strideA = 0;
        g.B = out_bar; g.ldb = ldn; g.strideB = 0;
        g.C = G; g.ldc = ldn; g.strideC = 0;
        g.E = G; g.lde = ldn; g.strideE = 0;
        g.alpha = gp->y_std; g.M = (int)n; g.N = (int)Nc; g.K = (int)S; g.batch = 1; g.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(g, st));
    }
    const long strideH = (long)(d + 1) * ldn;
    if (w.ks > 1) {
        const long kc = n / w.ks;
        BO_TRY(gemm(d + 1, (int)Nc, (int)kc, w.ks, 1.0, gp->xt_aug, np, kc, G, ldn, kc * ldn, 0.0, nullptr, 0, 0, w.H, ldn,
                    strideH, BO_TRI_NONE, 0, st));
        return bo_kgrad_finalize(h, w.H, 0, ldn, Xc, Nc, 1.0, grad_X, w.ks, strideH, st);
    }
    BO_TRY(gemm(d + 1, (int)Nc, (int)n, 1, 1.0, gp->xt_aug, np, 0, G, ldn, 0, 0.0, nullptr, 0, 0, w.H, ldn, strideH, BO_TRI_NONE,
                0, st));
    return bo_kgrad_finalize(h, w.H, strideH, ldn, Xc, Nc, 1.0, grad_X, 1, 0, st);
}

// ------------------------------------------------------------------------------------------------
// building blocks exported for the parity tests
// ------------------------------------------------------------------------------------------------
int bo_dgemm(bo_ctx* ctx, int M, int N, int K, int batch, double alpha, const double* A, long lda, long strideA,
             const double* B, long ldb, long strideB, double beta, double* C, long ldc, long strideC, int tri,
             void* stream) {
    if (!ctx || !A || !B || !C) return BO_ERR_INVALID;
    return gemm(M, N, K, batch, alpha, A, lda, strideA, B, ldb, strideB, beta, beta != 0.0 ? C : nullptr, ldc, strideC, C,
                ldc, strideC, tri, 0, (cudaStream_t)stream);
}

int bo_kernel_matrix(bo_ctx* ctx, const bo_gp* gp, long Nc, const double* Xc, double* out, void* work,
                     size_t work_bytes, void* stream) {
    if (!ctx || !valid_gp(gp) || Nc <= 0 || !Xc || !out || !work) return BO_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const long n = gp->n, M = gp->M, Kp = bo_kp(gp->d), ldn = bo_pad8(Nc);
    Bump b(work, work_bytes);
    double* baug = b.doubles(M * Kp * ldn);
    if (!b.ok) return BO_ERR_WORKSPACE;
    const BoKernelHyper h = hyper_of(gp);
    BO_TRY(bo_aug_cols(h, Xc, Nc, baug, Kp, ldn, Kp * ldn, st));
    return kmat(gp, gp->aug_train, (int)n, baug, ldn, Nc, out, ldn, n * ldn, 0, nullptr, st);
}

size_t bo_potrf_workspace_bytes(int n, int batch) { return sizeof(double) * (size_t)bo_linalg_scratch_doubles(n, batch); }

int bo_potrf(bo_ctx* ctx, int n, int batch, double* A, long lda, long strideA, double* inv, void* work,
             size_t work_bytes, int32_t* info, void* stream) {
    if (!ctx || !A || !work || !info || n <= 0 || batch <= 0) return BO_ERR_INVALID;
    if (work_bytes < bo_potrf_workspace_bytes(n, batch)) return BO_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    BO_CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int32_t) * batch, st));
    BO_TRY(bo_potrf_blocked(A, lda, strideA, n, batch, inv, lda, strideA, (double*)work, info, st));
    if (inv) BO_TRY(bo_trtri_blocked(A, lda, strideA, n, batch, inv, lda, strideA, (double*)work, st));
    return BO_OK;
}

}  // extern "C"
