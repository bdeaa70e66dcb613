// bo_kernel_matrix_static: the covariance-matrix entry point for callers that cannot hand the library a workspace
// (CUDA-graph capture, serving loops or harnesses that forbid cudaMalloc).  It has no counterpart in the reference.
// Same kernels and same results as bo_kernel_matrix (api.cu); the only
// difference is where the augmented candidate operand [M][Kp x pad8(Nc)] lives: in the file-scope device array below
// instead of caller-owned memory.  Nothing here calls cudaMalloc / cudaFree or any allocating routine.
#include "../../include/botorch_b200.h"

#include "cand.cuh"
#include "common.cuh"

// Every GPU that loads the library holds this array (512 KiB) until the process exits.
static __device__ __align__(256) double g_static_scratch[BO_STATIC_SCRATCH_DOUBLES];

namespace {

constexpr int MAX_DEVICES = 64;
// Device address of g_static_scratch per device ordinal.  In host code the NAME g_static_scratch is only a stand-in for
// the symbol, not an address a kernel can use: the address has to come from cudaGetSymbolAddress, resolved once per
// device with that device current.
double* g_scratch_addr[MAX_DEVICES] = {};

int static_scratch(double** out) {
    int dev = -1;
    BO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return BO_ERR_UNSUPPORTED;
    if (!g_scratch_addr[dev]) {
        double* p = nullptr;
        BO_CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_static_scratch));
        if (!p) return BO_ERR_CUDA;
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return BO_OK;
}

}  // namespace

extern "C" int bo_kernel_matrix_static(bo_ctx* ctx, const bo_gp* gp, long Nc, const double* Xc, double* out, void* stream) {
    if (!ctx || !gp || gp->n <= 0 || gp->d <= 0 || gp->d > BO_MAX_D || gp->M <= 0 || !gp->aug_train || !gp->inv_ls ||
        !gp->center || (gp->kernel_id != BO_KERNEL_RBF && gp->kernel_id != BO_KERNEL_MATERN52) || Nc <= 0 || !Xc || !out)
        return BO_ERR_INVALID;
    const long n = gp->n, M = gp->M, Kp = bo_round_up(gp->d + 2, 4), ldn = bo_round_up(Nc, 8);
    if (M * Kp * ldn > (long)BO_STATIC_SCRATCH_DOUBLES) return BO_ERR_STATIC_SCRATCH;
    // the output must live on the device whose copy of the array is used (the current device)
    cudaPointerAttributes attr;
    int dev = -1;
    BO_CUDA_TRY(cudaGetDevice(&dev));
    BO_CUDA_TRY(cudaPointerGetAttributes(&attr, out));
    if (attr.type != cudaMemoryTypeDevice || attr.device != dev) return BO_ERR_INVALID;
    double* baug = nullptr;
    BO_TRY(static_scratch(&baug));
    cudaStream_t st = (cudaStream_t)stream;
    BoKernelHyper h;
    h.kernel_id = gp->kernel_id;
    h.d = gp->d;
    h.M = gp->M;
    h.inv_ls = gp->inv_ls;
    h.outputscale = gp->outputscale;
    h.center = gp->center;
    // bo_aug_cols writes all Kp x ldn entries of every model's block (padding columns and rows included), so nothing in
    // the shared array has to be cleared between calls even though it keeps the previous call's contents.
    BO_TRY(bo_aug_cols(h, Xc, Nc, baug, Kp, ldn, Kp * ldn, st));
    BoGemm g{};
    g.A = gp->aug_train; g.lda = Kp; g.strideA = n * Kp;
    g.B = baug; g.ldb = ldn; g.strideB = Kp * ldn;
    g.C = out; g.ldc = ldn; g.strideC = n * ldn;
    g.M = (int)n; g.N = (int)Nc; g.K = (int)Kp; g.batch = gp->M;
    g.epi = gp->kernel_id == BO_KERNEL_RBF ? BO_EPI_RBF : BO_EPI_MATERN52;
    g.os = gp->outputscale;
    g.ldc2 = ldn; g.strideC2 = n * ldn;
    return bo_gemm_launch(g, st);
}
