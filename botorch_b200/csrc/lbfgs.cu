// Device-resident batched projected L-BFGS for the multi-start candidate optimisation.
//
// The reference hands ONE joint problem over all restarts to SciPy's L-BFGS-B (generation/gen.py:46-298); the objective
// is a plain sum over restarts (gen.py:209), so the restarts are independent box-constrained problems of q*d
// variables.  Here every restart keeps its own limited-memory curvature model and the whole iterate stays in HBM: one
// CTA per restart runs the two-loop recursion on the free variables (projected quasi-Newton with an active set), a
// projected Armijo backtracking decides per restart, and the host only enqueues  propose -> fused forward+backward ->
// accept  rounds without synchronising.  All reductions are fixed-order (deterministic).
#include "../../include/botorch_b200.h"

#include "common.cuh"

namespace {

constexpr int LB_THREADS = 128;

__device__ __forceinline__ double block_sum(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();  // red may still be read by the previous call
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double s = 0.0;
    for (int w = 0; w < LB_THREADS / 32; w++) s += red[w];
    return s;
}
__device__ __forceinline__ double block_max(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double s = red[0];
    for (int w = 1; w < LB_THREADS / 32; w++) s = fmax(s, red[w]);
    return s;
}

__device__ __forceinline__ double clip(double v, double lo, double hi) { return fmin(fmax(v, lo), hi); }

// state: 0 = last trial accepted (or first round): update the curvature model and compute a new direction
//        1 = last trial rejected: same direction, halved step
//        2 = converged / stopped: the trial point is the current point
//        3 = line search exhausted: drop the model, new (steepest descent) direction, no curvature update
__global__ void __launch_bounds__(LB_THREADS) lbfgs_propose_kernel(const bo_lbfgs s) {
    extern __shared__ double sm[];  // q[D], coefficient a[m]
    __shared__ double red[LB_THREADS / 32];
    const int b = blockIdx.x, tid = threadIdx.x, D = s.D, m = s.m;
    double* qv = sm;
    double* acoef = sm + D;
    const long off = (long)b * D;
    const double *x = s.x + off, *g = s.g + off;
    double* p = s.p + off;
    double* xt = s.x_trial + off;
    const int st = s.state[b];
    if (st == 2) {
        for (int i = tid; i < D; i += LB_THREADS) xt[i] = x[i];
        return;
    }
    if (st == 0 || st == 3) {
        double* Sb = s.S + (long)b * m * D;
        double* Yb = s.Y + (long)b * m * D;
        double* rho = s.rho + (long)b * m;
        int count = s.hist_count[b], head = s.hist_head[b];
        if (st == 0 && s.n_iter[b] > 0) {  // curvature pair of the step just accepted
            const double *xp = s.x_prev + off, *gp = s.g_prev + off;
            double sy = 0.0, yy = 0.0;
            for (int i = tid; i < D; i += LB_THREADS) {
                const double si = x[i] - xp[i], yi = g[i] - gp[i];
                sy += si * yi;
                yy += yi * yi;
            }
            sy = block_sum(sy, red);
            yy = block_sum(yy, red);
            if (sy > 1e-10 * yy && yy > 0.0) {
                for (int i = tid; i < D; i += LB_THREADS) {
                    Sb[(long)head * D + i] = x[i] - xp[i];
                    Yb[(long)head * D + i] = g[i] - gp[i];
                }
                if (tid == 0) rho[head] = 1.0 / sy;
                head = (head + 1) % m;
                count = count < m ? count + 1 : m;
                __syncthreads();
            }
        }
        // free part of the gradient: a variable sitting on a bound whose gradient pushes outward is held fixed
        double gnorm2 = 0.0;
        for (int i = tid; i < D; i += LB_THREADS) {
            const double lo = s.lower[i], hi = s.upper[i];
            const bool act = (x[i] <= lo && g[i] > 0.0) || (x[i] >= hi && g[i] < 0.0);
            const double v = act ? 0.0 : g[i];
            qv[i] = v;
            gnorm2 += v * v;
        }
        gnorm2 = block_sum(gnorm2, red);
        // two-loop recursion over the stored pairs, newest first (fixed variables carry q = 0 and stay 0 in the result)
        for (int k = 0; k < count; k++) {
            const int j = (head - 1 - k + 2 * m) % m;
            double d = 0.0;
            for (int i = tid; i < D; i += LB_THREADS) d += Sb[(long)j * D + i] * qv[i];
            d = block_sum(d, red) * rho[j];
            if (tid == 0) acoef[j] = d;
            for (int i = tid; i < D; i += LB_THREADS) qv[i] -= d * Yb[(long)j * D + i];
            __syncthreads();
        }
        double gamma = 1.0;
        if (count > 0) {
            const int j = (head - 1 + m) % m;
            double yy = 0.0;
            for (int i = tid; i < D; i += LB_THREADS) yy += Yb[(long)j * D + i] * Yb[(long)j * D + i];
            yy = block_sum(yy, red);
            gamma = 1.0 / (rho[j] * yy);
        }
        for (int i = tid; i < D; i += LB_THREADS) qv[i] *= gamma;
        __syncthreads();
        for (int k = count - 1; k >= 0; k--) {
            const int j = (head - 1 - k + 2 * m) % m;
            double d = 0.0;
            for (int i = tid; i < D; i += LB_THREADS) d += Yb[(long)j * D + i] * qv[i];
            d = block_sum(d, red) * rho[j];
            const double c = acoef[j] - d;
            for (int i = tid; i < D; i += LB_THREADS) qv[i] += c * Sb[(long)j * D + i];
            __syncthreads();
        }
        // p = -H g on the free variables; fall back to steepest descent when it is not a descent direction
        double gtd = 0.0;
        for (int i = tid; i < D; i += LB_THREADS) {
            const double lo = s.lower[i], hi = s.upper[i];
            const bool act = (x[i] <= lo && g[i] > 0.0) || (x[i] >= hi && g[i] < 0.0);
            const double v = act ? 0.0 : -qv[i];
            qv[i] = v;
            gtd += v * g[i];
        }
        gtd = block_sum(gtd, red);
        const bool descent = gtd < 0.0 && isfinite(gtd);
        for (int i = tid; i < D; i += LB_THREADS) {
            const double lo = s.lower[i], hi = s.upper[i];
            const bool act = (x[i] <= lo && g[i] > 0.0) || (x[i] >= hi && g[i] < 0.0);
            p[i] = descent ? qv[i] : (act ? 0.0 : -g[i]);
        }
        if (tid == 0) {
            s.hist_count[b] = descent ? count : 0;  // a useless model is dropped
            s.hist_head[b] = descent ? head : 0;
            // first step of a fresh model: unit step would be far too long (L-BFGS-B scales it by 1/|g| as well)
            const bool fresh = !descent || count == 0;
            s.alpha[b] = fresh ? fmin(1.0, 1.0 / sqrt(fmax(gnorm2, 1e-300))) : 1.0;
        }
        __syncthreads();
    }
    const double alpha = s.alpha[b];
    for (int i = tid; i < D; i += LB_THREADS) xt[i] = clip(x[i] + alpha * p[i], s.lower[i], s.upper[i]);
}

// Projected Armijo test of the evaluated trial point; on success the trial becomes the iterate.
__global__ void __launch_bounds__(LB_THREADS) lbfgs_accept_kernel(const bo_lbfgs s, const double* f_trial,
                                                                  const double* g_trial, double c1, double ftol,
                                                                  double pgtol, int maxiter) {
    __shared__ double red[LB_THREADS / 32];
    const int b = blockIdx.x, tid = threadIdx.x, D = s.D;
    if (s.state[b] == 2) return;
    const long off = (long)b * D;
    double *x = s.x + off, *g = s.g + off, *xp = s.x_prev + off, *gp = s.g_prev + off;
    const double *xt = s.x_trial + off, *gt = g_trial + off;
    double dec = 0.0, moved = 0.0;
    for (int i = tid; i < D; i += LB_THREADS) {
        const double dx = xt[i] - x[i];
        dec += g[i] * dx;
        moved = fmax(moved, fabs(dx));
    }
    dec = block_sum(dec, red);
    moved = block_max(moved, red);
    const double f0 = s.f[b], f1 = f_trial[b];
    // the clipped step must be a descent direction (after projection a quasi-Newton step can point uphill: dec > 0 would
    // let an INCREASE of f pass the Armijo test and then look like convergence); such a trial is rejected and the step
    // shrinks, ending in the steepest-descent restart below
    const bool ok = isfinite(f1) && dec < 0.0 && f1 <= f0 + c1 * dec && moved > 0.0;
    if (ok) {
        double pg = 0.0;
        for (int i = tid; i < D; i += LB_THREADS) {
            xp[i] = x[i];
            gp[i] = g[i];
            const double xn = xt[i], gn = gt[i];
            x[i] = xn;
            g[i] = gn;
            pg = fmax(pg, fabs(clip(xn - gn, s.lower[i], s.upper[i]) - xn));  // projected gradient, sup norm
        }
        pg = block_max(pg, red);
        if (tid == 0) {
            s.f[b] = f1;
            const int it = ++s.n_iter[b];
            const bool small_step = (f0 - f1) <= ftol * fmax(fmax(fabs(f0), fabs(f1)), 1.0);
            s.state[b] = (pg <= pgtol || small_step || it >= maxiter) ? 2 : 0;
        }
    } else if (tid == 0) {
        const double a = s.alpha[b] * 0.5;
        s.alpha[b] = a;
        if (moved == 0.0 || a < 1e-12) {  // no progress possible along this direction
            if (s.hist_count[b] > 0 && moved > 0.0) {
                s.hist_count[b] = 0;  // retry once from steepest descent with an empty model
                s.hist_head[b] = 0;
                s.state[b] = 3;
            } else {
                s.state[b] = 2;
            }
        } else {
            s.state[b] = 1;
        }
    }
}

// first evaluation: the iterate's own objective and gradient
__global__ void lbfgs_init_kernel(const bo_lbfgs s, const double* f0, const double* g0) {
    const int b = blockIdx.x, D = s.D;
    for (int i = threadIdx.x; i < D; i += blockDim.x) s.g[(long)b * D + i] = g0[(long)b * D + i];
    if (threadIdx.x == 0) {
        s.f[b] = f0[b];
        s.state[b] = isfinite(f0[b]) ? 0 : 2;
        s.n_iter[b] = 0;
        s.hist_count[b] = 0;
        s.hist_head[b] = 0;
        s.alpha[b] = 1.0;
    }
}

bool valid(const bo_lbfgs* s) {
    return s && s->B > 0 && s->D > 0 && s->m > 0 && s->m <= 64 && s->x && s->g && s->f && s->x_prev && s->g_prev && s->S &&
           s->Y && s->rho && s->hist_count && s->hist_head && s->p && s->alpha && s->x_trial && s->state && s->lower &&
           s->upper && s->n_iter;
}

}  // namespace

extern "C" {

int bo_lbfgs_init(bo_ctx* ctx, const bo_lbfgs* s, const double* f0, const double* g0, void* stream) {
    if (!ctx || !valid(s) || !f0 || !g0) return BO_ERR_INVALID;
    lbfgs_init_kernel<<<s->B, LB_THREADS, 0, (cudaStream_t)stream>>>(*s, f0, g0);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_lbfgs_propose(bo_ctx* ctx, const bo_lbfgs* s, void* stream) {
    if (!ctx || !valid(s)) return BO_ERR_INVALID;
    const size_t smem = sizeof(double) * ((size_t)s->D + s->m);
    if (smem > 200 * 1024) return BO_ERR_UNSUPPORTED;
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        BO_CUDA_TRY(cudaFuncSetAttribute(lbfgs_propose_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    lbfgs_propose_kernel<<<s->B, LB_THREADS, smem, (cudaStream_t)stream>>>(*s);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_lbfgs_accept(bo_ctx* ctx, const bo_lbfgs* s, const double* f_trial, const double* g_trial, double c1, double ftol,
                    double pgtol, int maxiter, void* stream) {
    if (!ctx || !valid(s) || !f_trial || !g_trial) return BO_ERR_INVALID;
    lbfgs_accept_kernel<<<s->B, LB_THREADS, 0, (cudaStream_t)stream>>>(*s, f_trial, g_trial, c1, ftol, pgtol, maxiter);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

}  // extern "C"
