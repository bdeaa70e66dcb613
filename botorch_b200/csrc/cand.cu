// Per-candidate kernels of the fused qLogNEI path: Schur block + q x q Cholesky (with the
// psd_safe_cholesky jitter schedule), reparameterised samples f = mu + L z, log_fatplus, fatmax over q,
// logmeanexp over MC samples, and the hand-derived reverse pass of all of it.
// Reference arithmetic: utils/low_rank.py:118-162, acquisition/logei.py:509-534,
// utils/safe_math.py:149-187 (_inf_max_helper), :293-352 (log_fatplus, fatmax), :209-223 (logmeanexp).
#include "cand.cuh"
#include <cstdlib>

#include <cfloat>
#include <cmath>

namespace {

constexpr int CAND_THREADS = 128;
#ifndef BO_CAND_MINB4
#define BO_CAND_MINB4 4  // resident CTAs per SM the q <= 4 kernels are compiled for; 6 and 8 (80 / 64 registers, spills) measured within noise at C4
#endif
constexpr int PANEL_ROWS = 64;

// ---------------------------------------------------------------------------------------------
// covariance functions of a squared scaled distance
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double kern_value(int kernel_id, double r2, double os) {
    r2 = r2 > 0.0 ? r2 : 0.0;
    if (kernel_id == BO_KERNEL_RBF) return os * exp(-0.5 * r2);
    const double sqrt5 = 2.23606797749978969641;
    const double r = sqrt(r2 > 1e-30 ? r2 : 1e-30);
    return os * ((sqrt5 * r + 1.0) + (5.0 / 3.0) * (r * r)) * exp(-sqrt5 * r);
}
// g(r) with  d k / d x_a = -g * (x_a - x_b) / l^2
__device__ __forceinline__ double kern_dfactor(int kernel_id, double r2, double os) {
    r2 = r2 > 0.0 ? r2 : 0.0;
    if (kernel_id == BO_KERNEL_RBF) return os * exp(-0.5 * r2);
    if (r2 < 1e-30) return 0.0;  // gpytorch clamps the distance: no gradient below the clamp
    const double sqrt5 = 2.23606797749978969641;
    const double r = sqrt(r2);
    return os * (5.0 / 3.0) * (1.0 + sqrt5 * r) * exp(-sqrt5 * r);
}

// ---------------------------------------------------------------------------------------------
// reductions of safe_math.py
// ---------------------------------------------------------------------------------------------
// log(tau*(softplus(u) + 0.1/(1+u^2))), u = z/tau   (torch softplus: identity above threshold 20)
__device__ __forceinline__ double log_fatplus(double z, double tau, double inv_tau) {
    const double u = z * inv_tau;
    // log(tau * (softplus(u) + 0.1 / (1 + u^2))) as the reference evaluates it.  tau = 1e-6 puts |u| in the 1e4..1e6 range
    // for the bulk of the samples: softplus is u above torch's threshold 20 and exp(u) underflows to exactly 0 below
    // -746, so both tails share ONE division and ONE log -- a warp whose lanes sit on both sides of zero executes this
    // code once instead of one tail after the other; only -746 <= u <= 20 pays for exp and log1p.
    double sp = (u > 20.0) ? u : 0.0;
    if (u >= -746.0 && u <= 20.0) sp = log1p(exp(u));
    return log(tau * (sp + 0.1 * (1.0 / fma(u, u, 1.0))));
}
// d log_fatplus / dz
__device__ __forceinline__ double dlog_fatplus(double z, double tau, double inv_tau) {
    const double u = z * inv_tau;
    double sp = (u > 20.0) ? u : 0.0, sig = (u > 20.0) ? 1.0 : 0.0;
    if (u >= -746.0 && u <= 20.0) {
        const double e = exp(u);
        sp = log1p(e);
        sig = e / (e + 1.0);
    }
    const double c = 1.0 / fma(u, u, 1.0);
    const double val = sp + 0.1 * c;
    if (!(val > 0.0)) return 0.0;
    return (sig - 0.2 * u * c * c) / (val * tau);
}
__device__ __forceinline__ double pareto2(double u) { return 2.0 / (2.0 + 2.0 * u + u * u); }

// fat = False (acquisition/logei.py:122,532): log_softplus (utils/safe_math.py:226-247, float64 cut-offs -35 / 32) for the
// improvement and smooth_amax = tau * logsumexp(x / tau) (safe_math.py:250-273) over the q points
__device__ __forceinline__ double log_softplus_ref(double z, double tau, double inv_tau) {
    const double u = z * inv_tau;
    if (!(u > -35.0)) return u + log(tau);
    return log((u > 32.0) ? z : tau * log1p(exp(u)));
}
__device__ __forceinline__ double dlog_softplus_ref(double z, double tau, double inv_tau) {
    const double u = z * inv_tau;
    if (!(u > -35.0)) return inv_tau;
    if (u > 32.0) return 1.0 / z;
    const double e = exp(u);
    return (e / (1.0 + e)) / (tau * log1p(e));
}
// log-improvement of one sample: reduction BO_REDUCE_LOG (fat) or BO_REDUCE_LOG_SOFT
__device__ __forceinline__ double log_improvement(int reduction, double z, double tau, double inv_tau) {
    return reduction == BO_REDUCE_LOG_SOFT ? log_softplus_ref(z, tau, inv_tau) : log_fatplus(z, tau, inv_tau);
}
__device__ __forceinline__ double dlog_improvement(int reduction, double z, double tau, double inv_tau) {
    return reduction == BO_REDUCE_LOG_SOFT ? dlog_softplus_ref(z, tau, inv_tau) : dlog_fatplus(z, tau, inv_tau);
}

// value and derivative in one pass (the forward kernel with bo_nei::local_derivs): the same operations in the same order
// as the two functions above, sharing exp(u), 1 / (1 + u^2) and the softplus -- the derivative costs one more division
__device__ __forceinline__ double log_improvement_vg(int reduction, double z, double tau, double inv_tau, double& d) {
    const double u = z * inv_tau;
    if (reduction == BO_REDUCE_LOG_SOFT) {
        if (!(u > -35.0)) { d = inv_tau; return u + log(tau); }
        if (u > 32.0) { d = 1.0 / z; return log(z); }
        const double e = exp(u), l1 = log1p(e);
        d = (e / (1.0 + e)) / (tau * l1);
        return log(tau * l1);
    }
    double sp = (u > 20.0) ? u : 0.0, sig = (u > 20.0) ? 1.0 : 0.0;
    if (u >= -746.0 && u <= 20.0) {
        const double e = exp(u);
        sp = log1p(e);
        sig = e / (e + 1.0);
    }
    const double c = 1.0 / fma(u, u, 1.0);
    const double val = sp + 0.1 * c;
    d = (val > 0.0) ? (sig - 0.2 * u * c * c) / (val * tau) : 0.0;
    return log(tau * val);
}

// fatmax over x[0..q) (alpha = 2) with the infinity handling of _inf_max_helper.  QT >= q is a compile-time bound:
// every loop is fully unrolled with a (j < q) guard so that the per-sample arrays stay in registers.
// The temperature enters as a multiplication by 1/tau (one rounding away from the reference's division: 1e-16 relative
// on u, far inside the 1e-9 parity bar) -- an FP64 division costs as much as the rest of the per-element work.
template <int QT>
__device__ __forceinline__ double fatmax_q(const double (&x)[QT], int q, double tau) {
    double M = x[0];
#pragma unroll
    for (int j = 1; j < QT; j++)
        if (j < q) M = fmax(M, x[j]);  // NaN-ignoring like amax is not needed: NaN flagged earlier
    if (isinf(M)) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < QT; j++)
            if (j < q && x[j] == M) s += x[j];
        return s;
    }
    const double inv_tau = 1.0 / tau;
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < QT; j++)
        if (j < q) sum += pareto2((M - x[j]) * inv_tau);
    return M + tau * log(sum);
}
// fatmax value (returned) and its gradient w.r.t. x (g[0..q)) in one pass over the q Pareto terms
template <int QT>
__device__ __forceinline__ double fatmax_q_grad(const double (&x)[QT], int q, double tau, double (&g)[QT]) {
    double M = x[0];
#pragma unroll
    for (int j = 1; j < QT; j++)
        if (j < q) M = fmax(M, x[j]);
    if (isinf(M)) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < QT; j++) {
            g[j] = (j < q && x[j] == M) ? 1.0 : 0.0;
            if (j < q && x[j] == M) s += x[j];
        }
        return s;
    }
    const double inv_tau = 1.0 / tau;
    double sum = 0.0;
    int nties = 0;
#pragma unroll
    for (int j = 0; j < QT; j++) {
        g[j] = 0.0;
        if (j < q) {
            const double u = (M - x[j]) * inv_tau;
            const double p = pareto2(u);
            sum += p;
            nties += (x[j] == M);
            g[j] = isinf(u) ? 0.0 : p * p * (1.0 + u);
        }
    }
    const double inv_sum = 1.0 / sum;
    double wsum = 0.0;
#pragma unroll
    for (int j = 0; j < QT; j++) {
        g[j] *= inv_sum;
        wsum += g[j];
    }
    const double rest = (1.0 - wsum) / nties;  // amax subgradient, split evenly among ties (torch.amax)
#pragma unroll
    for (int j = 0; j < QT; j++)
        if (j < q && x[j] == M) g[j] += rest;
    return M + tau * log(sum);
}

// smooth_amax over x[0..q): tau * logsumexp(x / tau) with the infinity handling of _inf_max_helper (safe_math.py:149-187)
template <int QT>
__device__ __forceinline__ double smoothmax_q(const double (&x)[QT], int q, double tau) {
    const double inv_tau = 1.0 / tau;
    double M = x[0] * inv_tau;
#pragma unroll
    for (int j = 1; j < QT; j++)
        if (j < q) M = fmax(M, x[j] * inv_tau);
    if (isinf(M)) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < QT; j++)
            if (j < q && x[j] * inv_tau == M) s += x[j] * inv_tau;
        return s * tau;
    }
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < QT; j++)
        if (j < q) sum += exp(x[j] * inv_tau - M);
    return (M + log(sum)) * tau;
}
// value and gradient (softmax weights; unit gradient on the infinite maxima)
template <int QT>
__device__ __forceinline__ double smoothmax_q_grad(const double (&x)[QT], int q, double tau, double (&g)[QT]) {
    const double inv_tau = 1.0 / tau;
    double M = x[0] * inv_tau;
#pragma unroll
    for (int j = 1; j < QT; j++)
        if (j < q) M = fmax(M, x[j] * inv_tau);
    if (isinf(M)) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < QT; j++) {
            const bool hit = j < q && x[j] * inv_tau == M;
            g[j] = hit ? 1.0 : 0.0;
            if (hit) s += x[j] * inv_tau;
        }
        return s * tau;
    }
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < QT; j++) {
        g[j] = (j < q) ? exp(x[j] * inv_tau - M) : 0.0;
        sum += g[j];
    }
    const double inv_sum = 1.0 / sum;
#pragma unroll
    for (int j = 0; j < QT; j++) g[j] *= inv_sum;
    return (M + log(sum)) * tau;
}
template <int QT>
__device__ __forceinline__ double qmax_log(int reduction, const double (&x)[QT], int q, double tau) {
    return reduction == BO_REDUCE_LOG_SOFT ? smoothmax_q<QT>(x, q, tau) : fatmax_q<QT>(x, q, tau);
}
template <int QT>
__device__ __forceinline__ double qmax_log_grad(int reduction, const double (&x)[QT], int q, double tau, double (&g)[QT]) {
    return reduction == BO_REDUCE_LOG_SOFT ? smoothmax_q_grad<QT>(x, q, tau, g) : fatmax_q_grad<QT>(x, q, tau, g);
}

// Sum P per-lane values over the 32 lanes of a warp with ~P shuffles instead of 5 P: at every butterfly step each lane
// keeps one half of its values and sends the other half to its partner, so the value count halves while the lane
// groups that share a value shrink.  idx[] tracks which original value a slot holds.  The order of the additions is
// fixed by the lane layout, so the result is deterministic.  After the five steps every lane owns NOUT slots.
template <int N>
__device__ __forceinline__ void halve_step(double* v, int* idx, int lane, int off) {
    constexpr int H = (N + 1) / 2;
    const bool hi = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < N - H; i++) {
        const double send = hi ? v[i] : v[i + H];
        const double keep = hi ? v[i + H] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        idx[i] = hi ? idx[i + H] : idx[i];
    }
    if constexpr (N & 1) v[H - 1] += __shfl_xor_sync(0xffffffffu, v[H - 1], off);
}
template <int P>
struct WarpMultiReduce {
    static constexpr int N1 = (P + 1) / 2, N2 = (N1 + 1) / 2, N3 = (N2 + 1) / 2, N4 = (N3 + 1) / 2, NOUT = (N4 + 1) / 2;
    // v[P] in, out[idx] = warp sum for every value (each written by the lanes that end up owning it; duplicates agree)
    static __device__ __forceinline__ void run(double* v, int lane, double* out) {
        int idx[P];
#pragma unroll
        for (int i = 0; i < P; i++) idx[i] = i;
        halve_step<P>(v, idx, lane, 16);
        halve_step<N1>(v, idx, lane, 8);
        halve_step<N2>(v, idx, lane, 4);
        halve_step<N3>(v, idx, lane, 2);
        halve_step<N4>(v, idx, lane, 1);
#pragma unroll
        for (int i = 0; i < NOUT; i++) out[idx[i]] = v[i];
    }
};

__device__ __forceinline__ double block_reduce_max(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    double r = red[0];
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) r = fmax(r, red[i]);
    return r;
}
__device__ __forceinline__ double block_reduce_sum(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    double r = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++) r += red[i];
    return r;
}

// ---------------------------------------------------------------------------------------------
// augmented operands
// ---------------------------------------------------------------------------------------------
__global__ void aug_rows_kernel(BoKernelHyper h, const double* X, long npts, double* out, long Kp, long strideOut) {
    const int z = blockIdx.y;
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= npts) return;
    const double* il = h.inv_ls + (long)z * h.d;
    double* o = out + (long)z * strideOut + i * Kp;
    double nrm = 0.0;
    for (int t = 0; t < h.d; t++) {
        const double v = (X[i * h.d + t] - h.center[t]) * il[t];
        o[t] = -2.0 * v;
        nrm += v * v;
    }
    o[h.d] = nrm;
    o[h.d + 1] = 1.0;
    for (long t = h.d + 2; t < Kp; t++) o[t] = 0.0;
}

__global__ void aug_cols_kernel(BoKernelHyper h, const double* X, long npts, double* out, long Kp, long ldo,
                                long strideOut) {
    const int z = blockIdx.y;
    const long c = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (c >= ldo) return;
    double* o = out + (long)z * strideOut + c;
    if (c >= npts) {  // padded columns: all-zero operand
        for (long t = 0; t < Kp; t++) o[t * ldo] = 0.0;
        return;
    }
    const double* il = h.inv_ls + (long)z * h.d;
    double nrm = 0.0;
    for (int t = 0; t < h.d; t++) {
        const double v = (X[c * h.d + t] - h.center[t]) * il[t];
        o[(long)t * ldo] = v;
        nrm += v * v;
    }
    o[(long)h.d * ldo] = 1.0;
    o[(long)(h.d + 1) * ldo] = nrm;
    for (long t = h.d + 2; t < Kp; t++) o[t * ldo] = 0.0;
}

}  // namespace

namespace {

__device__ __forceinline__ void pair_decode(int pair, int& j, int& k) {
    j = (int)((sqrt(8.0 * pair + 1.0) - 1.0) * 0.5);
    while ((j + 1) * (j + 2) / 2 <= pair) j++;
    while (j * (j + 1) / 2 > pair) j--;
    k = pair - j * (j + 1) / 2;
}

// G[j][k] = sum_i Mtx[i][c0+j] * Mtx[i][c0+k]  for the q columns of one t-batch element.
// Deterministic: fixed row -> group assignment and fixed-order final sum.
__device__ void gram_block(const double* __restrict__ Mtx, long ld, int rows, int c0, int q, double* panel,
                           double* part, double* G) {
    const int T = blockDim.x, tid = threadIdx.x;
    const int P = q * (q + 1) / 2;
    const int ngroups = (T / P) > 1 ? (T / P) : 1;
    for (int i = tid; i < P * ngroups; i += T) part[i] = 0.0;
    __syncthreads();
    for (int r0 = 0; r0 < rows; r0 += PANEL_ROWS) {
        const int pr = (rows - r0 < PANEL_ROWS) ? rows - r0 : PANEL_ROWS;
        for (int idx = tid; idx < pr * q; idx += T) {
            const int r = idx / q, j = idx % q;
            panel[idx] = Mtx[(long)(r0 + r) * ld + c0 + j];
        }
        __syncthreads();
        for (int slot = tid; slot < P * ngroups; slot += T) {
            const int grp = slot / P, pair = slot % P;
            int j, k;
            pair_decode(pair, j, k);
            double acc = 0.0;
            for (int r = grp; r < pr; r += ngroups) acc += panel[r * q + j] * panel[r * q + k];
            part[slot] += acc;
        }
        __syncthreads();
    }
    for (int pair = tid; pair < P; pair += T) {
        int j, k;
        pair_decode(pair, j, k);
        double s = 0.0;
        for (int grp = 0; grp < ngroups; grp++) s += part[grp * P + pair];
        G[j * q + k] = s;
        G[k * q + j] = s;
    }
    __syncthreads();
}

// Register-blocked variant for q <= QT (QT in {4, 8}): every thread walks its own rows (r = tid, tid + T, ...), keeps all
// QT(QT+1)/2 products in registers, then a fixed xor-shuffle tree and a fixed-order sum over warps: deterministic,
// no shared-memory staging and ~10x fewer instructions than gram_block for the q = 8 case of config C3.
template <int QT>
__device__ void gram_block_reg(const double* __restrict__ Mtx, long ld, int rows, int c0, int q, double* red, double* G) {
    constexpr int P = QT * (QT + 1) / 2;
    const int T = blockDim.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    double acc[P];
#pragma unroll
    for (int p = 0; p < P; p++) acc[p] = 0.0;
    const bool vec = (q == QT) && ((QT & 1) == 0) && ((c0 & 1) == 0) && ((ld & 1) == 0);
#pragma unroll 2
    for (int r = tid; r < rows; r += T) {
        double v[QT];
        const double* row = Mtx + (long)r * ld + c0;
        if (vec) {
#pragma unroll
            for (int j = 0; j < QT; j += 2) {
                const double2 t2 = *reinterpret_cast<const double2*>(row + j);
                v[j] = t2.x;
                v[j + 1] = t2.y;
            }
        } else {
#pragma unroll
            for (int j = 0; j < QT; j++) v[j] = (j < q) ? row[j] : 0.0;
        }
        int p = 0;
#pragma unroll
        for (int j = 0; j < QT; j++)
#pragma unroll
            for (int k = 0; k <= j; k++) acc[p++] += v[j] * v[k];
    }
    WarpMultiReduce<P>::run(acc, lane, red + warp * P);
    __syncthreads();
    for (int p = tid; p < P; p += T) {
        int j = 0;
        while ((j + 1) * (j + 2) / 2 <= p) j++;
        const int k = p - j * (j + 1) / 2;
        double s_ = 0.0;
        for (int w = 0; w < (T >> 5); w++) s_ += red[w * P + p];
        if (j < q) {
            G[j * q + k] = s_;
            G[k * q + j] = s_;
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void gram_any(const double* Mtx, long ld, int rows, int c0, int q, double* panel, double* part,
                                         double* G) {
    if (q <= 4) gram_block_reg<4>(Mtx, ld, rows, c0, q, part, G);
    else if (q <= 8) gram_block_reg<8>(Mtx, ld, rows, c0, q, part, G);
    else gram_block(Mtx, ld, rows, c0, q, panel, part, G);
}

// in-place lower Cholesky of the q x q matrix A (row-major); returns false on a non-positive pivot
__device__ bool chol_small(double* A, int q) {
    for (int j = 0; j < q; j++) {
        double d = A[j * q + j];
        for (int k = 0; k < j; k++) d -= A[j * q + k] * A[j * q + k];
        if (!(d > 0.0)) return false;
        d = sqrt(d);
        A[j * q + j] = d;
        for (int i = j + 1; i < q; i++) {
            double s = A[i * q + j];
            for (int k = 0; k < j; k++) s -= A[i * q + k] * A[j * q + k];
            A[i * q + j] = s / d;
        }
    }
    return true;
}

// Cholesky of the q x q matrix whose row `lane` sits in r[] (lanes >= q idle), right-looking, one pivot per step: the
// pivot and the column entries travel by shuffle, so the q steps cost a few hundred cycles instead of a serial
// O(q^3) chain through shared memory.  The subtraction order per entry (k ascending) is the one of chol_small.
// Returns false (uniformly) on a non-positive pivot.  On success r[0..lane] is row `lane` of L.
template <int QT>
__device__ __forceinline__ bool chol_warp(double (&r)[QT], int q, int lane) {
#pragma unroll
    for (int j = 0; j < QT; j++) {
        if (j < q) {
            const double d = __shfl_sync(0xffffffffu, r[j], j);
            if (!(d > 0.0)) return false;
            const double sd = sqrt(d);
            const double lij = (lane == j) ? sd : r[j] / sd;
            r[j] = lij;
#pragma unroll
            for (int k = j + 1; k < QT; k++) {
                if (k < q) {
                    const double lkj = __shfl_sync(0xffffffffu, lij, k);
                    if (lane >= k) r[k] -= lij * lkj;
                }
            }
        }
    }
    return true;
}

// scratch of the gram reductions: T (generic path, q > 8) or one 36-entry slot per warp (register-blocked path)
__host__ __device__ inline int part_doubles(int q, int T) {
    const int P = q * (q + 1) / 2;
    const int a = T > P ? T : P, b = (T / 32) * 36;
    return a > b ? a : b;
}
struct CandSmem {
    double *panel, *part, *GV, *GB, *Lq, *mu, *fm;
};
__device__ __forceinline__ CandSmem carve(double* sm, int q, int T, int S) {
    const int P = q * (q + 1) / 2;
    CandSmem c;
    c.panel = sm;
    c.part = c.panel + PANEL_ROWS * q;
    c.GV = c.part + part_doubles(q, T);
    c.GB = c.GV + q * q;
    c.Lq = c.GB + q * q;
    c.mu = c.Lq + q * q;
    c.fm = c.mu + q;
    (void)S;
    return c;
}
size_t cand_smem_bytes(int q, int T, int S) {
    const int P = q * (q + 1) / 2;
    (void)P;
    return sizeof(double) * ((size_t)PANEL_ROWS * q + part_doubles(q, T) + 3 * q * q + q + S);
}

__device__ __forceinline__ double pair_r2(const double* X, int d, const double* il, int a, int b) {
    double r2 = 0.0;
    for (int t = 0; t < d; t++) {
        const double v = (X[(long)a * d + t] - X[(long)b * d + t]) * il[t];
        r2 += v * v;
    }
    return r2;
}

template <int QT, bool PRE>
__global__ void __launch_bounds__(CAND_THREADS, (QT <= 4 ? BO_CAND_MINB4 : (QT <= 8 ? 4 : 1))) cand_fwd_kernel(const BoCandArgs a) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red[CAND_THREADS / 32];
    __shared__ int s_info;
    const int q = a.q, b = blockIdx.x, z = blockIdx.y, tid = threadIdx.x, T = blockDim.x;
    const CandSmem c = carve(sm, q, T, a.S);
    const int c0 = b * q;
    const double* V = a.V ? a.V + (long)z * a.strideV : nullptr;
    const double* BL = a.BL + (long)z * a.strideBL;
    double* F = a.F + (long)z * a.strideF;
    // mean row: row n of V, or (combined-root path) its own row produced by the sample GEMM
    const double* mu_row = a.mu_row ? a.mu_row + (long)z * a.strideMu : V + (long)a.n * a.ldn;

    if (V) gram_any(V, a.ldn, a.n, c0, q, c.panel, c.part, c.GV);
    else {
        for (int i = tid; i < q * q; i += T) c.GV[i] = 0.0;
        __syncthreads();
    }
    if (a.gram != nullptr) {
        // the GEMM that produced BL left per-row-tile partial Grams of every 8-column group: sum them in tile order
        const long groups = ((long)a.B * q + 7) / 8;
        const double* Gp = a.gram + (long)z * a.strideGram + (long)(c0 / 8) * 64;
        const int o8 = c0 % 8;
        // warp w sums the tiles w, w + W, w + 2W, ... (independent loads in flight), then the W partials are added in warp
        // order: a fixed summation tree
        const int warp = tid >> 5, lane = tid & 31, W = T >> 5;
        for (int idx = lane; idx < q * q; idx += 32) {
            const int j = idx / q, k = idx % q;
            const int hi = j > k ? j : k, lo = j > k ? k : j;  // one triangle: the result is exactly symmetric
            const int pr = (o8 + hi) * 8 + o8 + lo;
            double s_ = 0.0;
#pragma unroll 4
            for (int rt = warp; rt < a.gram_tiles; rt += W) s_ += Gp[(long)rt * groups * 64 + pr];
            c.panel[warp * q * q + idx] = s_;
        }
        __syncthreads();
        for (int idx = tid; idx < q * q; idx += T) {
            double s_ = 0.0;
            for (int w2 = 0; w2 < W; w2++) s_ += c.panel[w2 * q * q + idx];
            c.GB[idx] = s_;
        }
        __syncthreads();
    } else {
        gram_any(BL, a.ldn, a.n_b, c0, q, c.panel, c.part, c.GB);
    }

    const double ys2 = a.y_std * a.y_std;
    const double os = a.h.outputscale ? a.h.outputscale[z] : 1.0;
    const double* il = a.h.inv_ls + (long)z * a.h.d;
    for (int idx = tid; idx < q * q; idx += T) {
        const int j = idx / q, k = idx % q;
        const double kv = kern_value(a.h.kernel_id, pair_r2(a.X, a.h.d, il, c0 + j, c0 + k), os);
        c.Lq[idx] = ys2 * (kv - c.GV[idx]) - c.GB[idx];  // Schur block  br - bl_chol bl_chol^T
    }
    if (tid < q) c.mu[tid] = a.y_mean + a.y_std * (a.mean_const[z] + mu_row[c0 + tid]);
    __syncthreads();
    if (tid < 32) {  // psd_safe_cholesky on warp 0: plain attempt, then jitter 1e-8 * 10^i, i = 0..5; lane i owns row i
        const int lane = tid;
        double row0[QT], r[QT];
        bool has_nan = false;
#pragma unroll
        for (int k = 0; k < QT; k++) {
            row0[k] = (lane < q && k < q) ? c.Lq[lane * q + k] : 0.0;
            has_nan |= isnan(row0[k]);
        }
        has_nan = __any_sync(0xffffffffu, has_nan);
        int code = -2;
        if (has_nan) {
            code = -1;
        } else {
            double jit = a.schur_fixed ? a.schur_jitter : 0.0;
            const int last = a.schur_fixed ? 0 : 6;  // fixed jitter (joint-factorisation fallback): one attempt
            for (int attempt = 0; attempt <= last; attempt++) {
#pragma unroll
                for (int k = 0; k < QT; k++) r[k] = row0[k] + ((k == lane) ? jit : 0.0);
                if (chol_warp<QT>(r, q, lane)) { code = attempt; break; }
                jit = (attempt == 0) ? 1e-8 : jit * 10.0;
            }
        }
        if (lane < q) {
#pragma unroll
            for (int k = 0; k < QT; k++)
                if (k < q) c.Lq[lane * q + k] = (code >= 0 && k <= lane) ? r[k] : 0.0;
        }
        if (lane == 0) s_info = code;
    }
    __syncthreads();

    int bad = 0;
    double local_max = -INFINITY;
    const bool plain = (a.reduction == BO_REDUCE_PLAIN);
    const double inv_tau_relu = 1.0 / a.tau_relu;
    // The sample rows are read one iteration ahead (BO_CAND_PREFETCH): the store to row s at the end of an iteration would
    // otherwise keep the compiler from hoisting the load of row s + T above it (same pointer), and every iteration would
    // start by waiting out a full global-load latency.
    double fnext[QT], znext[QT], bnext = 0.0;
#pragma unroll
    for (int j = 0; j < QT; j++) {
        fnext[j] = (j < q && tid < a.S && !a.f_zero) ? F[(long)tid * a.ldn + c0 + j] : 0.0;
        znext[j] = (j < q && tid < a.S) ? a.Zq[(long)tid * q + j] : 0.0;
    }
    if (tid < a.S) bnext = a.best_f[(long)z * a.Sp + tid];
    for (int s = tid; s < a.S; s += T) {
        double li[QT], zq[QT], fcur[QT];
        const double bf = bnext;
#pragma unroll
        for (int j = 0; j < QT; j++) { fcur[j] = fnext[j]; zq[j] = znext[j]; }
        if (s + T < a.S) {
#pragma unroll
            for (int j = 0; j < QT; j++) {
                fnext[j] = (j < q && !a.f_zero) ? F[(long)(s + T) * a.ldn + c0 + j] : 0.0;
                znext[j] = (j < q) ? a.Zq[(long)(s + T) * q + j] : 0.0;
            }
            bnext = a.best_f[(long)z * a.Sp + s + T];
        }
#pragma unroll
        for (int j = 0; j < QT; j++) {
            li[j] = 0.0;
            if (j < q) {
                double v = fcur[j] + c.mu[j];
#pragma unroll
                for (int k = 0; k < QT; k++)
                    if (k <= j) v += c.Lq[j * q + k] * zq[k];
                if (!PRE) F[(long)s * a.ldn + c0 + j] = v;
                if (!isfinite(v)) bad = 1;
                if (PRE) li[j] = log_improvement_vg(a.reduction, a.obj_w * v - bf, a.tau_relu, inv_tau_relu, fcur[j]);  // fcur <- d li / d z
                else li[j] = plain ? fmax(a.obj_w * v - bf, 0.0) : log_improvement(a.reduction, a.obj_w * v - bf, a.tau_relu, inv_tau_relu);
            }
        }
        double m;
        if (PRE) {
            // bo_nei::local_derivs: everything of the reverse pass that depends on this sample alone is formed here, while
            // li is in registers -- d qmax_s / d f[s][j] = (fatmax weight) x (d log_fatplus) -- and replaces the sample in
            // its row; the reverse kernel multiplies by the logmeanexp weight exp(m_s - lse) and the upstream gradient.
            double gm[QT];
            m = qmax_log_grad<QT>(a.reduction, li, q, a.tau_max, gm);
#pragma unroll
            for (int j = 0; j < QT; j++) {
                if (j < q) F[(long)s * a.ldn + c0 + j] = gm[j] * (isinf(li[j]) ? 0.0 : fcur[j]);
            }
            a.msave[((long)z * a.B + b) * a.Sp + s] = m;
        } else if (plain) {  // (obj - best_f).clamp_min(0).amax(q)   (monte_carlo.py:410,536-547)
            m = li[0];
#pragma unroll
            for (int j = 1; j < QT; j++)
                if (j < q) m = fmax(m, li[j]);
        } else {
            m = qmax_log<QT>(a.reduction, li, q, a.tau_max);
        }
        c.fm[s] = m;
        local_max = fmax(local_max, m);
    }
    const int anybad = __syncthreads_or(bad);
    double lse, val;
    if (plain) {  // mean over the MC samples; lse keeps the plain sum
        double part = 0.0;
        for (int s = tid; s < a.S; s += T) part += c.fm[s];
        lse = block_reduce_sum(part, red);
        val = lse / (double)a.S;
    } else {
        const double Mx = block_reduce_max(local_max, red);
        if (isinf(Mx)) {
            lse = Mx;
        } else {
            double part = 0.0;
            for (int s = tid; s < a.S; s += T) part += exp(c.fm[s] - Mx);
            lse = Mx + log(block_reduce_sum(part, red));
        }
        val = lse - log((double)a.S);
    }
    if (tid == 0) {
        const long o = (long)z * a.B + b;
        a.acq[o] = val;
        a.lse[o] = lse;
        a.info[o] = anybad ? -1 : s_info;
    }
    for (int idx = tid; idx < q * q; idx += T) a.Lq[((long)z * a.B + b) * q * q + idx] = c.Lq[idx];
}

}  // namespace

namespace {

// Reverse pass of cand_fwd_kernel for one (model, t-batch element).
//   in : F holds the samples f[s][j];  Lq = br_chol;  lse;  grad_acq
//   out: F <- d loss / d f ;  Sbar = adjoint of the Schur block (symmetric);
//        Vbar row n <- y_std * d loss / d mu ;  gradX <- contribution of k(X, X)
template <int QT>
__global__ void __launch_bounds__(CAND_THREADS, (QT <= 4 ? BO_CAND_MINB4 : (QT <= 8 ? 4 : 1))) cand_bwd_kernel(const BoCandArgs a) {
    extern __shared__ __align__(16) double sm[];
    const int q = a.q, b = blockIdx.x, z = blockIdx.y, tid = threadIdx.x, T = blockDim.x;
    const int c0 = b * q;
    const long o = (long)z * a.B + b;
    double* L = sm;            // q*q
    double* Lbar = L + q * q;  // q*q
    double* Pm = Lbar + q * q; // q*q
    double* Y = Pm + q * q;    // q*q
    double* mubar = Y + q * q; // q
    double* F = a.F + (long)z * a.strideF;
    const double g = a.df_external ? 0.0 : a.grad_acq[o];
    const double lse = a.lse[o];
    for (int idx = tid; idx < q * q; idx += T) L[idx] = a.Lq[o * q * q + idx];

    // 1+2. d loss / d f, written in place; mu_bar[j] = sum_s df[s][j] and Lbar[j][k] = sum_s df[s][j] Zq[s][k] (k <= j)
    //      accumulated in registers on the fly, then a fixed shuffle tree + fixed-order sum over warps (deterministic)
    constexpr int PT = QT * (QT + 1) / 2;
    double* red = mubar + q;  // (T/32) x (QT + PT)
    double am[QT], al[PT];
    const bool plain = (a.reduction == BO_REDUCE_PLAIN);
    const double inv_tau_relu = 1.0 / a.tau_relu;
#pragma unroll
    for (int j = 0; j < QT; j++) am[j] = 0.0;
#pragma unroll
    for (int p = 0; p < PT; p++) al[p] = 0.0;
    double fnext[QT], mnext = 0.0;  // sample rows (and saved maxima) read one iteration ahead, as in the forward kernel
    const double* msv = a.local_derivs ? a.msave + ((long)z * a.B + b) * a.Sp : nullptr;
#pragma unroll
    for (int j = 0; j < QT; j++) fnext[j] = (j < q && tid < a.S) ? F[(long)tid * a.ldn + c0 + j] : 0.0;
    if (msv && tid < a.S) mnext = msv[tid];
    for (int s = tid; s < a.S; s += T) {
        double f[QT], li[QT], gm[QT], zq[QT], fcur[QT];
        const double mcur = mnext;
#pragma unroll
        for (int j = 0; j < QT; j++) fcur[j] = fnext[j];
        if (s + T < a.S) {
#pragma unroll
            for (int j = 0; j < QT; j++) fnext[j] = (j < q) ? F[(long)(s + T) * a.ldn + c0 + j] : 0.0;
            if (msv) mnext = msv[s + T];
        }
        if (a.df_external) {  // F already holds d loss / d f: only the accumulations below
            int p2 = 0;
#pragma unroll
            for (int j = 0; j < QT; j++) {
                const double df = fcur[j];
                am[j] += df;
#pragma unroll
                for (int k = 0; k <= j; k++) al[p2++] += df * ((k < q) ? a.Zq[(long)s * q + k] : 0.0);
            }
            continue;
        }
        if (a.local_derivs) {  // the forward left d qmax_s / d f in the row and qmax_s in msave: scale and accumulate
            const double m = mcur;
            const double w = isinf(lse) ? ((m == lse) ? 1.0 : 0.0) : exp(m - lse);
            const double gw = a.obj_w * g * w;
            int p2 = 0;
#pragma unroll
            for (int j = 0; j < QT; j++) {
                double df = 0.0;
                if (j < q) {
                    df = gw * fcur[j];
                    F[(long)s * a.ldn + c0 + j] = df;
                }
                am[j] += df;
#pragma unroll
                for (int k = 0; k <= j; k++) al[p2++] += df * ((k < q) ? a.Zq[(long)s * q + k] : 0.0);
            }
            continue;
        }
        const double bf = a.best_f[(long)z * a.Sp + s];
#pragma unroll
        for (int j = 0; j < QT; j++) {
            f[j] = 0.0; li[j] = 0.0; zq[j] = 0.0;
            if (j < q) {
                f[j] = fcur[j];
                li[j] = plain ? fmax(a.obj_w * f[j] - bf, 0.0) : log_improvement(a.reduction, a.obj_w * f[j] - bf, a.tau_relu, inv_tau_relu);
                zq[j] = a.Zq[(long)s * q + j];
            }
        }
        double m, w;
        int ties = 0;
        if (plain) {  // mean over samples; amax spreads its gradient evenly over tied maxima; clamp_min(0) passes x >= 0
            m = li[0];
#pragma unroll
            for (int j = 1; j < QT; j++)
                if (j < q) m = fmax(m, li[j]);
#pragma unroll
            for (int j = 0; j < QT; j++)
                if (j < q && li[j] == m) ties++;
            w = 1.0 / (double)a.S;
        } else {
            m = qmax_log_grad<QT>(a.reduction, li, q, a.tau_max, gm);
            // logmeanexp over samples: softmax weight; +-inf maximum -> unit gradient on the infinite entries
            w = isinf(lse) ? ((m == lse) ? 1.0 : 0.0) : exp(m - lse);
        }
        int p = 0;
#pragma unroll
        for (int j = 0; j < QT; j++) {
            double df = 0.0;
            if (j < q) {
                if (plain) {
                    df = (li[j] == m && a.obj_w * f[j] - bf >= 0.0) ? a.obj_w * g * w / (double)ties : 0.0;
                } else {
                    const double dli = isinf(li[j]) ? 0.0 : dlog_improvement(a.reduction, a.obj_w * f[j] - bf, a.tau_relu, inv_tau_relu);
                    df = a.obj_w * g * w * gm[j] * dli;
                }
                F[(long)s * a.ldn + c0 + j] = df;
            }
            am[j] += df;
#pragma unroll
            for (int k = 0; k <= j; k++) al[p++] += df * zq[k];
        }
    }
    {
        const int warp = tid >> 5, lane = tid & 31;
        double both[QT + PT];
#pragma unroll
        for (int j = 0; j < QT; j++) both[j] = am[j];
#pragma unroll
        for (int p = 0; p < PT; p++) both[QT + p] = al[p];
        WarpMultiReduce<QT + PT>::run(both, lane, red + warp * (QT + PT));
    }
    __syncthreads();
    for (int idx = tid; idx < q + q * q; idx += T) {
        if (idx < q) {
            double s_ = 0.0;
            for (int w2 = 0; w2 < (T >> 5); w2++) s_ += red[w2 * (QT + PT) + idx];
            mubar[idx] = s_;
        } else {
            const int j = (idx - q) / q, k = (idx - q) % q;
            double s_ = 0.0;
            if (k <= j)
                for (int w2 = 0; w2 < (T >> 5); w2++) s_ += red[w2 * (QT + PT) + QT + j * (j + 1) / 2 + k];
            Lbar[j * q + k] = s_;
        }
    }
    __syncthreads();
    if (a.info[o] < 0) {  // forward failed: samples were not a function of the Schur block
        for (int idx = tid; idx < q * q; idx += T) Lbar[idx] = 0.0;
        __syncthreads();
    }

    // 3. Cholesky adjoint:  Sbar = sym( L^-T Phi(L^T Lbar) L^-1 ),  Phi = lower triangle with halved diagonal
    for (int idx = tid; idx < q * q; idx += T) {
        const int r = idx / q, cc = idx % q;
        double s_ = 0.0;
        if (cc <= r) {
            for (int k = r; k < q; k++) s_ += L[k * q + r] * Lbar[k * q + cc];
            if (cc == r) s_ *= 0.5;
        }
        Pm[idx] = s_;
    }
    __syncthreads();
    const bool ok = (a.info[o] >= 0);
    if (tid < q && ok) {  // solve L^T Y = Pm, one column per thread
        const int cc = tid;
        for (int i = q - 1; i >= 0; i--) {
            double s_ = Pm[i * q + cc];
            for (int k = i + 1; k < q; k++) s_ -= L[k * q + i] * Y[k * q + cc];
            Y[i * q + cc] = s_ / L[i * q + i];
        }
    }
    __syncthreads();
    if (tid < q && ok) {  // solve W L = Y, one row per thread (result back into Pm)
        const int r = tid;
        for (int j = q - 1; j >= 0; j--) {
            double s_ = Y[r * q + j];
            for (int k = j + 1; k < q; k++) s_ -= Pm[r * q + k] * L[k * q + j];
            Pm[r * q + j] = s_ / L[j * q + j];
        }
    }
    __syncthreads();
    double* Sbar = a.Sbar + o * q * q;
    for (int idx = tid; idx < q * q; idx += T) {
        const int r = idx / q, cc = idx % q;
        const double v = ok ? 0.5 * (Pm[r * q + cc] + Pm[cc * q + r]) : 0.0;
        Y[idx] = v;  // keep the symmetric adjoint in smem for step 4
        Sbar[idx] = v;
    }
    if (tid < q) a.Vbar_mu[(long)z * a.strideVbar + c0 + tid] = a.y_std * mubar[tid];
    if (a.post != nullptr) {
        // this t-batch's q rows of its 8 x 8 column-group matrix: u_coef * Sbar on the diagonal block, zero elsewhere
        double* Pg = a.post + (long)z * a.stridePost + (long)(c0 / 8) * 64;
        const int o8 = c0 % 8;
        for (int idx = tid; idx < q * 8; idx += T) {
            const int r = idx / 8, cc = idx % 8, k = cc - o8;
            const double v = (k >= 0 && k < q && ok) ? a.u_coef * 0.5 * (Pm[r * q + k] + Pm[k * q + r]) : 0.0;
            Pg[(o8 + r) * 8 + cc] = v;
        }
    }
    __syncthreads();

    // 4. gradient through k(X, X):  Schur = ys^2 (k(x_j,x_k) - ...) - ...   =>  kbar = ys^2 Sbar (symmetric)
    const double ys2 = a.y_std * a.y_std;
    const double os = a.h.outputscale ? a.h.outputscale[z] : 1.0;
    const double* il = a.h.inv_ls + (long)z * a.h.d;
    const int d = a.h.d;
    double* gX = a.gradX + o * q * d;
    // pair coefficients once (q^2 distance + transcendental evaluations), then the q*d gradient entries are plain sums
    for (int idx = tid; idx < q * q; idx += T) {
        const int j = idx / q, k = idx % q;
        Pm[idx] = (j == k) ? 0.0
                           : -2.0 * ys2 * Y[idx] * kern_dfactor(a.h.kernel_id, pair_r2(a.X, d, il, c0 + j, c0 + k), os);
    }
    __syncthreads();
    for (int idx = tid; idx < q * d; idx += T) {
        const int j = idx / d, t = idx % d;
        const double xj = a.X[(long)(c0 + j) * d + t];
        double acc = 0.0;
        for (int k = 0; k < q; k++) acc += Pm[j * q + k] * (xj - a.X[(long)(c0 + k) * d + t]);
        gX[idx] = acc * il[t] * il[t];
    }

    // 5. combined-root path: Ubar = u_coef * U Sbar for this t-batch's q columns, in place over U (one row per thread and
    //    iteration, 16-byte accesses); saves the separate pass and overlaps with the other CTAs' reduction phases
    if constexpr (QT <= 8) if (a.Uinplace != nullptr) {
        double* Ub = a.Uinplace + (long)z * a.strideU + c0;
        const bool vec = (q == QT) && ((QT & 1) == 0) && ((c0 & 1) == 0) && ((a.ldn & 1) == 0);
        __syncthreads();  // Y (the symmetric Sbar) is read below as a broadcast operand
#pragma unroll 2
        for (int r = tid; r < a.n_b; r += T) {
            double* row = Ub + (long)r * a.ldn;
            double v[QT], w[QT];
            if (vec) {
#pragma unroll
                for (int j = 0; j < QT; j += 2) {
                    const double2 t2 = *reinterpret_cast<const double2*>(row + j);
                    v[j] = t2.x;
                    v[j + 1] = t2.y;
                }
            } else {
#pragma unroll
                for (int j = 0; j < QT; j++) v[j] = (j < q) ? row[j] : 0.0;
            }
#pragma unroll
            for (int j = 0; j < QT; j++) {
                double s_ = 0.0;
#pragma unroll
                for (int k = 0; k < QT; k++)
                    if (k < q && j < q) s_ += v[k] * Y[k * q + j];
                w[j] = a.u_coef * s_;
            }
            if (vec) {
#pragma unroll
                for (int j = 0; j < QT; j += 2) *reinterpret_cast<double2*>(row + j) = make_double2(w[j], w[j + 1]);
            } else {
#pragma unroll
                for (int j = 0; j < QT; j++)
                    if (j < q) row[j] = w[j];
            }
        }
    }
}
size_t cand_bwd_smem_bytes(int q, int QT) {
    return sizeof(double) * (size_t)(4 * q * q + q + (CAND_THREADS / 32) * (QT + QT * (QT + 1) / 2));
}

__global__ void colblock_update_kernel(double* Y, long strideY, const double* Xm, long strideX, const double* Tm,
                                       int rows, int B, int q, long ldn, double coef, int accumulate) {
    extern __shared__ __align__(16) double st[];  // q*q block of T
    const int b = blockIdx.x, z = blockIdx.z;
    const double* T = Tm + ((long)z * B + b) * q * q;
    for (int i = threadIdx.x; i < q * q; i += blockDim.x) st[i] = T[i];
    __syncthreads();
    const int c0 = b * q;
    const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
    const int r_lo = blockIdx.y * rows_per, r_hi = min(rows, r_lo + rows_per);
    double* Yz = Y + (long)z * strideY;
    const double* Xz = Xm + (long)z * strideX;
    for (int idx = threadIdx.x; idx < (r_hi - r_lo) * q; idx += blockDim.x) {
        const int i = r_lo + idx / q, j = idx % q;
        double s = 0.0;
        for (int k = 0; k < q; k++) s += Xz[(long)i * ldn + c0 + k] * st[k * q + j];
        double* y = Yz + (long)i * ldn + c0 + j;
        *y = (accumulate ? *y : 0.0) + coef * s;
    }
}

// In-place variant: Y[z][i][b*q + :] <- coef * Y[z][i][b*q + :] * T[z][b].  G = next power of two >= q lanes share one
// row block (lane j of the group produces column j), so a warp touches 32/G rows per instruction with the same
// coalescing as the out-of-place kernel; all loads of a group precede its stores (__syncwarp), groups never span warps.
__global__ void colblock_inplace_kernel(double* Y, long strideY, const double* Tm, int rows, int B, int q, int G, long ldn,
                                        double coef) {
    extern __shared__ __align__(16) double st[];  // q*q block of T
    const int b = blockIdx.x, z = blockIdx.z;
    const double* T = Tm + ((long)z * B + b) * q * q;
    for (int i = threadIdx.x; i < q * q; i += blockDim.x) st[i] = T[i];
    __syncthreads();
    double* Yz = Y + (long)z * strideY + (long)b * q;
    const int j = threadIdx.x % G, sub = threadIdx.x / G, per = blockDim.x / G;
    const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
    const int r_lo = blockIdx.y * rows_per, r_hi = min(rows, r_lo + rows_per);
    for (int i0 = r_lo; i0 < r_hi; i0 += per) {  // uniform trip count per warp: __syncwarp below is safe
        const int i = i0 + sub;
        const bool live = i < r_hi && j < q;
        double s_ = 0.0;
        double* row = Yz + (long)i * ldn;
        if (live)
            for (int k = 0; k < q; k++) s_ += row[k] * st[k * q + j];
        __syncwarp();
        if (live) row[j] = coef * s_;
    }
}

// gradX[z][b][j][t] += scale * sum_i Kbar[i][c] * g(r_ic) * (x_i[t] - xc[t]) / l_t^2
__global__ void __launch_bounds__(CAND_THREADS) kgrad_kernel(BoKernelHyper h, const double* Xpts, int npts,
                                                             const double* Kbar, long strideK, long ldn,
                                                             const double* X, int B, int q, double scale,
                                                             double* gradX) {
    extern __shared__ __align__(16) double sm[];
    const int b = blockIdx.x, z = blockIdx.y, tid = threadIdx.x, T = blockDim.x;
    const int d = h.d, c0 = b * q;
    constexpr int R = 128;
    double* w = sm;              // R*q
    double* xc = w + R * q;      // q*d candidate points
    double* il2 = xc + q * d;    // d
    double* part = il2 + d;      // ngroups*q*d
    const double* il = h.inv_ls + (long)z * d;
    const double os = h.outputscale ? h.outputscale[z] : 1.0;
    const double* Kz = Kbar + (long)z * strideK;
    for (int i = tid; i < q * d; i += T) xc[i] = X[(long)c0 * d + i];
    for (int i = tid; i < d; i += T) il2[i] = il[i] * il[i];
    const int QD = q * d;
    const int ngroups = (T / QD) > 1 ? (T / QD) : 1;
    for (int i = tid; i < ngroups * QD; i += T) part[i] = 0.0;
    __syncthreads();
    for (int r0 = 0; r0 < npts; r0 += R) {
        const int pr = (npts - r0 < R) ? npts - r0 : R;
        for (int idx = tid; idx < pr * q; idx += T) {
            const int r = idx / q, j = idx % q;
            const double* xi = Xpts + (long)(r0 + r) * d;
            double r2 = 0.0;
            for (int t = 0; t < d; t++) {
                const double v = (xi[t] - xc[j * d + t]) * il[t];
                r2 += v * v;
            }
            w[idx] = Kz[(long)(r0 + r) * ldn + c0 + j] * kern_dfactor(h.kernel_id, r2, os);
        }
        __syncthreads();
        for (int slot = tid; slot < ngroups * QD; slot += T) {
            const int grp = slot / QD, jt = slot % QD;
            const int j = jt / d, t = jt % d;
            double acc = 0.0;
            for (int r = grp; r < pr; r += ngroups) acc += w[r * q + j] * (Xpts[(long)(r0 + r) * d + t] - xc[jt]);
            part[slot] += acc;
        }
        __syncthreads();
    }
    double* gX = gradX + ((long)z * B + b) * QD;
    for (int jt = tid; jt < QD; jt += T) {
        double s = 0.0;
        for (int grp = 0; grp < ngroups; grp++) s += part[grp * QD + jt];
        gX[jt] += scale * s * il2[jt % d];
    }
}
// gradX[z][c][t] += scale * il2[z][t] * (H[z][t][c] - x_c[t] * H[z][d][c]),   H = [train_X^T; 1] (Kbar .* g)
// H may come as `ksplit` partial sums over row ranges of the contraction (split-K, slice stride strideKs): they are
// added here in slice order, so the result does not depend on scheduling.
__global__ void kgrad_finalize_kernel(BoKernelHyper h, const double* H, long strideH, long ldn, const double* X, long N,
                                      double scale, double* gradX, int ksplit, long strideKs) {
    const int z = blockIdx.y, d = h.d;
    const long idx = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (idx >= N * d) return;
    const long c = idx / d;
    const int t = (int)(idx % d);
    const double il = h.inv_ls[(long)z * d + t];
    const double* Hz = H + (long)z * strideH;
    double ht = 0.0, hd = 0.0;
    for (int k = 0; k < ksplit; k++) {
        ht += Hz[k * strideKs + (long)t * ldn + c];
        hd += Hz[k * strideKs + (long)d * ldn + c];
    }
    gradX[(long)z * N * d + idx] += scale * il * il * (ht - X[idx] * hd);
}
// The same for per-row-tile partials of a fused contraction (BoGemm::hfuse: dozens of slices instead of <= 8): one thread per
// (input dimension, column) with the column fastest, so that every slice row is read coalesced.
__global__ void kgrad_finalize_cols_kernel(BoKernelHyper h, const double* H, long strideH, long ldn, const double* X, long N,
                                           double scale, double* gradX, int ksplit, long strideKs) {
    const int z = blockIdx.z, d = h.d, t = blockIdx.y;
    const long c = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (c >= N) return;
    const double il = h.inv_ls[(long)z * d + t];
    const double* Hz = H + (long)z * strideH;
    double ht = 0.0, hd = 0.0;
#pragma unroll 4
    for (int k = 0; k < ksplit; k++) {
        ht += Hz[k * strideKs + (long)t * ldn + c];
        hd += Hz[k * strideKs + (long)d * ldn + c];
    }
    gradX[(long)z * N * d + c * d + t] += scale * il * il * (ht - X[c * d + t] * hd);
}
size_t kgrad_smem_bytes(int q, int d, int T) {
    const int QD = q * d;
    const int ngroups = (T / QD) > 1 ? (T / QD) : 1;
    return sizeof(double) * ((size_t)128 * q + QD + d + (size_t)ngroups * QD);
}

}  // namespace

namespace {

// joint posterior blocks of one t-batch element from V = [inv(L) K(X_tr, X); alpha . K]
__global__ void __launch_bounds__(CAND_THREADS) cand_posterior_kernel(const BoCandArgs a, double* mean, double* cov) {
    extern __shared__ __align__(16) double sm[];
    const int q = a.q, b = blockIdx.x, z = blockIdx.y, tid = threadIdx.x, T = blockDim.x;
    const CandSmem c = carve(sm, q, T, 0);
    const int c0 = b * q;
    const double* V = a.V + (long)z * a.strideV;
    gram_any(V, a.ldn, a.n, c0, q, c.panel, c.part, c.GV);
    const double ys2 = a.y_std * a.y_std;
    const double os = a.h.outputscale ? a.h.outputscale[z] : 1.0;
    const double* il = a.h.inv_ls + (long)z * a.h.d;
    const long o = (long)z * a.B + b;
    for (int idx = tid; idx < q * q; idx += T) {
        const int j = idx / q, k = idx % q;
        const double kv = kern_value(a.h.kernel_id, pair_r2(a.X, a.h.d, il, c0 + j, c0 + k), os);
        cov[o * q * q + idx] = ys2 * (kv - c.GV[idx]);
    }
    if (tid < q) mean[o * q + tid] = a.y_mean + a.y_std * (a.mean_const[z] + V[(long)a.n * a.ldn + c0 + tid]);
}

// Reverse pass of cand_posterior_kernel for one (model, t-batch element):
//   T = cov_bar + cov_bar^T (adjoint of the symmetric block), row n of Vbar <- y_std * mean_bar,
//   gradX <- contribution of k(X, X):  cov = ys^2 (k(x_j, x_k) - ...)
__global__ void __launch_bounds__(CAND_THREADS) cand_posterior_bwd_kernel(const BoCandArgs a, const double* mean_bar,
                                                                          const double* cov_bar, double* Tout) {
    extern __shared__ __align__(16) double sm[];
    const int q = a.q, b = blockIdx.x, z = blockIdx.y, tid = threadIdx.x, T = blockDim.x;
    const int c0 = b * q, d = a.h.d;
    const long o = (long)z * a.B + b;
    double* Ts = sm;  // q*q
    for (int idx = tid; idx < q * q; idx += T) {
        const int j = idx / q, k = idx % q;
        const double v = cov_bar[o * q * q + j * q + k] + cov_bar[o * q * q + k * q + j];
        Ts[idx] = v;
        Tout[o * q * q + idx] = v;
    }
    if (tid < q) a.Vbar_mu[(long)z * a.strideVbar + c0 + tid] = a.y_std * mean_bar[o * q + tid];
    __syncthreads();
    const double ys2 = a.y_std * a.y_std;
    const double os = a.h.outputscale ? a.h.outputscale[z] : 1.0;
    const double* il = a.h.inv_ls + (long)z * d;
    double* gX = a.gradX + o * q * d;
    for (int idx = tid; idx < q * d; idx += T) {
        const int j = idx / d, t = idx % d;
        double acc = 0.0;
        for (int k = 0; k < q; k++) {
            if (k == j) continue;
            const double gf = kern_dfactor(a.h.kernel_id, pair_r2(a.X, d, il, c0 + j, c0 + k), os);
            acc += ys2 * Ts[j * q + k] * (-gf) * (a.X[(long)(c0 + j) * d + t] - a.X[(long)(c0 + k) * d + t]) * il[t] *
                   il[t];
        }
        gX[idx] = acc;
    }
}

// Column statistics of a row-major panel.  Block = 32 columns x COL_LANES row lanes: a warp reads 32 consecutive doubles
// of one row (256 B, coalesced), lane group `ty` owns rows ty, ty + COL_LANES, ... and the lanes are combined in fixed
// order through shared memory, so the result depends on neither the grid nor the scheduling.  (The first version had one
// thread per column and 128-column blocks: 128 CTAs for a 16 384-candidate chunk, 0.19 of HBM peak, and TWO CTAs for the
// 256 sample columns of bo_nei_setup -- profiles/r02_kernel_rooflines.md.)
constexpr int COL_LANES = 8;
__global__ void __launch_bounds__(32 * COL_LANES) col_mean_var_kernel(const double* V, long strideV, long ldn, int n,
                                                                      long Nc, double y_mean, double y_std,
                                                                      const double* mean_const, const double* outputscale,
                                                                      double* mean, double* var) {
    __shared__ double part[COL_LANES][33];
    const int z = blockIdx.y, tx = threadIdx.x, ty = threadIdx.y;
    const long c = blockIdx.x * 32L + tx;
    const double* Vz = V + (long)z * strideV;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;  // 4 independent chains per lane, fixed order
    if (c < Nc) {
        int i = ty;
        for (; i + 3 * COL_LANES < n; i += 4 * COL_LANES) {
            const double v0 = Vz[(long)i * ldn + c], v1 = Vz[(long)(i + COL_LANES) * ldn + c];
            const double v2 = Vz[(long)(i + 2 * COL_LANES) * ldn + c], v3 = Vz[(long)(i + 3 * COL_LANES) * ldn + c];
            s0 += v0 * v0; s1 += v1 * v1; s2 += v2 * v2; s3 += v3 * v3;
        }
        for (; i < n; i += COL_LANES) { const double v = Vz[(long)i * ldn + c]; s0 += v * v; }
    }
    part[ty][tx] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (ty != 0 || c >= Nc) return;
    double ss = 0.0;
    #pragma unroll
    for (int l = 0; l < COL_LANES; l++) ss += part[l][tx];
    const double os = outputscale ? outputscale[z] : 1.0;
    mean[(long)z * Nc + c] = y_mean + y_std * (mean_const[z] + Vz[(long)n * ldn + c]);
    var[(long)z * Nc + c] = y_std * y_std * (os - ss);
}

constexpr int MAX_LANES = 32;
__global__ void __launch_bounds__(32 * MAX_LANES) col_max_kernel(const double* Fb, long strideF, long ldf,
                                                                 const double* mean_b, long strideMb, int rows, int cols,
                                                                 double* out, long strideOut, double w) {
    __shared__ double part[MAX_LANES][33];
    const int z = blockIdx.y, tx = threadIdx.x, ty = threadIdx.y;
    const int s = blockIdx.x * 32 + tx;
    const double* Fz = Fb + (long)z * strideF;
    const double* mb = mean_b + (long)z * strideMb;
    double m = -INFINITY;
    if (s < cols)
        for (int i = ty; i < rows; i += MAX_LANES) m = fmax(m, w * (mb[i] + Fz[(long)i * ldf + s]));  // objective value w * f
    part[ty][tx] = m;
    __syncthreads();
    if (ty != 0 || s >= cols) return;
    #pragma unroll
    for (int l = 1; l < MAX_LANES; l++) m = fmax(m, part[l][tx]);  // max is exact in any order
    out[(long)z * strideOut + s] = m;
}

// per-model first-occurrence argmax of mean + sqrt(var) * z  (z may be null -> argmax of mean)
__global__ void __launch_bounds__(1024) ts_argmax_kernel(long Nc, const double* mean, const double* var,
                                                         const double* zn, double* out_val, long long* out_idx) {
    __shared__ double sv[32];
    __shared__ long long si[32];
    const int m = blockIdx.x;
    double best = -INFINITY;
    long long bi = 0x7fffffffffffffffLL;
    for (long c = threadIdx.x; c < Nc; c += blockDim.x) {
        double v = mean[(long)m * Nc + c];
        if (zn) v += sqrt(fmax(var[(long)m * Nc + c], 0.0)) * zn[(long)m * Nc + c];
        if (v > best || (v == best && c < bi)) { best = v; bi = c; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sv[w] = best; si[w] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < (int)(blockDim.x >> 5); i++)
            if (sv[i] > best || (sv[i] == best && si[i] < bi)) { best = sv[i]; bi = si[i]; }
        out_val[m] = best;
        out_idx[m] = bi;
    }
}

}  // namespace

#define BO_LAUNCH_CHECK()        \
    do {                         \
        g_bo_launch_count++;     \
        BO_CUDA_TRY(cudaGetLastError()); \
    } while (0)

int bo_aug_rows(const BoKernelHyper& h, const double* X, long npts, double* out, long Kp, long strideOut,
                cudaStream_t stream) {
    if (npts <= 0) return BO_OK;
    dim3 grid((unsigned)((npts + 127) / 128), h.M);
    aug_rows_kernel<<<grid, 128, 0, stream>>>(h, X, npts, out, Kp, strideOut);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_aug_cols(const BoKernelHyper& h, const double* X, long npts, double* out, long Kp, long ldo, long strideOut,
                cudaStream_t stream) {
    if (npts <= 0) return BO_OK;
    dim3 grid((unsigned)((ldo + 127) / 128), h.M);
    aug_cols_kernel<<<grid, 128, 0, stream>>>(h, X, npts, out, Kp, ldo, strideOut);
    BO_LAUNCH_CHECK();
    return BO_OK;
}

static int set_smem_limit(const void* fn, size_t bytes) {
    if (bytes > 48 * 1024) BO_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return BO_OK;
}

namespace {

// One CTA per t-batch element: per-sample log improvement (+ log feasibility weight), q-reduction, sample reduction.
template <int QT>
__global__ void __launch_bounds__(CAND_THREADS) mc_reduce_fwd_kernel(const BoReduceArgs a) {
    extern __shared__ __align__(16) double sm[];  // fm[S]
    __shared__ double red[CAND_THREADS / 32];
    const int q = a.q, b = blockIdx.x, tid = threadIdx.x, T = blockDim.x, c0 = b * q;
    const bool plain = (a.reduction == BO_REDUCE_PLAIN);
    const double inv_tau_relu = 1.0 / a.tau_relu;
    double local_max = -INFINITY;
    for (int s = tid; s < a.S; s += T) {
        double li[QT];
        const double bf = a.best_f[s];
#pragma unroll
        for (int j = 0; j < QT; j++) {
            li[j] = 0.0;
            if (j < q) {
                const double v = a.F[(long)s * a.ldn + c0 + j];
                const double lw = a.lw ? a.lw[(long)s * a.ldn + c0 + j] : 0.0;
                li[j] = plain ? fmax(a.obj_w * v - bf, 0.0) * (a.lw ? exp(lw) : 1.0)
                              : log_improvement(a.reduction, a.obj_w * v - bf, a.tau_relu, inv_tau_relu) + lw;
            }
        }
        double m;
        if (plain) {
            m = li[0];
#pragma unroll
            for (int j = 1; j < QT; j++)
                if (j < q) m = fmax(m, li[j]);
        } else {
            m = qmax_log<QT>(a.reduction, li, q, a.tau_max);
        }
        sm[s] = m;
        local_max = fmax(local_max, m);
    }
    __syncthreads();
    double lse, val;
    if (plain) {
        double part = 0.0;
        for (int s = tid; s < a.S; s += T) part += sm[s];
        lse = block_reduce_sum(part, red);
        val = lse / (double)a.S;
    } else {
        const double Mx = block_reduce_max(local_max, red);
        if (isinf(Mx)) {
            lse = Mx;
        } else {
            double part = 0.0;
            for (int s = tid; s < a.S; s += T) part += exp(sm[s] - Mx);
            lse = Mx + log(block_reduce_sum(part, red));
        }
        val = lse - log((double)a.S);
    }
    if (tid == 0) { a.acq[b] = val; a.lse[b] = lse; }
}

// Reverse of mc_reduce_fwd_kernel: dF[s][c] = d loss / d f, lw_bar[s][c] = d loss / d lw
template <int QT>
__global__ void __launch_bounds__(CAND_THREADS) mc_reduce_bwd_kernel(const BoReduceArgs a) {
    const int q = a.q, b = blockIdx.x, tid = threadIdx.x, T = blockDim.x, c0 = b * q;
    const bool plain = (a.reduction == BO_REDUCE_PLAIN);
    const double inv_tau_relu = 1.0 / a.tau_relu;
    const double g = a.grad_acq[b], lse = a.lse[b];
    for (int s = tid; s < a.S; s += T) {
        double f[QT], li[QT], lwv[QT], gm[QT];
        const double bf = a.best_f[s];
#pragma unroll
        for (int j = 0; j < QT; j++) {
            f[j] = 0.0; li[j] = 0.0; lwv[j] = 0.0; gm[j] = 0.0;
            if (j < q) {
                f[j] = a.F[(long)s * a.ldn + c0 + j];
                lwv[j] = a.lw ? a.lw[(long)s * a.ldn + c0 + j] : 0.0;
                li[j] = plain ? fmax(a.obj_w * f[j] - bf, 0.0) * (a.lw ? exp(lwv[j]) : 1.0)
                              : log_improvement(a.reduction, a.obj_w * f[j] - bf, a.tau_relu, inv_tau_relu) + lwv[j];
            }
        }
        double m, w;
        int ties = 0;
        if (plain) {
            m = li[0];
#pragma unroll
            for (int j = 1; j < QT; j++)
                if (j < q) m = fmax(m, li[j]);
#pragma unroll
            for (int j = 0; j < QT; j++)
                if (j < q && li[j] == m) ties++;
            w = 1.0 / (double)a.S;
        } else {
            m = qmax_log_grad<QT>(a.reduction, li, q, a.tau_max, gm);
            w = isinf(lse) ? ((m == lse) ? 1.0 : 0.0) : exp(m - lse);
        }
#pragma unroll
        for (int j = 0; j < QT; j++) {
            if (j >= q) continue;
            double df, dl;
            if (plain) {  // value = relu(obj - bf) * exp(lw); amax spreads its gradient evenly over ties
                const double hit = (li[j] == m) ? g * w / (double)ties : 0.0;
                const double ind = a.lw ? exp(lwv[j]) : 1.0;
                df = (a.obj_w * f[j] - bf >= 0.0) ? hit * a.obj_w * ind : 0.0;
                dl = hit * li[j];
            } else {
                const double dli = isinf(li[j]) ? 0.0 : dlog_improvement(a.reduction, a.obj_w * f[j] - bf, a.tau_relu, inv_tau_relu);
                dl = isinf(li[j]) ? 0.0 : g * w * gm[j];
                df = a.obj_w * dl * dli;
            }
            a.dF[(long)s * a.ldn + c0 + j] = df;
            if (a.lw_bar) a.lw_bar[(long)s * a.ldn + c0 + j] = dl;
        }
    }
}

}  // namespace

int bo_mc_reduce_fwd_launch(const BoReduceArgs& a, cudaStream_t stream) {
    if (a.q < 1 || a.q > BO_MAX_Q) return BO_ERR_UNSUPPORTED;
    if (a.B <= 0) return BO_OK;
    const size_t smem = sizeof(double) * (size_t)a.S;
    if (smem > 200 * 1024) return BO_ERR_UNSUPPORTED;
    if (a.q <= 4) {
        BO_TRY(set_smem_limit((const void*)mc_reduce_fwd_kernel<4>, smem));
        mc_reduce_fwd_kernel<4><<<a.B, CAND_THREADS, smem, stream>>>(a);
    } else if (a.q <= 8) {
        BO_TRY(set_smem_limit((const void*)mc_reduce_fwd_kernel<8>, smem));
        mc_reduce_fwd_kernel<8><<<a.B, CAND_THREADS, smem, stream>>>(a);
    } else {
        BO_TRY(set_smem_limit((const void*)mc_reduce_fwd_kernel<BO_MAX_Q>, smem));
        mc_reduce_fwd_kernel<BO_MAX_Q><<<a.B, CAND_THREADS, smem, stream>>>(a);
    }
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_mc_reduce_bwd_launch(const BoReduceArgs& a, cudaStream_t stream) {
    if (a.q < 1 || a.q > BO_MAX_Q) return BO_ERR_UNSUPPORTED;
    if (a.B <= 0) return BO_OK;
    if (a.q <= 4) mc_reduce_bwd_kernel<4><<<a.B, CAND_THREADS, 0, stream>>>(a);
    else if (a.q <= 8) mc_reduce_bwd_kernel<8><<<a.B, CAND_THREADS, 0, stream>>>(a);
    else mc_reduce_bwd_kernel<BO_MAX_Q><<<a.B, CAND_THREADS, 0, stream>>>(a);
    BO_LAUNCH_CHECK();
    return BO_OK;
}

// Threads per (model, t-batch) CTA.  With thousands of small problems (SAAS: 256 models x 64 t-batches) narrower CTAs win a
// little: the serial sections (q x q Cholesky on one warp, block reductions) idle fewer warps and more problems are in
// flight per SM.  Measured at C4 (16 384 problems, step time): 128 threads 1.62 ms, 64 threads 1.45 ms, 32 threads
// 1.35 ms with the sample rows read one iteration ahead (before that: 1.60 / 1.55 / 1.56): one warp per problem has no
// block barrier to wait at and 8 samples per thread to hide each load behind.
// BO_CAND_THREADS overrides (32 / 64 / 128) for A/B measurements.
static int cand_threads(const BoCandArgs& a) {
    static const int forced = [] { const char* e = getenv("BO_CAND_THREADS"); return e ? atoi(e) : 0; }();
    if (forced == 32 || forced == 64 || forced == 128) return forced;
    const long problems = (long)a.B * a.h.M;
    if (problems >= 8192 && a.q <= 4) return 32;
    return (problems >= 4096 && a.q <= 8) ? 64 : CAND_THREADS;
}

int bo_cand_fwd(const BoCandArgs& a, cudaStream_t stream) {
    if (a.q < 1 || a.q > BO_MAX_Q || a.h.d > BO_MAX_D) return BO_ERR_UNSUPPORTED;
    if (a.B <= 0) return BO_OK;
    const int CT = cand_threads(a);
    const size_t smem = cand_smem_bytes(a.q, CAND_THREADS, a.S);
    if (smem > 200 * 1024) return BO_ERR_UNSUPPORTED;
    dim3 grid(a.B, a.h.M);
    if (a.local_derivs && a.q > 8) return BO_ERR_INVALID;  // cand_args clears it for q > 8
    if (a.q <= 4 && a.local_derivs) {
        BO_TRY(set_smem_limit((const void*)cand_fwd_kernel<4, true>, smem));
        cand_fwd_kernel<4, true><<<grid, CT, smem, stream>>>(a);
    } else if (a.q <= 4) {
        BO_TRY(set_smem_limit((const void*)cand_fwd_kernel<4, false>, smem));
        cand_fwd_kernel<4, false><<<grid, CT, smem, stream>>>(a);
    } else if (a.q <= 8 && a.local_derivs) {
        BO_TRY(set_smem_limit((const void*)cand_fwd_kernel<8, true>, smem));
        cand_fwd_kernel<8, true><<<grid, CT, smem, stream>>>(a);
    } else if (a.q <= 8) {
        BO_TRY(set_smem_limit((const void*)cand_fwd_kernel<8, false>, smem));
        cand_fwd_kernel<8, false><<<grid, CT, smem, stream>>>(a);
    } else if (a.q <= 16) {
        BO_TRY(set_smem_limit((const void*)cand_fwd_kernel<16, false>, smem));
        cand_fwd_kernel<16, false><<<grid, CT, smem, stream>>>(a);
    } else {
        BO_TRY(set_smem_limit((const void*)cand_fwd_kernel<BO_MAX_Q, false>, smem));
        cand_fwd_kernel<BO_MAX_Q, false><<<grid, CT, smem, stream>>>(a);
    }
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_cand_bwd(const BoCandArgs& a, cudaStream_t stream) {
    if (a.q < 1 || a.q > BO_MAX_Q || a.h.d > BO_MAX_D) return BO_ERR_UNSUPPORTED;
    if (a.B <= 0) return BO_OK;
    dim3 grid(a.B, a.h.M);
    const int CT = cand_threads(a);
    if (a.q <= 4) cand_bwd_kernel<4><<<grid, CT, cand_bwd_smem_bytes(a.q, 4), stream>>>(a);
    else if (a.q <= 8) cand_bwd_kernel<8><<<grid, CT, cand_bwd_smem_bytes(a.q, 8), stream>>>(a);
    else if (a.q <= 12) cand_bwd_kernel<12><<<grid, CAND_THREADS, cand_bwd_smem_bytes(a.q, 12), stream>>>(a);
    else {
        const size_t smem = cand_bwd_smem_bytes(a.q, BO_MAX_Q);
        BO_TRY(set_smem_limit((const void*)cand_bwd_kernel<BO_MAX_Q>, smem));
        cand_bwd_kernel<BO_MAX_Q><<<grid, CAND_THREADS, smem, stream>>>(a);
    }
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_colblock_update(double* Y, long strideY, const double* Xm, long strideX, const double* T, int rows, int B,
                       int q, int M, long ldn, double coef, int accumulate, cudaStream_t stream) {
    if (rows <= 0 || B <= 0) return BO_OK;
    int ysplit = (rows + 255) / 256;  // ~256 rows per CTA
    if (ysplit > 65535) ysplit = 65535;
    dim3 grid(B, ysplit, M);
    colblock_update_kernel<<<grid, 256, sizeof(double) * q * q, stream>>>(Y, strideY, Xm, strideX, T, rows, B, q, ldn,
                                                                         coef, accumulate);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_colblock_inplace(double* Y, long strideY, const double* T, int rows, int B, int q, int M, long ldn, double coef,
                        cudaStream_t stream) {
    if (rows <= 0 || B <= 0) return BO_OK;
    int ysplit = (rows + 255) / 256;
    if (ysplit > 65535) ysplit = 65535;
    dim3 grid(B, ysplit, M);
    int G = 1;
    while (G < q) G <<= 1;
    colblock_inplace_kernel<<<grid, 256, sizeof(double) * q * q, stream>>>(Y, strideY, T, rows, B, q, G, ldn, coef);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_kgrad(const BoKernelHyper& h, const double* Xpts, int npts, const double* Kbar, long strideK, long ldn,
             const double* X, int B, int q, double scale, double* gradX, cudaStream_t stream) {
    if (B <= 0 || npts <= 0) return BO_OK;
    const size_t smem = kgrad_smem_bytes(q, h.d, CAND_THREADS);
    if (smem > 200 * 1024) return BO_ERR_UNSUPPORTED;
    BO_TRY(set_smem_limit((const void*)kgrad_kernel, smem));
    dim3 grid(B, h.M);
    kgrad_kernel<<<grid, CAND_THREADS, smem, stream>>>(h, Xpts, npts, Kbar, strideK, ldn, X, B, q, scale, gradX);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_kgrad_finalize(const BoKernelHyper& h, const double* H, long strideH, long ldn, const double* X, long N,
                      double scale, double* gradX, int ksplit, long strideKs, cudaStream_t stream) {
    if (N <= 0) return BO_OK;
    if (ksplit > 8) {  // many slices: coalesced form
        dim3 gridc((unsigned)((N + 127) / 128), (unsigned)h.d, (unsigned)h.M);
        kgrad_finalize_cols_kernel<<<gridc, 128, 0, stream>>>(h, H, strideH, ldn, X, N, scale, gradX, ksplit, strideKs);
        BO_LAUNCH_CHECK();
        return BO_OK;
    }
    dim3 grid((unsigned)((N * h.d + 255) / 256), h.M);
    kgrad_finalize_kernel<<<grid, 256, 0, stream>>>(h, H, strideH, ldn, X, N, scale, gradX, ksplit, strideKs);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_cand_posterior(const BoCandArgs& a, double* mean, double* cov, cudaStream_t stream) {
    if (a.q < 1 || a.q > BO_MAX_Q) return BO_ERR_UNSUPPORTED;
    if (a.B <= 0) return BO_OK;
    const size_t smem = cand_smem_bytes(a.q, CAND_THREADS, 0);
    BO_TRY(set_smem_limit((const void*)cand_posterior_kernel, smem));
    dim3 grid(a.B, a.h.M);
    cand_posterior_kernel<<<grid, CAND_THREADS, smem, stream>>>(a, mean, cov);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_cand_posterior_bwd(const BoCandArgs& a, const double* mean_bar, const double* cov_bar, double* Tout,
                          cudaStream_t stream) {
    if (a.q < 1 || a.q > BO_MAX_Q || a.h.d > BO_MAX_D) return BO_ERR_UNSUPPORTED;
    if (a.B <= 0) return BO_OK;
    dim3 grid(a.B, a.h.M);
    cand_posterior_bwd_kernel<<<grid, CAND_THREADS, sizeof(double) * a.q * a.q, stream>>>(a, mean_bar, cov_bar, Tout);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_col_mean_var(const double* V, long strideV, long ldn, int n, long Nc, int M, double y_mean, double y_std,
                    const double* mean_const, const double* outputscale, double* mean, double* var,
                    cudaStream_t stream) {
    if (Nc <= 0) return BO_OK;
    dim3 grid((unsigned)((Nc + 31) / 32), M);
    col_mean_var_kernel<<<grid, dim3(32, COL_LANES), 0, stream>>>(V, strideV, ldn, n, Nc, y_mean, y_std, mean_const, outputscale, mean,
                                                  var);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_col_max(const double* Fb, long strideF, long ldf, const double* mean_b, long strideMb, int rows, int cols,
               int M, double* out, long strideOut, double obj_w, cudaStream_t stream) {
    dim3 grid((cols + 31) / 32, M);
    col_max_kernel<<<grid, dim3(32, MAX_LANES), 0, stream>>>(Fb, strideF, ldf, mean_b, strideMb, rows, cols, out, strideOut, obj_w);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
int bo_ts_argmax_launch(int M, long Nc, const double* mean, const double* var, const double* z, double* out_val,
                        long long* out_idx, cudaStream_t stream) {
    if (M <= 0 || Nc <= 0) return BO_ERR_INVALID;
    ts_argmax_kernel<<<M, 1024, 0, stream>>>(Nc, mean, var, z, out_val, out_idx);
    BO_LAUNCH_CHECK();
    return BO_OK;
}
