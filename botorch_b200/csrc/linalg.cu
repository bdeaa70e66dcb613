// Blocked right-looking Cholesky and triangular inverse for the one-time GP setup
// (K + s2 I = L L^T, inv(L); posterior covariance root L_b at X_baseline).
// Replaces torch.linalg.cholesky / solve_triangular(L, I) behind gpytorch's prediction caches
// (reference call site models/gpytorch.py:446 -> DefaultPredictionStrategy.mean_cache/covar_cache)
// and behind cached_cholesky.py:115-120 (_compute_root_decomposition).
// Diagonal NB x NB blocks are factored and inverted inside one CTA in shared memory; panels and
// trailing updates are DMMA GEMMs (gemm.cu).
#include "linalg.cuh"


namespace {

constexpr int NB = BO_NB;

// One CTA per batch member: Cholesky of the kb x kb diagonal block + its inverse (and inverse^T for the panel solve).
// 256 threads as a 16 x 16 grid; thread (ty, tx) keeps the 4 x 4 elements (ty + 16 a, tx + 16 b) of the NB x NB block in
// registers.  Step j works from column j in a double-buffered shared vector: every thread forms l = column * rsqrt(pivot),
// the owners of column j+1 update and PUBLISH that column first (look-ahead), then everyone applies the rest of the
// rank-1 update to its registers -- one barrier per step, and the publish latency hides behind the bulk update.
// The triangular inverse is blocked 16 -> 32 -> 64 with register-tiled merges.
// History (clock64 stamps, tools/diag_timing.cu): shared-memory version with three barriers per step 94 us; registers +
// one barrier 40 us; phase stamps then showed 433 cycles per column step and 13k cycles in un-pipelined merge loops.
static_assert(NB == 64, "potrf_diag_kernel is laid out for 64 x 64 diagonal blocks");
#ifdef BO_DIAG_TIMING  // tools/diag_timing.cu: clock64 stamps at the phase boundaries of one CTA
__device__ long long g_diag_stamps[16];
#define DIAG_STAMP(i) do { if (threadIdx.x == 0 && blockIdx.x == 0) g_diag_stamps[i] = clock64(); } while (0)
#else
#define DIAG_STAMP(i) do { } while (0)
#endif

// Owner work for column j1 (tile column BJ1 = j1 / 16), executed by the ONE warp that holds it (thread grid: tx = tid / 16,
// ty = tid % 16, so the 16 threads owning a column are a half warp): pivot by shuffle, l = column * rsqrt(pivot), final L
// values into the registers, scaled column (zeros at and above the pivot) into the shared vector of the next step.
// Keeping rsqrt and the scaling in one warp matters: the SM issues FP64 at 64 lanes per clock, and eight warps redundantly
// running rsqrt + 8 multiplies per step cost more pipe time than the rank-1 update itself.
template <int BJ1>
__device__ __forceinline__ void chol_publish(double (&a)[4][4], int j1, int kb, int ty, int tx, double* nxt, double* rdiag,
                                             volatile int* failflag) {
    const double piv = __shfl_sync(0xffffffffu, a[BJ1][BJ1], 16 * (j1 & 1) + (j1 & 15));
    const double rd = rsqrt(piv);  // (a float-seeded rsqrt + one cubic correction measured SLOWER: 620 vs 525 cycles per step)
    const double d = piv * rd;
    if (tx == (j1 & 15)) {
#pragma unroll
        for (int ai = 0; ai < 4; ai++) {
            const int i = ty + 16 * ai;
            double v = 0.0;
            if (i > j1) { v = a[ai][BJ1] * rd; a[ai][BJ1] = v; }
            else if (i == j1) a[ai][BJ1] = d;
            nxt[i] = v;
        }
        if (ty == 0) {
            rdiag[j1] = rd;
            if (!(piv > 0.0) && j1 < kb && *failflag == 0) *failflag = j1 + 1;
        }
    }
}

// One column step; BJ = register-tile column of column j, BJ1 = that of column j+1 (compile time: static register indexing).
// The warp owning column j+1 updates that tile column first and publishes it (look-ahead) before the bulk of its update.
template <int BJ, int BJ1>
__device__ __forceinline__ void chol_step(double (&a)[4][4], int j, int kb, int ty, int tx, double (*colbuf)[NB],
                                          double* rdiag, volatile int* failflag) {
    const double* buf = colbuf[j & 1];
    double li[4], lk[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        li[q] = q >= BJ ? buf[ty + 16 * q] : 0.0;  // register tiles left of / above tile column BJ are finished
        lk[q] = q >= BJ ? buf[tx + 16 * q] : 0.0;
    }
    bool own = false;
    if constexpr (BJ1 < 4) {
        own = (threadIdx.x >> 5) == (((j + 1) & 15) >> 1);  // warp-uniform
        if (own) {
#pragma unroll
            for (int ai = 0; ai < 4; ai++)
                if (ai >= BJ) a[ai][BJ1] -= li[ai] * lk[BJ1];
            chol_publish<BJ1>(a, j + 1, kb, ty, tx, colbuf[(j + 1) & 1], rdiag, failflag);
        }
    }
#pragma unroll
    for (int ai = 0; ai < 4; ai++)
#pragma unroll
        for (int bi = 0; bi < 4; bi++)
            if (ai >= BJ && bi >= BJ && !(own && bi == BJ1)) a[ai][bi] -= li[ai] * lk[bi];
    __syncthreads();
}

template <int BJ>
__device__ __forceinline__ void chol_block(double (&a)[4][4], int kb, int ty, int tx, double (*colbuf)[NB], double* rdiag,
                                           volatile int* failflag) {
    // a failed pivot only poisons later columns with NaN: looking at the flag once per 16 columns keeps the shared-memory
    // round trip + branch off the per-column dependent chain (uniform: written before the barrier that ended a step)
    if (*failflag) return;
    for (int jj = 0; jj < 15; jj++) chol_step<BJ, BJ>(a, BJ * 16 + jj, kb, ty, tx, colbuf, rdiag, failflag);
    chol_step<BJ, BJ + 1>(a, BJ * 16 + 15, kb, ty, tx, colbuf, rdiag, failflag);
}

__global__ void __launch_bounds__(256) potrf_diag_kernel(double* A, long lda, long strideA, int kb, int k0,
                                                          double* inv, long ldi, long strideI, double* invT,
                                                          int* info) {
    extern __shared__ __align__(16) double potrf_sm[];
    double (*S)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm);
    double (*X)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm + NB * (NB + 1));
    __shared__ double colbuf[2][NB];
    __shared__ double rdiag[NB];
    __shared__ double Tbuf[32 * 33];
    __shared__ int failflag;
    const int z = blockIdx.x, tid = threadIdx.x;
    const int tx = tid >> 4, ty = tid & 15;
    double* Az = A + (long)z * strideA;
    double a[4][4];
    DIAG_STAMP(0);
    if (tid == 0) failflag = 0;
    {   // coalesced rows into shared (all 16 loads of a thread in flight together), then each thread's 4 x 4 patch
        double tmp[16];
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const int idx = tid + 256 * u, i = idx >> 6, j = idx & 63;
            tmp[u] = (i < kb && j <= i) ? __ldg(Az + (long)i * lda + j) : ((i == j) ? 1.0 : 0.0);
        }
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const int idx = tid + 256 * u;
            S[idx >> 6][idx & 63] = tmp[u];
        }
    }
    __syncthreads();
#pragma unroll
    for (int ai = 0; ai < 4; ai++)
#pragma unroll
        for (int bi = 0; bi < 4; bi++) a[ai][bi] = S[ty + 16 * ai][tx + 16 * bi];
    if (tid < 32) chol_publish<0>(a, 0, kb, ty, tx, colbuf[0], rdiag, &failflag);
    __syncthreads();
    DIAG_STAMP(1);
    chol_block<0>(a, kb, ty, tx, colbuf, rdiag, &failflag);
    chol_block<1>(a, kb, ty, tx, colbuf, rdiag, &failflag);
    chol_block<2>(a, kb, ty, tx, colbuf, rdiag, &failflag);
    chol_block<3>(a, kb, ty, tx, colbuf, rdiag, &failflag);
    const int fail = failflag;
    __syncthreads();
    DIAG_STAMP(2);
    if (fail) {
        if (tid == 0 && info[z] == 0) info[z] = k0 + fail;
        // leave a benign unit block so downstream kernels stay finite; info reports the failure
#pragma unroll
        for (int ai = 0; ai < 4; ai++)
#pragma unroll
            for (int bi = 0; bi < 4; bi++) a[ai][bi] = (ty + 16 * ai == tx + 16 * bi) ? 1.0 : 0.0;
        if (tid < NB) rdiag[tid] = 1.0;
    }
#pragma unroll
    for (int ai = 0; ai < 4; ai++)
#pragma unroll
        for (int bi = 0; bi < 4; bi++) {
            const int i = ty + 16 * ai, j = tx + 16 * bi;
            S[i][j] = j <= i ? a[ai][bi] : 0.0;
            X[i][j] = 0.0;
        }
    __syncthreads();
    // inverse of the lower-triangular block, blocked 16 -> 32 -> 64:
    // (1) the four 16 x 16 diagonal sub-blocks by forward substitution, one thread per column, fully unrolled in registers;
    if (tid < NB) {
        const int o = tid & ~15, c = tid & 15;
        double x[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {
            double sacc = (i == c) ? 1.0 : 0.0;
#pragma unroll
            for (int k = 0; k < i; k++) sacc -= S[o + i][o + k] * x[k];
            x[i] = sacc * rdiag[o + i];
        }
#pragma unroll
        for (int i = 0; i < 16; i++) X[o + i][o + c] = x[i];
    }
    __syncthreads();
    // (2) merge adjacent blocks: X21 = -X22 (L21 X11).  Fixed trip counts over the (zero-padded) triangles so the loads
    //     pipeline; each thread owns an R x 2 patch of the outputs and reuses its operand loads across it.
    {   // 16 -> 32: two pairs, 2 x (16 x 16) outputs, one 1 x 2 patch per thread
        const int pr = tid >> 7, r = (tid >> 3) & 15, c0 = (tid & 7) * 2, o = pr * 32;
        double t0 = 0.0, t1 = 0.0;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const double l = S[o + 16 + r][o + k];
            t0 += l * X[o + k][o + c0];
            t1 += l * X[o + k][o + c0 + 1];
        }
        Tbuf[(pr * 16 + r) * 33 + c0] = t0;
        Tbuf[(pr * 16 + r) * 33 + c0 + 1] = t1;
        __syncthreads();
        t0 = t1 = 0.0;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const double x = X[o + 16 + r][o + 16 + k];
            t0 += x * Tbuf[(pr * 16 + k) * 33 + c0];
            t1 += x * Tbuf[(pr * 16 + k) * 33 + c0 + 1];
        }
        __syncthreads();
        X[o + 16 + r][o + c0] = -t0;
        X[o + 16 + r][o + c0 + 1] = -t1;
        __syncthreads();
    }
    {   // 32 -> 64: one pair, 32 x 32 outputs, one 2 x 2 patch per thread
        const int r0 = (tid >> 4) * 2, c0 = (tid & 15) * 2;
        double t[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll 8
        for (int k = 0; k < 32; k++) {
            const double l0 = S[32 + r0][k], l1 = S[32 + r0 + 1][k], x0 = X[k][c0], x1 = X[k][c0 + 1];
            t[0][0] += l0 * x0; t[0][1] += l0 * x1; t[1][0] += l1 * x0; t[1][1] += l1 * x1;
        }
        Tbuf[r0 * 33 + c0] = t[0][0]; Tbuf[r0 * 33 + c0 + 1] = t[0][1];
        Tbuf[(r0 + 1) * 33 + c0] = t[1][0]; Tbuf[(r0 + 1) * 33 + c0 + 1] = t[1][1];
        __syncthreads();
        t[0][0] = t[0][1] = t[1][0] = t[1][1] = 0.0;
#pragma unroll 8
        for (int k = 0; k < 32; k++) {
            const double x0 = X[32 + r0][32 + k], x1 = X[32 + r0 + 1][32 + k], u0 = Tbuf[k * 33 + c0], u1 = Tbuf[k * 33 + c0 + 1];
            t[0][0] += x0 * u0; t[0][1] += x0 * u1; t[1][0] += x1 * u0; t[1][1] += x1 * u1;
        }
        __syncthreads();
        X[32 + r0][c0] = -t[0][0]; X[32 + r0][c0 + 1] = -t[0][1];
        X[32 + r0 + 1][c0] = -t[1][0]; X[32 + r0 + 1][c0 + 1] = -t[1][1];
        __syncthreads();
    }
    DIAG_STAMP(3);
    double* Iz = inv ? inv + (long)z * strideI : nullptr;
    double* ITz = invT + (long)z * NB * NB;
    for (int idx = tid; idx < NB * NB; idx += blockDim.x) {
        const int i = idx >> 6, j = idx & 63;
        if (i < kb && j < kb) {
            Az[(long)i * lda + j] = S[i][j];
            if (Iz) Iz[(long)i * ldi + j] = X[i][j];
        }
        ITz[idx] = (i < kb && j < kb) ? X[j][i] : 0.0;  // invT[i][j] = inv[j][i]; rows/cols >= kb are zero
    }
    DIAG_STAMP(4);
}

__global__ void tril_zero_kernel(double* A, long lda, long strideA, int n) {
    double* Az = A + (long)blockIdx.z * strideA;
    const int i = blockIdx.y * blockDim.y + threadIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && j < n && j > i) Az[(long)i * lda + j] = 0.0;
}

__global__ void transpose_kernel(const double* in, long ldi, long strideI, double* out, long ldo, long strideO,
                                 int rows, int cols) {
    __shared__ double tile[32][33];
    const double* I = in + (long)blockIdx.z * strideI;
    double* O = out + (long)blockIdx.z * strideO;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int row = r0 + r, col = c0 + threadIdx.x;
        tile[r][threadIdx.x] = (row < rows && col < cols) ? I[(long)row * ldi + col] : 0.0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int orow = c0 + r, ocol = r0 + threadIdx.x;  // out is cols x rows
        if (orow < cols && ocol < rows) O[(long)orow * ldo + ocol] = tile[threadIdx.x][r];
    }
}

__global__ void copy2d_kernel(const double* src, long lds, long strideS, double* dst, long ldd, long strideD,
                              int rows, int cols) {
    const double* S = src + (long)blockIdx.z * strideS;
    double* D = dst + (long)blockIdx.z * strideD;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = blockIdx.y; i < rows; i += gridDim.y)
        if (j < cols) D[(long)i * ldd + j] = S[(long)i * lds + j];
}

__global__ void add_diag_kernel(double* A, long lda, long strideA, int n, const double* add) {
    const int z = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) A[(long)z * strideA + (long)i * lda + i] += add[z];
}

// A[z][i][i] += vec[z][i]
__global__ void add_diag_vec_kernel(double* A, long lda, long strideA, int n, const double* vec, long strideVec) {
    const int z = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) A[(long)z * strideA + (long)i * lda + i] += vec[(long)z * strideVec + i];
}
// dst[z][i][j] = (vec[z][i] + (add ? add[z] : 0)) * src[z][i][j]
__global__ void scale_rows_kernel(const double* src, double* dst, long ld, long stride, int rows, int cols,
                                  const double* vec, long strideVec, const double* add) {
    const int z = blockIdx.z, i = blockIdx.y;
    const double f = vec[(long)z * strideVec + i] + (add ? add[z] : 0.0);
    const long base = (long)z * stride + (long)i * ld;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < cols; j += gridDim.x * blockDim.x) dst[base + j] = f * src[base + j];
}

__global__ void fill_kernel(double* p, long count, double v) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < count; i += (long)gridDim.x * blockDim.x) p[i] = v;
}

// `count` doubles at the start of each of gridDim.y batch members `stride` apart (augmented layouts: one launch, not one per member)
__global__ void fill_strided_kernel(double* p, long stride, long count, double v) {
    double* q = p + (long)blockIdx.y * stride;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < count; i += (long)gridDim.x * blockDim.x) q[i] = v;
}

}  // namespace

long bo_linalg_scratch_doubles(int n, int batch) {
    const long np = bo_round_up(n, 8);
    // Bmat (2 NB x NB) | W^T x 2 (potrf panels, double-buffered for the look-ahead) | T (triangular inverse: the L21 X11 products of one recursion level, <= n^2/4 + slack)
    return (long)batch * ((long)2 * NB * NB + (long)np * NB + (long)NB * np + np * np / 2 + (long)NB * np);
}

int bo_fill(double* p, long count, double v, cudaStream_t stream) {
    if (count <= 0) return BO_OK;
    long blocks = (count + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    fill_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p, count, v);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_transpose(const double* in, long ldi, long strideI, double* out, long ldo, long strideO, int rows, int cols,
                 int batch, cudaStream_t stream) {
    if (rows <= 0 || cols <= 0) return BO_OK;
    dim3 grid((cols + 31) / 32, (rows + 31) / 32, batch), block(32, 8);
    transpose_kernel<<<grid, block, 0, stream>>>(in, ldi, strideI, out, ldo, strideO, rows, cols);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_copy2d(const double* src, long lds, long strideS, double* dst, long ldd, long strideD, int rows, int cols,
              int batch, cudaStream_t stream) {
    if (rows <= 0 || cols <= 0) return BO_OK;
    dim3 grid((cols + 127) / 128, rows < 1024 ? rows : 1024, batch);
    copy2d_kernel<<<grid, 128, 0, stream>>>(src, lds, strideS, dst, ldd, strideD, rows, cols);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_add_diag(double* A, long lda, long strideA, int n, int batch, const double* add, cudaStream_t stream) {
    if (!add || n <= 0) return BO_OK;
    dim3 grid((n + 127) / 128, batch);
    add_diag_kernel<<<grid, 128, 0, stream>>>(A, lda, strideA, n, add);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_add_diag_vec(double* A, long lda, long strideA, int n, int batch, const double* vec, long strideVec,
                    cudaStream_t stream) {
    if (!vec || n <= 0) return BO_OK;
    dim3 grid((n + 127) / 128, batch);
    add_diag_vec_kernel<<<grid, 128, 0, stream>>>(A, lda, strideA, n, vec, strideVec);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

int bo_scale_rows(const double* src, double* dst, long ld, long stride, int rows, int cols, int batch, const double* vec,
                  long strideVec, const double* add, cudaStream_t stream) {
    if (rows <= 0 || cols <= 0) return BO_OK;
    if (rows > 65535 || batch > 65535) return BO_ERR_UNSUPPORTED;
    dim3 grid((cols + 255) / 256 > 8 ? 8 : (cols + 255) / 256, rows, batch);
    scale_rows_kernel<<<grid, 256, 0, stream>>>(src, dst, ld, stride, rows, cols, vec, strideVec, add);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

namespace {
// Side stream + events for the look-ahead of bo_potrf_blocked (one set per device, created on first use).
struct PotrfAux {
    cudaStream_t side = nullptr;
    cudaEvent_t fork = nullptr, join[2] = {nullptr, nullptr};
};
PotrfAux* potrf_aux() {
    static PotrfAux aux[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    PotrfAux& a = aux[dev];
    if (!a.side) {
        if (cudaStreamCreateWithFlags(&a.side, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&a.join[0], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&a.join[1], cudaEventDisableTiming) != cudaSuccess)
            return nullptr;
    }
    return &a;
}
}  // namespace

// Blocked Cholesky with the trailing update taken off the critical path.  Per block column k:
//   diag(k)   one CTA: factors A_kk, inverts it, emits inv(L_kk)^T
//   panel(k)  P_k <- P_k inv(L_kk)^T in place (16-row tiles over ~n/16 SMs) + the transposed copy for the rank-NB updates
//   next(k)   block column k+1 only: A[k+1.., k+1] -= P_k P_k[k+1]^T   -- all diag(k+1) / panel(k+1) wait for
//   trail(k)  A[k+2.., k+2..] -= P_k P_k^T on a SIDE stream (forked/joined with events, still ordered inside the caller's
//             stream): it only has to land before next(k+1) touches block column k+2.
// The dependent chain per block column is three SHORT launches; the O(n^2) update overlaps the next column's chain.
// (Folding next(k) into diag/panel as in-CTA products was measured and dropped: a lone CTA runs a 64^3 product at one
// SM's FP64 rate -- 16.6k cycles in the diagonal kernel against ~4 us for the 124-CTA launch.)
int bo_potrf_blocked(double* A, long lda, long strideA, int n, int batch, double* Xinv, long ldx, long strideX,
                     double* scratch, int* info, cudaStream_t stream) {
    const long np = bo_round_up(n, 8);
    constexpr size_t kPotrfSmem = sizeof(double) * 2 * NB * (NB + 1);
    static bool attr_set = false;
    if (!attr_set) {
        BO_CUDA_TRY(cudaFuncSetAttribute(potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kPotrfSmem));
        attr_set = true;
    }
    PotrfAux* aux = potrf_aux();
    if (!aux) return BO_ERR_CUDA;
    double* invT = scratch;                                   // [batch][NB x NB] (region sized 2 NB x NB)
    double* WT[2] = {invT + (long)batch * 2 * NB * NB, invT + (long)batch * 2 * NB * NB + (long)batch * NB * np};  // [batch][NB x np] x 2
    if (Xinv) {
        if (strideX == (long)n * ldx) {
            BO_TRY(bo_fill(Xinv, (long)batch * n * ldx, 0.0, stream));
        } else {
            // zero n rows of ldx per member (strideX exceeds n*ldx: augmented layouts); gridDim.y <= 65535
            const long count = (long)n * ldx;
            long blocks = (count + 255) / 256;
            if (blocks > 148 * 4) blocks = 148 * 4;
            for (int z0 = 0; z0 < batch; z0 += 65535) {
                const int nz = (batch - z0 < 65535) ? batch - z0 : 65535;
                fill_strided_kernel<<<dim3((unsigned)blocks, (unsigned)nz), 256, 0, stream>>>(Xinv + (long)z0 * strideX,
                                                                                              strideX, count, 0.0);
                g_bo_launch_count++;
                BO_CUDA_TRY(cudaGetLastError());
            }
        }
    }
    bool pending[2] = {false, false};
    int k = 0;
    for (int k0 = 0; k0 < n; k0 += NB, k++) {
        const int kb = (n - k0 < NB) ? n - k0 : NB;
        double* Akk = A + (long)k0 * lda + k0;
        potrf_diag_kernel<<<batch, 256, kPotrfSmem, stream>>>(Akk, lda, strideA, kb, k0,
                                                      Xinv ? Xinv + (long)k0 * ldx + k0 : nullptr, ldx, strideX, invT,
                                                      info);
        g_bo_launch_count++;
        BO_CUDA_TRY(cudaGetLastError());
        const int rows = n - k0 - kb;
        if (rows <= 0) break;
        double* P = A + (long)(k0 + kb) * lda + k0;
        // P <- P inv(L_kk)^T in place (BN = 64: each CTA overwrites exactly the rows it has read), P^T on the side
        BoGemm g{};
        g.A = P; g.lda = lda; g.strideA = strideA;
        g.B = invT; g.ldb = NB; g.strideB = (long)NB * NB;
        g.C = P; g.ldc = lda; g.strideC = strideA;
        g.CT = WT[k & 1]; g.ldct = np; g.strideCT = (long)NB * np;
        g.alpha = 1.0; g.beta = 0.0; g.M = rows; g.N = kb; g.K = kb; g.batch = batch;
        g.epi = BO_EPI_LINEAR; g.tile64 = 1;
        BO_TRY(bo_gemm_launch(g, stream));
        const int nb1 = rows < NB ? rows : NB;  // width of block column k+1
        const int rows2 = rows - NB;            // rows / columns right of it
        if (rows2 > 0) {  // trail(k) on the side stream, behind trail(k-1) and this panel
            BO_CUDA_TRY(cudaEventRecord(aux->fork, stream));
            BO_CUDA_TRY(cudaStreamWaitEvent(aux->side, aux->fork, 0));
            double* A22 = A + (long)(k0 + kb + NB) * lda + (k0 + kb + NB);
            BoGemm u{};
            u.A = P + (long)NB * lda; u.lda = lda; u.strideA = strideA;
            u.B = WT[k & 1] + NB; u.ldb = np; u.strideB = (long)NB * np;
            u.C = A22; u.ldc = lda; u.strideC = strideA;
            u.D = A22; u.ldd = lda; u.strideD = strideA;
            u.alpha = -1.0; u.beta = 1.0; u.M = rows2; u.N = rows2; u.K = kb; u.batch = batch;
            u.out_lower = 1; u.epi = BO_EPI_LINEAR;
            BO_TRY(bo_gemm_launch(u, aux->side));
            BO_CUDA_TRY(cudaEventRecord(aux->join[k & 1], aux->side));
            pending[k & 1] = true;
        }
        // next(k): block column k+1 must hold trail(k-1)'s contribution first (issued one iteration ago on the side stream)
        if (pending[(k + 1) & 1]) {
            BO_CUDA_TRY(cudaStreamWaitEvent(stream, aux->join[(k + 1) & 1], 0));
            pending[(k + 1) & 1] = false;
        }
        double* A11 = A + (long)(k0 + kb) * lda + (k0 + kb);
        BoGemm v{};
        v.A = P; v.lda = lda; v.strideA = strideA;
        v.B = WT[k & 1]; v.ldb = np; v.strideB = (long)NB * np;
        v.C = A11; v.ldc = lda; v.strideC = strideA;
        v.D = A11; v.ldd = lda; v.strideD = strideA;
        v.alpha = -1.0; v.beta = 1.0; v.M = rows; v.N = nb1; v.K = kb; v.batch = batch;
        v.epi = BO_EPI_LINEAR; v.tile64 = 1;
        BO_TRY(bo_gemm_launch(v, stream));
    }
    for (int s = 0; s < 2; s++)
        if (pending[s]) BO_CUDA_TRY(cudaStreamWaitEvent(stream, aux->join[s], 0));
    dim3 block(32, 8), grid((n + 31) / 32, (n + 7) / 8, batch);
    tril_zero_kernel<<<grid, block, 0, stream>>>(A, lda, strideA, n);
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

// Recursive-doubling triangular inverse.  The NB x NB diagonal blocks of X already hold inv(L_kk).  Level s (= NB, 2 NB,
// ...) merges adjacent s-blocks:  inv([[L11, 0], [L21, L22]]) = [[X11, 0], [-X22 (L21 X11), X22]].  All pairs of a level
// are independent, so a level is TWO batched DMMA launches over every pair (plus one ragged pair at the matrix edge)
// instead of two launches per block row with at most n/64 CTAs each: log2(n/NB) levels, full-width grids.
int bo_trtri_blocked(const double* L, long ldl, long strideL, int n, int batch, double* X, long ldx, long strideX,
                     double* scratch, cudaStream_t stream) {
    const long np = bo_round_up(n, 8);
    double* T = scratch + (long)batch * ((long)2 * NB * NB + 2 * np * NB);  // [batch][<= np*np/2 + NB*np]
    const long strideT = np * np / 2 + (long)NB * np;
    // one pair: rows r of the second block (r <= s), first block s x s at offset o
    auto pair = [&](long o, long s, long r, int nb_pairs, long pstrideL, long pstrideX, long pstrideT, int bcount,
                    long bstrideL, long bstrideX, long bstrideT) -> int {
        // T = L21 X11   (r x s, X11 lower triangular)
        BoGemm a{};
        a.A = L + (o + s) * ldl + o; a.lda = ldl; a.strideA = nb_pairs > 1 ? pstrideL : bstrideL;
        a.B = X + o * ldx + o; a.ldb = ldx; a.strideB = nb_pairs > 1 ? pstrideX : bstrideX;
        a.C = T; a.ldc = s; a.strideC = nb_pairs > 1 ? pstrideT : bstrideT;
        a.alpha = 1.0; a.M = (int)r; a.N = (int)s; a.K = (int)s; a.batch = nb_pairs > 1 ? nb_pairs : bcount;
        a.tri = BO_TRI_B_LOWER; a.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(a, stream));
        // X21 = -X22 T  (X22 lower triangular r x r)
        BoGemm b{};
        b.A = X + (o + s) * ldx + (o + s); b.lda = ldx; b.strideA = a.strideB;
        b.B = T; b.ldb = s; b.strideB = a.strideC;
        b.C = X + (o + s) * ldx + o; b.ldc = ldx; b.strideC = a.strideB;
        b.alpha = -1.0; b.M = (int)r; b.N = (int)s; b.K = (int)r; b.batch = a.batch;
        b.tri = BO_TRI_A_LOWER; b.epi = BO_EPI_LINEAR;
        return bo_gemm_launch(b, stream);
    };
    for (long s = NB; s < n; s *= 2) {
        const long full = n / (2 * s);       // pairs with two complete s-blocks
        const long rem = n - full * 2 * s;   // rows left after them
        if (batch == 1) {
            if (full > 0)
                BO_TRY(pair(0, s, s, (int)full, 2 * s * ldl + 2 * s, 2 * s * ldx + 2 * s, s * s, 1, 0, 0, 0));
            if (rem > s) {  // ragged pair: a complete first block and rem - s rows of the second
                T += full * s * s;
                BO_TRY(pair(full * 2 * s, s, rem - s, 1, 0, 0, 0, 1, 0, 0, 0));
                T -= full * s * s;
            }
        } else {  // several matrices: batch index = matrix, pairs one after the other
            for (long p = 0; p < full; p++) BO_TRY(pair(p * 2 * s, s, s, 1, 0, 0, 0, batch, strideL, strideX, strideT));
            if (rem > s) BO_TRY(pair(full * 2 * s, s, rem - s, 1, 0, 0, 0, batch, strideL, strideX, strideT));
        }
    }
    return BO_OK;
}
