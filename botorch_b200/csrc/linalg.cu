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

// One CTA per batch member: Cholesky of the kb x kb diagonal block + its inverse (and inverse^T).
// 256 threads as a 16 x 16 grid; thread (ty, tx) keeps the 4 x 4 elements (ty + 16 a, tx + 16 b) of the NB x NB block in
// registers.  Step j: the owners of column j publish it through a double-buffered shared vector (ONE barrier per step),
// every thread forms l = column / sqrt(pivot) and applies the rank-1 update to its registers.  The triangular inverse
// then runs with 4 lanes per column (partial dot products combined by shuffles).
// (The first version kept the block in shared memory with three barriers and a div/mod-indexed update per step and
// inverted with one serial thread per column: 94 us per block, 55% of a whole factorisation at n = 2000.)
static_assert(NB == 64, "potrf_diag_kernel is laid out for 64 x 64 diagonal blocks");
__global__ void __launch_bounds__(256) potrf_diag_kernel(double* A, long lda, long strideA, int kb, int k0,
                                                          double* inv, long ldi, long strideI, double* invT,
                                                          int* info) {
    extern __shared__ __align__(16) double potrf_sm[];
    double (*S)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm);
    double (*X)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm + NB * (NB + 1));
    __shared__ double colbuf[2][NB];
    __shared__ double rdiag[NB];
    __shared__ double Tbuf[32 * 32];
    const int z = blockIdx.x, tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;
    double* Az = A + (long)z * strideA;
    double a[4][4];
#pragma unroll
    for (int ai = 0; ai < 4; ai++)
#pragma unroll
        for (int bi = 0; bi < 4; bi++) {
            const int i = ty + 16 * ai, j = tx + 16 * bi;
            a[ai][bi] = (i < kb && j <= i) ? Az[(long)i * lda + j] : ((i == j) ? 1.0 : 0.0);
        }
    int fail = 0;
#pragma unroll
    for (int bj = 0; bj < 4; bj++) {
        for (int jj = 0; jj < 16; jj++) {
            const int j = bj * 16 + jj;
            double* buf = colbuf[j & 1];
            if (tx == jj) {
#pragma unroll
                for (int ai = 0; ai < 4; ai++) buf[ty + 16 * ai] = a[ai][bj];
            }
            __syncthreads();
            const double ajj = buf[j];
            if (!(ajj > 0.0) && j < kb) { fail = j + 1; break; }  // uniform: every thread reads the same pivot
            const double rd = rsqrt(ajj);  // one MUFU + Newton chain instead of sqrt followed by a division
            const double d = ajj * rd;
            if (tid == 0) rdiag[j] = rd;
            double li[4], lk[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int i = ty + 16 * q, k = tx + 16 * q;
                li[q] = i > j ? buf[i] * rd : 0.0;
                lk[q] = k > j ? buf[k] * rd : 0.0;
            }
#pragma unroll
            for (int ai = 0; ai < 4; ai++)
#pragma unroll
                for (int bi = 0; bi < 4; bi++) a[ai][bi] -= li[ai] * lk[bi];
            if (tx == jj) {
#pragma unroll
                for (int ai = 0; ai < 4; ai++) {
                    const int i = ty + 16 * ai;
                    if (i == j) a[ai][bj] = d;
                    else if (i > j) a[ai][bj] = li[ai];
                }
            }
        }
        if (fail) break;
    }
    __syncthreads();
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
    // (2) merge adjacent blocks: X21 = -X22 (L21 X11), every output element on its own thread
#pragma unroll
    for (int sz = 16; sz < NB; sz *= 2) {
        const int per = sz * sz, total = per * (NB / (2 * sz));
        for (int e = tid; e < total; e += 256) {  // T = L21 X11
            const int pr = e / per, r = (e % per) / sz, c = e % sz, o = pr * 2 * sz;
            double acc = 0.0;
            for (int k = c; k < sz; k++) acc += S[o + sz + r][o + k] * X[o + k][o + c];
            Tbuf[e] = acc;
        }
        __syncthreads();
        for (int e = tid; e < total; e += 256) {  // X21 = -X22 T
            const int pr = e / per, r = (e % per) / sz, c = e % sz, o = pr * 2 * sz;
            double acc = 0.0;
            for (int k = 0; k <= r; k++) acc += X[o + sz + r][o + sz + k] * Tbuf[pr * per + k * sz + c];
            X[o + sz + r][o + c] = -acc;
        }
        __syncthreads();
    }
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

}  // namespace

long bo_linalg_scratch_doubles(int n, int batch) {
    const long np = bo_round_up(n, 8);
    // invT | W | W^T (potrf panels) | T (triangular inverse: the L21 X11 products of one recursion level, <= n^2/4 + slack)
    return (long)batch * ((long)NB * NB + (long)np * NB + (long)NB * np + np * np / 2 + (long)NB * np);
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
    double* invT = scratch;                              // [batch][NB*NB]
    double* WT = invT + (long)batch * NB * NB + (long)batch * np * NB;  // [batch][NB x np]
    if (Xinv) {
        for (int z = 0; z < batch; z++)  // zero n rows of ldx (strideX may exceed n*ldx: augmented layouts)
            BO_TRY(bo_fill(Xinv + (long)z * strideX, (long)n * ldx, 0.0, stream));
    }
    for (int k0 = 0; k0 < n; k0 += NB) {
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
        // P <- P * inv(L_kk)^T in place (64-wide tiles: each CTA overwrites exactly the rows it has read), W^T = P^T on the side
        BoGemm g{};
        g.A = P; g.lda = lda; g.strideA = strideA;
        g.B = invT; g.ldb = NB; g.strideB = (long)NB * NB;
        g.C = P; g.ldc = lda; g.strideC = strideA;
        g.CT = WT; g.ldct = np; g.strideCT = (long)NB * np;
        g.alpha = 1.0; g.beta = 0.0; g.M = rows; g.N = kb; g.K = kb; g.batch = batch;
        g.epi = BO_EPI_LINEAR; g.tile64 = 1;
        BO_TRY(bo_gemm_launch(g, stream));
        // A22 -= P P^T   (lower block-triangle only)
        double* A22 = A + (long)(k0 + kb) * lda + (k0 + kb);
        BoGemm u{};
        u.A = P; u.lda = lda; u.strideA = strideA;
        u.B = WT; u.ldb = np; u.strideB = (long)NB * np;
        u.C = A22; u.ldc = lda; u.strideC = strideA;
        u.D = A22; u.ldd = lda; u.strideD = strideA;
        u.alpha = -1.0; u.beta = 1.0; u.M = rows; u.N = rows; u.K = kb; u.batch = batch;
        u.out_lower = 1; u.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(u, stream));
    }
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
    double* T = scratch + (long)batch * ((long)NB * NB + 2 * np * NB);  // [batch][<= np*np/2 + NB*np]
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
