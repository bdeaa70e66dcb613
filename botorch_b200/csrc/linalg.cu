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
__global__ void __launch_bounds__(256) potrf_diag_kernel(double* A, long lda, long strideA, int kb, int k0,
                                                          double* inv, long ldi, long strideI, double* invT,
                                                          int* info) {
    extern __shared__ __align__(16) double potrf_sm[];
    double (*S)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm);
    double (*X)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm + NB * (NB + 1));
    const int z = blockIdx.x, tid = threadIdx.x;
    double* Az = A + (long)z * strideA;
    for (int idx = tid; idx < NB * NB; idx += blockDim.x) {
        const int i = idx / NB, j = idx % NB;
        S[i][j] = (i < kb && j <= i) ? Az[(long)i * lda + j] : ((i == j) ? 1.0 : 0.0);
        X[i][j] = 0.0;
    }
    __syncthreads();
    int fail = 0;
    for (int j = 0; j < kb; j++) {
        const double ajj = S[j][j];
        if (!(ajj > 0.0)) { fail = j + 1; break; }  // uniform: every thread reads the same pivot
        __syncthreads();
        const double d = sqrt(ajj);
        const double rd = 1.0 / d;
        for (int i = j + tid; i < kb; i += blockDim.x) S[i][j] = (i == j) ? d : S[i][j] * rd;
        __syncthreads();
        const int m = kb - j - 1;  // trailing block size
        for (int idx = tid; idx < m * m; idx += blockDim.x) {
            const int i = j + 1 + idx / m, k = j + 1 + idx % m;
            if (k <= i) S[i][k] -= S[i][j] * S[k][j];
        }
        __syncthreads();
    }
    if (fail) {
        if (tid == 0 && info[z] == 0) info[z] = k0 + fail;
        // leave a benign unit block so downstream kernels stay finite; info reports the failure
        __syncthreads();
        for (int idx = tid; idx < NB * NB; idx += blockDim.x) {
            const int i = idx / NB, j = idx % NB;
            S[i][j] = (i == j) ? 1.0 : 0.0;
        }
        __syncthreads();
    }
    // inverse of the lower-triangular block, one column per thread (forward substitution)
    if (tid < kb) {
        const int c = tid;
        X[c][c] = 1.0 / S[c][c];
        for (int i = c + 1; i < kb; i++) {
            double s = 0.0;
            for (int k = c; k < i; k++) s += S[i][k] * X[k][c];
            X[i][c] = -s / S[i][i];
        }
    }
    __syncthreads();
    double* Iz = inv ? inv + (long)z * strideI : nullptr;
    double* ITz = invT + (long)z * NB * NB;
    for (int idx = tid; idx < kb * kb; idx += blockDim.x) {
        const int i = idx / kb, j = idx % kb;
        Az[(long)i * lda + j] = (j <= i) ? S[i][j] : 0.0;
        if (Iz) Iz[(long)i * ldi + j] = X[i][j];
    }
    for (int idx = tid; idx < NB * NB; idx += blockDim.x) {
        const int i = idx / NB, j = idx % NB;
        ITz[idx] = X[j][i];  // invT[i][j] = inv[j][i]; rows/cols >= kb are zero
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
    return (long)batch * ((long)NB * NB + (long)np * NB + (long)NB * np);
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
    double* W = invT + (long)batch * NB * NB;            // [batch][np x NB]
    double* WT = W + (long)batch * np * NB;              // [batch][NB x np]
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
        // W = P * inv(L_kk)^T  (and W^T for the rank-kb update)
        BoGemm g{};
        g.A = P; g.lda = lda; g.strideA = strideA;
        g.B = invT; g.ldb = NB; g.strideB = (long)NB * NB;
        g.C = W; g.ldc = NB; g.strideC = np * NB;
        g.CT = WT; g.ldct = np; g.strideCT = (long)NB * np;
        g.alpha = 1.0; g.beta = 0.0; g.M = rows; g.N = kb; g.K = kb; g.batch = batch;
        g.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(g, stream));
        BO_TRY(bo_copy2d(W, NB, np * NB, P, lda, strideA, rows, kb, batch, stream));
        // A22 -= W W^T   (lower block-triangle only)
        double* A22 = A + (long)(k0 + kb) * lda + (k0 + kb);
        BoGemm u{};
        u.A = W; u.lda = NB; u.strideA = np * NB;
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

int bo_trtri_blocked(const double* L, long ldl, long strideL, int n, int batch, double* X, long ldx, long strideX,
                     double* scratch, cudaStream_t stream) {
    const long np = bo_round_up(n, 8);
    double* T = scratch + (long)batch * NB * NB + (long)batch * np * NB;  // reuse the W^T region: [batch][NB x np]
    for (int r0 = NB; r0 < n; r0 += NB) {
        const int rb = (n - r0 < NB) ? n - r0 : NB;
        // T = L[r0:r0+rb, 0:r0] * X[0:r0, 0:r0]      (X lower triangular)
        BoGemm a{};
        a.A = L + (long)r0 * ldl; a.lda = ldl; a.strideA = strideL;
        a.B = X; a.ldb = ldx; a.strideB = strideX;
        a.C = T; a.ldc = np; a.strideC = (long)NB * np;
        a.alpha = 1.0; a.M = rb; a.N = r0; a.K = r0; a.batch = batch;
        a.tri = BO_TRI_B_LOWER; a.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(a, stream));
        // X[r0:r0+rb, 0:r0] = -inv(L_ii) * T
        BoGemm b{};
        b.A = X + (long)r0 * ldx + r0; b.lda = ldx; b.strideA = strideX;
        b.B = T; b.ldb = np; b.strideB = (long)NB * np;
        b.C = X + (long)r0 * ldx; b.ldc = ldx; b.strideC = strideX;
        b.alpha = -1.0; b.M = rb; b.N = r0; b.K = rb; b.batch = batch;
        b.tri = BO_TRI_A_LOWER; b.epi = BO_EPI_LINEAR;
        BO_TRY(bo_gemm_launch(b, stream));
    }
    return BO_OK;
}
