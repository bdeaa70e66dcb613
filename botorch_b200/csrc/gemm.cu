// Batched FP64 GEMM for sm_100a on the DMMA.8x8x4 tensor path (mma.sync m8n8k4 f64 -- the only
// FP64 tensor-core instruction Blackwell has; tcgen05 has no f64 kind), cp.async multi-stage
// shared-memory pipeline, triangular k-range skipping and fused epilogues (linear combine,
// RBF / Matern-5/2 covariance from an augmented squared-distance contraction).
//
// Replaces, on the qLogNEI hot path, the cuBLAS/MKL calls behind
//   gpytorch Kernel.covar_dist / sq_dist              (reference call site models/gpytorch.py:446)
//   gpytorch exact_predictive_mean / _covar           (fast_pred_var, models/utils/assorted.py:292)
//   torch.linalg.solve_triangular in low_rank.py:131  (as a multiply with the cached inverse factor)
//   new_Lq.matmul(base_samples) in low_rank.py:156
#include "common.cuh"
#include <cstdlib>
#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint, libcuda is not linked)

long g_bo_launch_count = 0;

// Optional per-launch CUDA-event timing of the GEMM kernel (bench.py's roofline leg): events are recorded on the
// launching stream around every dgemm_dmma_kernel launch and summed by bo_prof_read after a synchronise.
#include <vector>
namespace {
struct ProfRec { cudaEvent_t a, b; };
int g_prof_on = 0;  // 0 off, 1 every GEMM launch, 2 only the TMA-staged kernel (the dominant one: bench.py's roofline)
std::vector<ProfRec> g_prof_recs;
size_t g_prof_used = 0;
double g_prof_flops = 0.0;
}  // namespace

extern "C" void bo_prof_enable(int on) {
    g_prof_on = on;
    g_prof_used = 0;
    g_prof_flops = 0.0;
}
// total milliseconds, launches and algorithmic flops of the GEMM launches since bo_prof_enable(1); caller synchronises
extern "C" int bo_prof_read(double* ms, long* launches, double* flops) {
    double tot = 0.0;
    for (size_t i = 0; i < g_prof_used; i++) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, g_prof_recs[i].a, g_prof_recs[i].b) != cudaSuccess) return BO_ERR_CUDA;
        tot += t;
    }
    if (ms) *ms = tot;
    if (launches) *launches = (long)g_prof_used;
    if (flops) *flops = g_prof_flops;
    return BO_OK;
}

// DMMA.8x8x4 issue-rate microbenchmark: the FP64 tensor roofline denominator, measured in the run that quotes it
// (bench.py times the launch with CUDA events).  8 independent accumulators per warp, 8 warps per CTA, 4 CTAs per SM.
namespace {
__global__ void __launch_bounds__(256) dmma_peak_kernel(double* out, int iters, double s) {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = 0.0;
    const double a = s * (threadIdx.x & 31), b = 1.0 - s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1])
                         : "d"(a), "d"(b));
    }
    double r = 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++) r += c[i][0] + c[i][1];
    if (r == 123.456) out[0] = r;  // keeps the chain alive without a store per thread
}
}  // namespace

// launches the microbenchmark on `stream`; *flops receives the FP64 flop count of the launch (2 x 256 FMA per warp-DMMA)
extern "C" int bo_dmma_peak_launch(int iters, double* scratch, double* flops, void* stream) {
    if (iters <= 0 || !scratch || !flops) return BO_ERR_INVALID;
    int dev = 0, sms = 148;
    BO_CUDA_TRY(cudaGetDevice(&dev));
    BO_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = sms * 4;
    dmma_peak_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(scratch, iters, 0.999);
    BO_CUDA_TRY(cudaGetLastError());
    *flops = 2.0 * 256.0 * 8.0 * (double)iters * 8.0 * (double)grid;
    return BO_OK;
}

namespace {

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

template <int BM_, int BN_, int BK_, int WM_, int WN_, int STAGES_>
struct Cfg {
    static constexpr int BM = BM_, BN = BN_, BK = BK_, WM = WM_, WN = WN_, STAGES = STAGES_;
    static constexpr int WARPS_M = BM / WM, WARPS_N = BN / WN;
    static constexpr int THREADS = WARPS_M * WARPS_N * 32;
    static constexpr int LDA_S = BK + 4;  // row pitch == 4 (mod 16) doubles: conflict-free LDS.64 fragments
    static constexpr int LDB_S = BN + 4;
    static constexpr int A_STAGE = BM * LDA_S, B_STAGE = BK * LDB_S;
    static constexpr size_t SMEM = (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double);
    static_assert(BK % 4 == 0 && (BK + 4) % 16 == 4, "BK");
    static_assert(BN % 16 == 0, "BN");
};

__device__ __forceinline__ double epilogue_value(const BoGemm& p, int z, double acc, int row, int col) {
    if (p.epi == BO_EPI_LINEAR) {
        double v = (p.alpha_vec ? p.alpha * p.alpha_vec[z] : p.alpha) * acc;
        if (p.D != nullptr) v += p.beta * p.D[(long)z * p.strideD + (long)row * p.ldd + col];
        if (p.E != nullptr) v *= p.E[(long)z * p.strideE + (long)row * p.lde + col];
        return v;
    }
    if (p.epi == BO_EPI_SINCOS) {
        double sn, cs;
        sincos(acc, &sn, &cs);
        const double sc = p.os ? p.alpha * sqrt(p.os[z]) : p.alpha;
        p.C2[(long)z * p.strideC2 + (long)row * p.ldc2 + col] = sc * cs;
        return sc * sn;
    }
    double r2 = acc > 0.0 ? acc : 0.0;
    const bool on_diag = (row == col);
    if (p.diag_exact && on_diag) r2 = 0.0;
    const double os = p.os ? p.os[z] : 1.0;
    double v;
    if (p.epi == BO_EPI_RBF) {
        v = os * exp(-0.5 * r2);
    } else {
        const double sqrt5 = 2.23606797749978969641;
        double r = sqrt(r2 > 1e-30 ? r2 : 1e-30);
        const double e = exp(-sqrt5 * r);
        v = os * ((sqrt5 * r + 1.0) + (5.0 / 3.0) * (r * r)) * e;
        if (p.C2)  // derivative factor of Matern-5/2 (zero below gpytorch's distance clamp)
            p.C2[(long)z * p.strideC2 + (long)row * p.ldc2 + col] = r2 < 1e-30 ? 0.0 : os * (5.0 / 3.0) * (1.0 + sqrt5 * r) * e;
    }
    if (p.diag_add && on_diag) v += p.diag_add[z];
    return v;
}

template <class C>
__global__ void __launch_bounds__(C::THREADS) dgemm_dmma_kernel(const BoGemm p) {
    extern __shared__ __align__(16) double smem[];
    double* As = smem;
    double* Bs = smem + (size_t)C::STAGES * C::A_STAGE;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int wm0 = (warp / C::WARPS_N) * C::WM;
    const int wn0 = (warp % C::WARPS_N) * C::WN;
    const int z = blockIdx.z;
    // heavy (large-m) row tiles first when A is lower triangular: better tail balance
    // grid = (column tiles, row tiles): CTAs are dispatched x-fastest, so all column tiles of the heaviest row tile go
    // first and the last wave holds only the lightest tiles (longest-processing-time order for triangular A)
    const int mt = (p.tri == BO_TRI_A_LOWER) ? (gridDim.y - 1 - blockIdx.y) : blockIdx.y;
    const int m0 = mt * C::BM;
    const int n0 = blockIdx.x * C::BN;
    if (p.out_lower && n0 > m0 + C::BM - 1) return;

    const double* __restrict__ A = p.A + (long)z * p.strideA;
    const double* __restrict__ B = p.B + (long)z * p.strideB;

    int kb = 0, ke = p.K;
    if (p.tri == BO_TRI_A_LOWER) ke = min(p.K, m0 + C::BM + p.tri_off);
    else if (p.tri == BO_TRI_A_UPPER) kb = m0 > p.tri_off ? ((m0 - p.tri_off) / C::BK) * C::BK : 0;
    else if (p.tri == BO_TRI_B_LOWER) kb = (n0 / C::BK) * C::BK;
    const int ntiles = ke > kb ? (ke - kb + C::BK - 1) / C::BK : 0;

    // Per-thread copy plan, fixed for the whole k loop: every thread moves NA 16-byte chunks of the A tile and NB of the B
    // tile per stage.  Source pointers, validity and shared-memory addresses are computed once; a k-tile then costs one
    // pointer add per chunk (the generic bounds logic only runs for the ragged last tile).
    constexpr int ACH = C::BK / 2, BCH = C::BN / 2;  // 16-byte chunks per A row / B row
    constexpr int NA = (C::BM * ACH) / C::THREADS, NB = (C::BK * BCH) / C::THREADS;
    static_assert((C::BM * ACH) % C::THREADS == 0 && (C::BK * BCH) % C::THREADS == 0, "copy plan");
    const double* a_src[NA];
    const double* b_src[NB];
    unsigned a_dst[NA], b_dst[NB];
    int a_bytes[NA], b_bytes[NB];
    const unsigned as0 = (unsigned)__cvta_generic_to_shared(As), bs0 = (unsigned)__cvta_generic_to_shared(Bs);
#pragma unroll
    for (int i = 0; i < NA; i++) {
        const int id = tid + i * C::THREADS, r = id / ACH, cc = id % ACH, row = m0 + r;
        a_bytes[i] = row < p.M ? 16 : 0;
        a_src[i] = A + (long)(row < p.M ? row : 0) * p.lda + 2 * cc + kb;
        a_dst[i] = as0 + (unsigned)((r * C::LDA_S + 2 * cc) * sizeof(double));
    }
#pragma unroll
    for (int i = 0; i < NB; i++) {
        const int id = tid + i * C::THREADS, r = id / BCH, cc = id % BCH, col = n0 + 2 * cc;
        b_bytes[i] = max(0, min(16, (p.N - col) * 8));
        b_src[i] = B + (long)(kb + r) * p.ldb + (b_bytes[i] > 0 ? col : 0);
        b_dst[i] = bs0 + (unsigned)((r * C::LDB_S + 2 * cc) * sizeof(double));
    }
    const long b_step = (long)C::BK * p.ldb;
    auto cp16 = [](unsigned dst, const void* src, int bytes) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(bytes));
    };
    auto load_tile = [&](int tile, int stage) {
        const int k0 = kb + tile * C::BK;
        const unsigned a_off = (unsigned)(stage * C::A_STAGE * sizeof(double));
        const unsigned b_off = (unsigned)(stage * C::B_STAGE * sizeof(double));
        if (k0 + C::BK <= ke) {  // interior tile: no k bounds
#pragma unroll
            for (int i = 0; i < NA; i++) cp16(a_dst[i] + a_off, a_src[i] + (long)tile * C::BK, a_bytes[i]);
#pragma unroll
            for (int i = 0; i < NB; i++) cp16(b_dst[i] + b_off, b_src[i] + tile * b_step, b_bytes[i]);
        } else {  // ragged last tile: clip (zero-fill) beyond ke
#pragma unroll
            for (int i = 0; i < NA; i++) {
                const int id = tid + i * C::THREADS, cc = id % ACH, k = k0 + 2 * cc;
                const int bytes = a_bytes[i] ? max(0, min(16, (ke - k) * 8)) : 0;
                cp16(a_dst[i] + a_off, bytes > 0 ? a_src[i] + (long)tile * C::BK : A, bytes);
            }
#pragma unroll
            for (int i = 0; i < NB; i++) {
                const int id = tid + i * C::THREADS, r = id / BCH;
                const int bytes = (k0 + r < ke) ? b_bytes[i] : 0;
                cp16(b_dst[i] + b_off, bytes > 0 ? b_src[i] + tile * b_step : B, bytes);
            }
        }
    };

    constexpr int TM = C::WM / 8, TN = C::WN / 8;
    double acc[TM][TN][2];
#pragma unroll
    for (int i = 0; i < TM; i++)
#pragma unroll
        for (int j = 0; j < TN; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
    for (int s = 0; s < C::STAGES - 1; s++) {
        if (s < ntiles) load_tile(s, s);
        cp_async_commit();
    }

    for (int kt = 0; kt < ntiles; kt++) {
        cp_async_wait<C::STAGES - 2>();
        __syncthreads();
        {
            const int nxt = kt + C::STAGES - 1;
            if (nxt < ntiles) load_tile(nxt, nxt % C::STAGES);
            cp_async_commit();
        }
        const double* as = As + (size_t)(kt % C::STAGES) * C::A_STAGE + (wm0 + g) * C::LDA_S + t;
        const double* bs = Bs + (size_t)(kt % C::STAGES) * C::B_STAGE + t * C::LDB_S + wn0 + g;
#pragma unroll
        for (int kk = 0; kk < C::BK; kk += 4) {
            double a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; i++) a[i] = as[i * 8 * C::LDA_S + kk];
#pragma unroll
            for (int j = 0; j < TN; j++) b[j] = bs[kk * C::LDB_S + j * 8];
#pragma unroll
            for (int i = 0; i < TM; i++)
#pragma unroll
                for (int j = 0; j < TN; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    cp_async_wait<0>();

    double* __restrict__ Cz = p.C + (long)z * p.strideC;
    double* __restrict__ CTz = p.CT ? p.CT + (long)z * p.strideCT : nullptr;
    constexpr bool kStageFits = (size_t)C::BM * (C::BN + 2) <= (size_t)C::STAGES * (C::A_STAGE + C::B_STAGE);
    if constexpr (kStageFits) if (p.epi != BO_EPI_LINEAR) {
        // Covariance epilogues (exp per element): stage the tile through shared memory and run ONE rolled copy of the
        // transcendental code with row-contiguous stores.  The fully unrolled per-fragment form is ~32 inlined exp()
        // bodies that each warp executes once -- ncu showed the K = d+2 launches stalled on instruction fetch.
        constexpr int LDC_S = C::BN + 2;
        __syncthreads();
        double* Cs = smem;
#pragma unroll
        for (int i = 0; i < TM; i++)
#pragma unroll
            for (int j = 0; j < TN; j++)
                *reinterpret_cast<double2*>(Cs + (wm0 + i * 8 + g) * LDC_S + wn0 + j * 8 + 2 * t) =
                    make_double2(acc[i][j][0], acc[i][j][1]);
        __syncthreads();
        if (p.epi == BO_EPI_RBF && !p.diag_exact && p.diag_add == nullptr && CTz == nullptr && m0 + C::BM <= p.M &&
            n0 + C::BN <= p.N) {
            // interior tile of the plain RBF cross-covariance (the K = d+2 launch of every forward): two columns per
            // thread and iteration, 16-byte loads / stores, four independent exp chains in flight, no per-element bounds or
            // mode tests -- the generic loop below spent more issue slots on those than on the exp itself (ncu: 55 % issue
            // utilisation, FP64 pipe 14 %)
            const double os = p.os ? p.os[z] : 1.0;
            constexpr int HALF = C::BN / 2;
#pragma unroll 4
            for (int e = tid; e < C::BM * HALF; e += C::THREADS) {
                const int r = e / HALF, c2 = (e % HALF) * 2;
                const double2 a = *reinterpret_cast<const double2*>(Cs + r * LDC_S + c2);
                const double v0 = os * exp(-0.5 * (a.x > 0.0 ? a.x : 0.0));
                const double v1 = os * exp(-0.5 * (a.y > 0.0 ? a.y : 0.0));
                *reinterpret_cast<double2*>(Cz + (long)(m0 + r) * p.ldc + n0 + c2) = make_double2(v0, v1);
            }
            return;
        }
#pragma unroll 2
        for (int e = tid; e < C::BM * C::BN; e += C::THREADS) {
            const int r = e / C::BN, c = e % C::BN;
            const int row = m0 + r, col = n0 + c;
            if (row >= p.M || col >= p.N) continue;
            const double v = epilogue_value(p, z, Cs[r * LDC_S + c], row, col);
            Cz[(long)row * p.ldc + col] = v;
            if (CTz) CTz[(long)col * p.ldct + row] = v;
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < TM; i++) {
        const int row = m0 + wm0 + i * 8 + g;
        if (row >= p.M) continue;
#pragma unroll
        for (int j = 0; j < TN; j++) {
            const int col = n0 + wn0 + j * 8 + 2 * t;
            if (col >= p.N) continue;
            const double v0 = epilogue_value(p, z, acc[i][j][0], row, col);
            if (col + 1 < p.N) {
                const double v1 = epilogue_value(p, z, acc[i][j][1], row, col + 1);
                *reinterpret_cast<double2*>(Cz + (long)row * p.ldc + col) = make_double2(v0, v1);
                if (CTz) {
                    CTz[(long)col * p.ldct + row] = v0;
                    CTz[(long)(col + 1) * p.ldct + row] = v1;
                }
            } else {
                Cz[(long)row * p.ldc + col] = v0;
                if (CTz) CTz[(long)col * p.ldct + row] = v0;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// TMA-staged variant: operand tiles are moved by the tensor memory accelerator (cp.async.bulk.tensor + mbarrier
// full/empty pairs) issued by ONE producer warp; the consumer warps only wait on a stage's `full` barrier, run their
// DMMA.8x8x4 sequence out of shared memory and release the stage -- no CTA-wide barrier and no address arithmetic in the
// main loop.  FP64 has no tcgen05 kind, so the math stays on mma.sync m8n8k4; only the data movement changes.
//
// Shared-memory layout (dense, no padding, no swizzle) chosen so that every fragment load of a warp is ONE contiguous
// 256-byte segment (conflict-free): the tensor maps present the operands as rank-4 tensors
//   A: {4 (k within a k4 group), M rows, K/4 groups, batch}   box {4, BM, BK/4, 1}  -> As[k4][row][4]
//   B: {8 (column within an 8-column group), K rows, N/8 groups, batch}   box {8, BK, BN/8, 1}  -> Bs[n8][k][8]
// so lane (g, t) of a warp reads A[row0 + g][k4*4 + t] at ((k4*BM + row0 + g)*4 + t) and B[k + t][col0 + g] at
// ((n8*BK + k + t)*8 + g).  Out-of-range rows / k rows / batch are zero-filled by the TMA unit itself.
// ------------------------------------------------------------------------------------------------------------------
template <int BM_, int BN_, int BK_, int WM_, int WN_, int STAGES_, int MINB_>
struct TCfg {
    static constexpr int BM = BM_, BN = BN_, BK = BK_, WM = WM_, WN = WN_, STAGES = STAGES_, MINB = MINB_;
    static constexpr int WARPS_M = BM / WM, WARPS_N = BN / WN;
    static constexpr int CONSUMERS = WARPS_M * WARPS_N;
    static constexpr int THREADS = (CONSUMERS + 1) * 32;  // + the producer warp
    static constexpr int A_STAGE = BM * BK, B_STAGE = BK * BN;  // doubles
    static constexpr unsigned STAGE_BYTES = (unsigned)((A_STAGE + B_STAGE) * sizeof(double));
    static constexpr size_t SMEM = (size_t)STAGES * STAGE_BYTES + 2 * STAGES * sizeof(unsigned long long) + 128;
    static_assert((size_t)CONSUMERS * (WN / 8) * 64 * sizeof(double) <= (size_t)STAGES * STAGE_BYTES, "gram partials");
    static_assert(BK % 4 == 0 && BN % 8 == 0 && BM <= 256 && BK <= 256 && BN / 8 <= 256, "TMA box limits");
};

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_4d(unsigned dst, const CUtensorMap* map, unsigned bar, int c0, int c1, int c2,
                                            int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];\n" ::
            "r"(dst), "l"(reinterpret_cast<unsigned long long>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

// MODE: 0 plain, 1 with the gram epilogue (BoGemm::gram), 2 with the mid-loop column-group transform (BoGemm::post).
// Separate instantiations: the fused pieces must not cost the plain kernel registers.
template <class C, int MODE>
__global__ void __launch_bounds__(C::THREADS, C::MINB)
dgemm_dmma_tma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2, const BoGemm p,
                      const int za_step, const int zb_step) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // 128-byte aligned tile area (TMA destination alignment), then the barriers
    const unsigned raw = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (128u - (raw & 127u)) & 127u;
    double* tiles = reinterpret_cast<double*>(smem_raw + pad);
    const unsigned tiles_s = raw + pad;
    const unsigned bars_s = tiles_s + (unsigned)C::STAGES * C::STAGE_BYTES;  // full[STAGES], empty[STAGES]

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int z = blockIdx.z;
    const int mt = (p.tri == BO_TRI_A_LOWER) ? (gridDim.y - 1 - blockIdx.y) : blockIdx.y;
    const int m0 = mt * C::BM;
    const int n0 = blockIdx.x * C::BN;
    if (p.out_lower && n0 > m0 + C::BM - 1) return;

    int kb = 0, ke = p.K;
    if (p.tri == BO_TRI_A_LOWER) ke = min(p.K, m0 + C::BM + p.tri_off);
    else if (p.tri == BO_TRI_A_UPPER) kb = m0 > p.tri_off ? ((m0 - p.tri_off) / C::BK) * C::BK : 0;
    else if (p.tri == BO_TRI_B_LOWER) kb = (n0 / C::BK) * C::BK;
    // fused adjoint (p.post): tiles of [post_k0, ke) first (through tmA / tmB), then the tiles of [0, post_k0) through the
    // second pair of maps, whose k extent ends at post_k0 (the tile that straddles it is zero-filled beyond)
    constexpr bool has_post = (MODE == 2 || MODE == 3);
    if (has_post) kb = p.post_k0;
    const int n_first = ke > kb ? (ke - kb + C::BK - 1) / C::BK : 0;
    const int ntiles = n_first + (has_post ? (p.post_k0 + C::BK - 1) / C::BK : 0);

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; s++) {
            mbar_init(bars_s + 8u * s, 1);                          // full: the producer's expect_tx arrival
            mbar_init(bars_s + 8u * (C::STAGES + s), C::CONSUMERS);  // empty: one arrival per consumer warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();

    constexpr int TM = C::WM / 8, TN = C::WN / 8;
    double acc[TM][TN][2];
    const int wm0 = (warp / C::WARPS_N) * C::WM;
    const int wn0 = (warp % C::WARPS_N) * C::WN;

    if (warp == C::CONSUMERS) {
        // ---- producer: one lane streams the k-tiles of this CTA's (row tile, column tile) through the ring ----
        if (lane == 0) {
            for (int kt = 0; kt < ntiles; kt++) {
                const int s = kt % C::STAGES;
                if (kt >= C::STAGES) mbar_wait(bars_s + 8u * (C::STAGES + s), ((kt / C::STAGES) - 1) & 1);
                const unsigned full = bars_s + 8u * s;
                mbar_expect_tx(full, C::STAGE_BYTES);
                const unsigned a_dst = tiles_s + (unsigned)s * C::STAGE_BYTES;
                const unsigned b_dst = a_dst + (unsigned)(C::A_STAGE * sizeof(double));
                if (kt < n_first) {
                    const int k0 = kb + kt * C::BK;
                    tma_load_4d(a_dst, &tmA, full, 0, m0, k0 / 4, z * za_step);
                    tma_load_4d(b_dst, &tmB, full, 0, k0, n0 / 8, z * zb_step);
                } else {
                    const int k0 = (kt - n_first) * C::BK;
                    tma_load_4d(a_dst, &tmA2, full, 0, m0, k0 / 4, z * za_step);
                    tma_load_4d(b_dst, &tmB2, full, 0, k0, n0 / 8, z * zb_step);
                }
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < TM; i++)
#pragma unroll
            for (int j = 0; j < TN; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
        for (int kt = 0; kt < ntiles; kt++) {
            if (has_post && kt == n_first) {
                // acc <- acc * P per 8-column group, on the tensor path: the accumulator block (8 rows x 8 columns, lane
                // (g, t) holds columns 2t, 2t+1 of row g) is re-laid as the A operand of two m8n8k4 steps by 4 shuffles
                // (column c lives in lane (g, c / 2), slot c % 2); P supplies the B fragments
                const double* P = p.post + (long)z * p.stridePost;
                const int src_lo = (lane & ~3) | (t >> 1), src_hi = src_lo | 2;
#pragma unroll
                for (int j = 0; j < TN; j++) {
                    const int col0 = n0 + wn0 + j * 8;
                    if (col0 >= p.N) continue;
                    const double* Pg = P + (long)(col0 / 8) * 64;
                    const double pb0 = Pg[t * 8 + g], pb1 = Pg[(4 + t) * 8 + g];
#pragma unroll
                    for (int i = 0; i < TM; i++) {
                        const double x00 = __shfl_sync(0xffffffffu, acc[i][j][0], src_lo);
                        const double x01 = __shfl_sync(0xffffffffu, acc[i][j][1], src_lo);
                        const double x10 = __shfl_sync(0xffffffffu, acc[i][j][0], src_hi);
                        const double x11 = __shfl_sync(0xffffffffu, acc[i][j][1], src_hi);
                        double r0 = 0.0, r1 = 0.0;
                        dmma884(r0, r1, (t & 1) ? x01 : x00, pb0);
                        dmma884(r0, r1, (t & 1) ? x11 : x10, pb1);
                        acc[i][j][0] = r0;
                        acc[i][j][1] = r1;
                    }
                }
            }
            const int s = kt % C::STAGES;
            mbar_wait(bars_s + 8u * s, (kt / C::STAGES) & 1);
            const double* as = tiles + (size_t)s * (C::A_STAGE + C::B_STAGE) + (wm0 + g) * 4 + t;
            const double* bs = tiles + (size_t)s * (C::A_STAGE + C::B_STAGE) + C::A_STAGE + ((wn0 / 8) * C::BK + t) * 8 + g;
            // k-tiles that cross the diagonal of a triangular A: the 8-row blocks whose entries are all zero over a k4 step
            // are skipped (half of the work of those tiles; ~3% of a triangular product with 64-row tiles)
            const int k0 = kb + kt * C::BK;
            const bool edge = kt < n_first && ((p.tri == BO_TRI_A_UPPER && k0 < m0 + C::BM - p.tri_off) ||
                                               (p.tri == BO_TRI_A_LOWER && k0 + C::BK > m0 + p.tri_off + 8));
            if (!edge) {
#pragma unroll
                for (int kk = 0; kk < C::BK; kk += 4) {
                    double a[TM], b[TN];
#pragma unroll
                    for (int i = 0; i < TM; i++) a[i] = as[(kk / 4) * C::BM * 4 + i * 32];
#pragma unroll
                    for (int j = 0; j < TN; j++) b[j] = bs[(j * C::BK + kk) * 8];
#pragma unroll
                    for (int i = 0; i < TM; i++)
#pragma unroll
                        for (int j = 0; j < TN; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                }
            } else {
#pragma unroll
                for (int kk = 0; kk < C::BK; kk += 4) {
                    double b[TN];
#pragma unroll
                    for (int j = 0; j < TN; j++) b[j] = bs[(j * C::BK + kk) * 8];
#pragma unroll
                    for (int i = 0; i < TM; i++) {
                        const int r0 = m0 + wm0 + i * 8;
                        const bool zero = (p.tri == BO_TRI_A_UPPER) ? (k0 + kk + 4 <= r0 - p.tri_off)
                                                                    : (k0 + kk > r0 + 7 + p.tri_off);
                        if (zero) continue;
                        const double av = as[(kk / 4) * C::BM * 4 + i * 32];
#pragma unroll
                        for (int j = 0; j < TN; j++) dmma884(acc[i][j][0], acc[i][j][1], av, b[j]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bars_s + 8u * (C::STAGES + s));
        }
    }

    if constexpr (MODE == 1) {
        // Gram C_g^T C_g (8 x 8) of every 8-column group over this tile's rows >= gram_row0, on the tensor path: for a k4
        // step over rows r..r+3 of an 8-row block both MMA operands are C[r + t][column g], which lives in lane
        // (r % 8 + t, g / 2), slot g % 2 of the accumulator block -- 2 shuffles and a select per step.  Per-warp partials
        // go through shared memory (the ring is idle by now) and are added over the warps stacked in M in a fixed order.
        double* gpart = tiles;
        __syncthreads();  // every warp is done reading the ring
        if (warp < C::CONSUMERS) {
#pragma unroll
            for (int j = 0; j < TN; j++) {
                double g0 = 0.0, g1 = 0.0;
#pragma unroll
                for (int i = 0; i < TM; i++) {
                    const int row = m0 + wm0 + i * 8 + g;
                    const bool valid = row >= p.gram_row0 && row < p.M;
                    const double c0m = valid ? acc[i][j][0] : 0.0, c1m = valid ? acc[i][j][1] : 0.0;
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        const int src = ((hh * 4 + t) << 2) | (g >> 1);
                        const double y0 = __shfl_sync(0xffffffffu, c0m, src);
                        const double y1 = __shfl_sync(0xffffffffu, c1m, src);
                        const double av = (g & 1) ? y1 : y0;
                        dmma884(g0, g1, av, av);
                    }
                }
                *reinterpret_cast<double2*>(gpart + (warp * TN + j) * 64 + g * 8 + 2 * t) = make_double2(g0, g1);
            }
        }
        __syncthreads();
        double* Gz = p.gram + (long)z * p.strideGram + (long)mt * ((p.N + 7) / 8) * 64;
        for (int e = tid; e < (C::BN / 8) * 64; e += C::THREADS) {
            const int jg = e / 64, pr = e % 64;
            if (n0 + jg * 8 >= p.N) continue;
            const int wn = jg / TN, j = jg % TN;
            double s_ = 0.0;
#pragma unroll
            for (int wmi = 0; wmi < C::WARPS_M; wmi++) s_ += gpart[((wmi * C::WARPS_N + wn) * TN + j) * 64 + pr];
            Gz[(long)(n0 / 8 + jg) * 64 + pr] = s_;
        }
        __syncthreads();  // the covariance epilogue below may stage through the same memory
    }

    if constexpr (MODE == 3) {
        // Fused gradient contraction: G = epilogue(acc) replaces the accumulators in place (LINEAR epilogue with the
        // elementwise factor E), then Hpart[f][col] = sum_{r in tile} hA[f][m0 + r] G[r][col] on the DMMA pipe: the A
        // operand comes from global memory (a (d+1) x 64 strip of hA), the B operand is the accumulator block re-laid by two
        // shuffles + a select per k4 step, as in the gram mode.  Per-warp partials meet in shared memory (the ring is idle)
        // and the warps stacked in M are added in a fixed order.
        double* hs = tiles;  // [CONSUMERS][3][TN][64]
        if (warp < C::CONSUMERS) {
#pragma unroll
            for (int i = 0; i < TM; i++) {
                const int row = m0 + wm0 + i * 8 + g;
#pragma unroll
                for (int j = 0; j < TN; j++) {
                    const int col = n0 + wn0 + j * 8 + 2 * t;
                    const bool rv = row < p.M;
                    acc[i][j][0] = (rv && col < p.N) ? epilogue_value(p, z, acc[i][j][0], row, col) : 0.0;
                    acc[i][j][1] = (rv && col + 1 < p.N) ? epilogue_value(p, z, acc[i][j][1], row, col + 1) : 0.0;
                }
            }
        }
        __syncthreads();  // every warp is done reading the ring
        // the (d+1) x 64 strip of hA of this row tile, staged once (pitch 68: conflict-free fragment reads)
        constexpr int HP = C::BM + 4;
        double* hstrip = tiles + C::CONSUMERS * 3 * TN * 64;
        for (int e = tid; e < BO_HFUSE_ROWS * C::BM; e += C::THREADS) {
            const int f = e / C::BM, r = e % C::BM;
            hstrip[f * HP + r] = (f < p.hrows && m0 + r < p.M) ? p.hA[(long)f * p.ldhA + m0 + r] : 0.0;
        }
        __syncthreads();
        if (warp < C::CONSUMERS) {
#pragma unroll
            for (int j = 0; j < TN; j++) {
                double h[3][2];
#pragma unroll
                for (int fb = 0; fb < 3; fb++) h[fb][0] = h[fb][1] = 0.0;
#pragma unroll
                for (int i = 0; i < TM; i++) {
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        const int src = ((hh * 4 + t) << 2) | (g >> 1);
                        const double y0 = __shfl_sync(0xffffffffu, acc[i][j][0], src);
                        const double y1 = __shfl_sync(0xffffffffu, acc[i][j][1], src);
                        const double bv = (g & 1) ? y1 : y0;  // G[row i*8 + hh*4 + t][col j*8 + g]
                        const int ar = wm0 + i * 8 + hh * 4 + t;
#pragma unroll
                        for (int fb = 0; fb < 3; fb++) dmma884(h[fb][0], h[fb][1], hstrip[(fb * 8 + g) * HP + ar], bv);
                    }
                }
#pragma unroll
                for (int fb = 0; fb < 3; fb++)
                    *reinterpret_cast<double2*>(hs + ((warp * 3 + fb) * TN + j) * 64 + g * 8 + 2 * t) =
                        make_double2(h[fb][0], h[fb][1]);
            }
        }
        __syncthreads();
        double* Hz = p.hpart + (long)z * p.strideHpart + (long)mt * BO_HFUSE_ROWS * p.ldc;
        for (int e = tid; e < BO_HFUSE_ROWS * C::BN; e += C::THREADS) {
            const int f = e / C::BN, c = e % C::BN;
            if (n0 + c >= p.N) continue;
            const int fb = f / 8, fg = f % 8, wn = (c / 8) / TN, j = (c / 8) % TN, cc = c % 8;
            double s_ = 0.0;
#pragma unroll
            for (int wmi = 0; wmi < C::WARPS_M; wmi++)
                s_ += hs[(((wmi * C::WARPS_N + wn) * 3 + fb) * TN + j) * 64 + fg * 8 + cc];
            Hz[(long)f * p.ldc + n0 + c] = s_;
        }
        if (p.C == nullptr) return;
        // G is still wanted: the accumulators already hold the epilogue values
        double* __restrict__ Gz = p.C + (long)z * p.strideC;
        if (warp < C::CONSUMERS) {
#pragma unroll
            for (int i = 0; i < TM; i++) {
                const int row = m0 + wm0 + i * 8 + g;
                if (row >= p.M) continue;
#pragma unroll
                for (int j = 0; j < TN; j++) {
                    const int col = n0 + wn0 + j * 8 + 2 * t;
                    if (col >= p.N) continue;
                    if (col + 1 < p.N) *reinterpret_cast<double2*>(Gz + (long)row * p.ldc + col) = make_double2(acc[i][j][0], acc[i][j][1]);
                    else Gz[(long)row * p.ldc + col] = acc[i][j][0];
                }
            }
        }
        return;
    }

    double* __restrict__ Cz = p.C + (long)z * p.strideC;
    double* __restrict__ CTz = p.CT ? p.CT + (long)z * p.strideCT : nullptr;
    constexpr bool kStageFits = (size_t)C::BM * (C::BN + 2) * sizeof(double) <= (size_t)C::STAGES * C::STAGE_BYTES;
    if constexpr (kStageFits) if (p.epi != BO_EPI_LINEAR) {
        // covariance epilogues: staged through shared memory, one rolled copy of the transcendental code (see above).
        // Every TMA write has landed (each consumer waited on every `full` phase), so the ring can be reused.
        constexpr int LDC_S = C::BN + 2;
        __syncthreads();
        double* Cs = tiles;
        if (warp < C::CONSUMERS) {
#pragma unroll
            for (int i = 0; i < TM; i++)
#pragma unroll
                for (int j = 0; j < TN; j++)
                    *reinterpret_cast<double2*>(Cs + (wm0 + i * 8 + g) * LDC_S + wn0 + j * 8 + 2 * t) =
                        make_double2(acc[i][j][0], acc[i][j][1]);
        }
        __syncthreads();
#pragma unroll 2
        for (int e = tid; e < C::BM * C::BN; e += C::THREADS) {
            const int r = e / C::BN, c = e % C::BN;
            const int row = m0 + r, col = n0 + c;
            if (row >= p.M || col >= p.N) continue;
            const double v = epilogue_value(p, z, Cs[r * LDC_S + c], row, col);
            Cz[(long)row * p.ldc + col] = v;
            if (CTz) CTz[(long)col * p.ldct + row] = v;
        }
        return;
    }
    if (warp >= C::CONSUMERS) return;
#pragma unroll
    for (int i = 0; i < TM; i++) {
        const int row = m0 + wm0 + i * 8 + g;
        if (row >= p.M) continue;
#pragma unroll
        for (int j = 0; j < TN; j++) {
            const int col = n0 + wn0 + j * 8 + 2 * t;
            if (col >= p.N) continue;
            const double v0 = epilogue_value(p, z, acc[i][j][0], row, col);
            if (col + 1 < p.N) {
                const double v1 = epilogue_value(p, z, acc[i][j][1], row, col + 1);
                *reinterpret_cast<double2*>(Cz + (long)row * p.ldc + col) = make_double2(v0, v1);
                if (CTz) {
                    CTz[(long)col * p.ldct + row] = v0;
                    CTz[(long)(col + 1) * p.ldct + row] = v1;
                }
            } else {
                Cz[(long)row * p.ldc + col] = v0;
                if (CTz) CTz[(long)col * p.ldct + row] = v0;
            }
        }
    }
}

using CfgBig = Cfg<128, 128, 16, 64, 32, 4>;
using CfgMid = Cfg<64, 64, 16, 32, 32, 4>;
using CfgSmall = Cfg<32, 32, 16, 16, 16, 3>;
// tuning candidates (selected with BO_GEMM_CFG=n for tools/gemm_bench.py)
using CfgT2 = Cfg<64, 64, 16, 32, 32, 3>;     // 4 warps, 4 CTAs/SM
using CfgT3 = Cfg<128, 64, 16, 32, 32, 3>;    // 8 warps of 32x32, 2 CTAs/SM
using CfgT4 = Cfg<64, 128, 16, 32, 32, 3>;    // 8 warps of 32x32, 2 CTAs/SM
using CfgT5 = Cfg<128, 128, 32, 64, 32, 3>;   // Big with BK=32: half the barriers
using CfgT6 = Cfg<128, 128, 16, 32, 32, 3>;   // 16 warps of 32x32 in one CTA
using CfgT7 = Cfg<64, 128, 16, 32, 64, 3>;    // 4 warps of 32x64
using CfgT8 = Cfg<64, 64, 32, 32, 32, 2>;     // BK=32, double buffered, 3 CTAs/SM
using CfgT9 = Cfg<64, 64, 32, 32, 32, 3>;     // BK=32, 3 stages, 2 CTAs/SM
using CfgT10 = Cfg<64, 64, 16, 32, 32, 2>;    // double buffered, 4 CTAs/SM (16 warps)
using CfgT11 = Cfg<64, 32, 16, 32, 32, 3>;    // 2 warps per CTA, up to 8 CTAs/SM
// latency-bound panel products of the blocked Cholesky (M x 64 x K <= 128, a few dozen 64-row tiles at most): a lone CTA
// per SM is bound by its own SM's DMMA issue rate and by one global round trip per k-tile (graph-replayed 64 x 64 tiles:
// 10.8 us per launch even for a single tile).  16-row tiles spread the panel over four times as many SMs and BK = 32 with
// three stages keeps most of the k-range in flight; BN = 64 keeps the in-place form legal (see BoGemm::tile64).
using CfgPanel = Cfg<16, 64, 32, 16, 16, 3>;
using CfgSkinny = Cfg<32, 128, 16, 32, 32, 2>; // M <= 32 (the (d+1)-row gradient contraction): one row tile, wide columns

// multiply-adds per output column: nonzero entries of A (algorithmic count; triangular operands count their nonzeros only)
double gemm_macs(const BoGemm& g) {
    double macs = (double)g.M * g.K;
    if (g.tri == BO_TRI_A_LOWER) {
        macs = 0.0;
        for (int m = 0; m < g.M; m++) macs += (m + 1 + g.tri_off < g.K) ? m + 1 + g.tri_off : g.K;
    } else if (g.tri == BO_TRI_A_UPPER) {
        macs = 0.0;
        for (int m = 0; m < g.M; m++) macs += (m < g.tri_off) ? g.K : ((g.K - (m - g.tri_off) > 0) ? g.K - (m - g.tri_off) : 0);
    } else if (g.tri == BO_TRI_B_LOWER || g.out_lower) {
        macs *= 0.5;
    }
    return macs;
}

template <class C>
int launch_cfg(const BoGemm& g, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        BO_CUDA_TRY(cudaFuncSetAttribute(dgemm_dmma_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)C::SMEM));
        attr_set = true;
    }
    dim3 grid((g.N + C::BN - 1) / C::BN, (g.M + C::BM - 1) / C::BM, g.batch);
    if (grid.y > 65535u || grid.z > 65535u) return BO_ERR_UNSUPPORTED;
    ProfRec* rec = nullptr;
    if (g_prof_on == 1) {
        if (g_prof_used == g_prof_recs.size()) {
            ProfRec r;
            BO_CUDA_TRY(cudaEventCreate(&r.a));
            BO_CUDA_TRY(cudaEventCreate(&r.b));
            g_prof_recs.push_back(r);
        }
        rec = &g_prof_recs[g_prof_used++];
        g_prof_flops += 2.0 * gemm_macs(g) * g.N * g.batch;
        BO_CUDA_TRY(cudaEventRecord(rec->a, stream));
    }
    dgemm_dmma_kernel<C><<<grid, C::THREADS, C::SMEM, stream>>>(g);
    if (rec) BO_CUDA_TRY(cudaEventRecord(rec->b, stream));
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// Persistent form of the TMA-staged kernel (LINEAR epilogue): a 1-D grid of (SMs x MINB) CTAs, each walking a fixed
// list of output tiles.  The producer warp streams the k-tiles of one tile after another through the same ring, so
// the first k-tiles of the NEXT tile land while the DMMA warps run the epilogue of the current one: no pipeline fill
// per tile, which is where the triangular products (many short tiles) lost ~10% against the dense case.
// Tiles are ordered heaviest first (row tile major) and dealt to the CTAs in alternating direction wave by wave
// (boustrophedon), a static longest-processing-time schedule: no atomics, identical work list on both warp roles.
// ------------------------------------------------------------------------------------------------------------------
struct TileGeo {
    int z, mt, m0, n0, kb, ke, n_first, ntiles;
    bool skip;
};

template <class C, int MODE>
__device__ __forceinline__ TileGeo tile_geometry(const BoGemm& p, long L, int gx, int gy, int gz) {
    TileGeo q;
    const int x = (int)(L % gx);
    const int zz = (int)((L / gx) % gz);
    const int y = (int)(L / ((long)gx * gz));
    q.z = zz;
    q.mt = (p.tri == BO_TRI_A_LOWER) ? (gy - 1 - y) : y;  // heavy (long k-range) row tiles first
    q.m0 = q.mt * C::BM;
    q.n0 = x * C::BN;
    q.skip = p.out_lower && q.n0 > q.m0 + C::BM - 1;
    q.kb = 0;
    q.ke = p.K;
    if (p.tri == BO_TRI_A_LOWER) q.ke = min(p.K, q.m0 + C::BM + p.tri_off);
    else if (p.tri == BO_TRI_A_UPPER) q.kb = q.m0 > p.tri_off ? ((q.m0 - p.tri_off) / C::BK) * C::BK : 0;
    else if (p.tri == BO_TRI_B_LOWER) q.kb = (q.n0 / C::BK) * C::BK;
    if (MODE == 2 || MODE == 3) q.kb = p.post_k0;
    q.n_first = q.ke > q.kb ? (q.ke - q.kb + C::BK - 1) / C::BK : 0;
    q.ntiles = q.n_first + ((MODE == 2 || MODE == 3) ? (p.post_k0 + C::BK - 1) / C::BK : 0);
    return q;
}

__device__ __forceinline__ void consumer_bar(int threads) {
    asm volatile("bar.sync 1, %0;\n" ::"r"(threads) : "memory");
}

template <class C, int MODE>
__global__ void __launch_bounds__(C::THREADS, C::MINB)
dgemm_dmma_tma_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                 const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                                 const BoGemm p, const int za_step, const int zb_step, const int gx, const int gy,
                                 const int gz) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const unsigned raw = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (128u - (raw & 127u)) & 127u;
    double* tiles = reinterpret_cast<double*>(smem_raw + pad);
    const unsigned tiles_s = raw + pad;
    const unsigned bars_s = tiles_s + (unsigned)C::STAGES * C::STAGE_BYTES;  // full[STAGES], empty[STAGES]
    // gram partials of the fused forward: their own region (the ring keeps filling during the epilogue)
    double* gpart = reinterpret_cast<double*>(smem_raw + pad + (size_t)C::STAGES * C::STAGE_BYTES +
                                              2 * C::STAGES * sizeof(unsigned long long));

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const long total = (long)gx * gy * gz;
    const long G = gridDim.x;
    const long c = blockIdx.x;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; s++) {
            mbar_init(bars_s + 8u * s, 1);
            mbar_init(bars_s + 8u * (C::STAGES + s), C::CONSUMERS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();

    constexpr int TM = C::WM / 8, TN = C::WN / 8;
    const int wm0 = (warp / C::WARPS_N) * C::WM;
    const int wn0 = (warp % C::WARPS_N) * C::WN;

    if (warp == C::CONSUMERS) {
        // ---- producer ----
        if (lane == 0) {
            long it = 0;  // k-tiles issued so far by this CTA (ring position)
            for (long w = 0; w * G < total; w++) {
                const long L = w * G + ((w & 1) ? (G - 1 - c) : c);
                if (L >= total) continue;
                const TileGeo q = tile_geometry<C, MODE>(p, L, gx, gy, gz);
                if (q.skip) continue;
                for (int kt = 0; kt < q.ntiles; kt++, it++) {
                    const int s = (int)(it % C::STAGES);
                    if (it >= C::STAGES) mbar_wait(bars_s + 8u * (C::STAGES + s), (unsigned)(((it / C::STAGES) - 1) & 1));
                    const unsigned full = bars_s + 8u * s;
                    mbar_expect_tx(full, C::STAGE_BYTES);
                    const unsigned a_dst = tiles_s + (unsigned)s * C::STAGE_BYTES;
                    const unsigned b_dst = a_dst + (unsigned)(C::A_STAGE * sizeof(double));
                    if (kt < q.n_first) {
                        const int k0 = q.kb + kt * C::BK;
                        tma_load_4d(a_dst, &tmA, full, 0, q.m0, k0 / 4, q.z * za_step);
                        tma_load_4d(b_dst, &tmB, full, 0, k0, q.n0 / 8, q.z * zb_step);
                    } else {
                        const int k0 = (kt - q.n_first) * C::BK;
                        tma_load_4d(a_dst, &tmA2, full, 0, q.m0, k0 / 4, q.z * za_step);
                        tma_load_4d(b_dst, &tmB2, full, 0, k0, q.n0 / 8, q.z * zb_step);
                    }
                }
            }
        }
        return;
    }

    // ---- consumers ----
    long it = 0;
    for (long w = 0; w * G < total; w++) {
        const long L = w * G + ((w & 1) ? (G - 1 - c) : c);
        if (L >= total) continue;
        const TileGeo q = tile_geometry<C, MODE>(p, L, gx, gy, gz);
        if (q.skip) continue;
        const int z = q.z, m0 = q.m0, n0 = q.n0;
        double acc[TM][TN][2];
#pragma unroll
        for (int i = 0; i < TM; i++)
#pragma unroll
            for (int j = 0; j < TN; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
        for (int kt = 0; kt < q.ntiles; kt++, it++) {
            if ((MODE == 2 || MODE == 3) && kt == q.n_first) {
                const double* P = p.post + (long)z * p.stridePost;
                const int src_lo = (lane & ~3) | (t >> 1), src_hi = src_lo | 2;
#pragma unroll
                for (int j = 0; j < TN; j++) {
                    const int col0 = n0 + wn0 + j * 8;
                    if (col0 >= p.N) continue;
                    const double* Pg = P + (long)(col0 / 8) * 64;
                    const double pb0 = Pg[t * 8 + g], pb1 = Pg[(4 + t) * 8 + g];
#pragma unroll
                    for (int i = 0; i < TM; i++) {
                        const double x00 = __shfl_sync(0xffffffffu, acc[i][j][0], src_lo);
                        const double x01 = __shfl_sync(0xffffffffu, acc[i][j][1], src_lo);
                        const double x10 = __shfl_sync(0xffffffffu, acc[i][j][0], src_hi);
                        const double x11 = __shfl_sync(0xffffffffu, acc[i][j][1], src_hi);
                        double r0 = 0.0, r1 = 0.0;
                        dmma884(r0, r1, (t & 1) ? x01 : x00, pb0);
                        dmma884(r0, r1, (t & 1) ? x11 : x10, pb1);
                        acc[i][j][0] = r0;
                        acc[i][j][1] = r1;
                    }
                }
            }
            const int s = (int)(it % C::STAGES);
            mbar_wait(bars_s + 8u * s, (unsigned)((it / C::STAGES) & 1));
            const double* as = tiles + (size_t)s * (C::A_STAGE + C::B_STAGE) + (wm0 + g) * 4 + t;
            const double* bs = tiles + (size_t)s * (C::A_STAGE + C::B_STAGE) + C::A_STAGE + ((wn0 / 8) * C::BK + t) * 8 + g;
            const int k0 = q.kb + kt * C::BK;
            const bool edge = kt < q.n_first && ((p.tri == BO_TRI_A_UPPER && k0 < m0 + C::BM - p.tri_off) ||
                                                 (p.tri == BO_TRI_A_LOWER && k0 + C::BK > m0 + p.tri_off + 8));
            if (!edge) {
#pragma unroll
                for (int kk = 0; kk < C::BK; kk += 4) {
                    double a[TM], b[TN];
#pragma unroll
                    for (int i = 0; i < TM; i++) a[i] = as[(kk / 4) * C::BM * 4 + i * 32];
#pragma unroll
                    for (int j = 0; j < TN; j++) b[j] = bs[(j * C::BK + kk) * 8];
#pragma unroll
                    for (int i = 0; i < TM; i++)
#pragma unroll
                        for (int j = 0; j < TN; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                }
            } else {
#pragma unroll
                for (int kk = 0; kk < C::BK; kk += 4) {
                    double b[TN];
#pragma unroll
                    for (int j = 0; j < TN; j++) b[j] = bs[(j * C::BK + kk) * 8];
#pragma unroll
                    for (int i = 0; i < TM; i++) {
                        const int r0 = m0 + wm0 + i * 8;
                        const bool zero = (p.tri == BO_TRI_A_UPPER) ? (k0 + kk + 4 <= r0 - p.tri_off)
                                                                    : (k0 + kk > r0 + 7 + p.tri_off);
                        if (zero) continue;
                        const double av = as[(kk / 4) * C::BM * 4 + i * 32];
#pragma unroll
                        for (int j = 0; j < TN; j++) dmma884(acc[i][j][0], acc[i][j][1], av, b[j]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bars_s + 8u * (C::STAGES + s));
        }

        if constexpr (MODE == 1) {
#pragma unroll
            for (int j = 0; j < TN; j++) {
                double g0 = 0.0, g1 = 0.0;
#pragma unroll
                for (int i = 0; i < TM; i++) {
                    const int row = m0 + wm0 + i * 8 + g;
                    const bool valid = row >= p.gram_row0 && row < p.M;
                    const double c0m = valid ? acc[i][j][0] : 0.0, c1m = valid ? acc[i][j][1] : 0.0;
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) {
                        const int src = ((hh * 4 + t) << 2) | (g >> 1);
                        const double y0 = __shfl_sync(0xffffffffu, c0m, src);
                        const double y1 = __shfl_sync(0xffffffffu, c1m, src);
                        const double av = (g & 1) ? y1 : y0;
                        dmma884(g0, g1, av, av);
                    }
                }
                *reinterpret_cast<double2*>(gpart + (warp * TN + j) * 64 + g * 8 + 2 * t) = make_double2(g0, g1);
            }
            consumer_bar(C::CONSUMERS * 32);
            double* Gz = p.gram + (long)z * p.strideGram + (long)q.mt * ((p.N + 7) / 8) * 64;
            for (int e = tid; e < (C::BN / 8) * 64; e += C::CONSUMERS * 32) {
                const int jg = e / 64, pr = e % 64;
                if (n0 + jg * 8 >= p.N) continue;
                const int wn = jg / TN, j = jg % TN;
                double s_ = 0.0;
#pragma unroll
                for (int wmi = 0; wmi < C::WARPS_M; wmi++) s_ += gpart[((wmi * C::WARPS_N + wn) * TN + j) * 64 + pr];
                Gz[(long)(n0 / 8 + jg) * 64 + pr] = s_;
            }
            consumer_bar(C::CONSUMERS * 32);  // the next tile's partials reuse gpart
        }

        double* __restrict__ Cz = p.C + (long)z * p.strideC;
        double* __restrict__ CTz = p.CT ? p.CT + (long)z * p.strideCT : nullptr;
#pragma unroll
        for (int i = 0; i < TM; i++) {
            const int row = m0 + wm0 + i * 8 + g;
            if (row >= p.M) continue;
#pragma unroll
            for (int j = 0; j < TN; j++) {
                const int col = n0 + wn0 + j * 8 + 2 * t;
                if (col >= p.N) continue;
                const double v0 = epilogue_value(p, z, acc[i][j][0], row, col);
                if (col + 1 < p.N) {
                    const double v1 = epilogue_value(p, z, acc[i][j][1], row, col + 1);
                    *reinterpret_cast<double2*>(Cz + (long)row * p.ldc + col) = make_double2(v0, v1);
                    if (CTz) {
                        CTz[(long)col * p.ldct + row] = v0;
                        CTz[(long)(col + 1) * p.ldct + row] = v1;
                    }
                } else {
                    Cz[(long)row * p.ldc + col] = v0;
                    if (CTz) CTz[(long)col * p.ldct + row] = v0;
                }
            }
        }
    }
}

// ---- host side of the TMA variant -------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return (EncodeTiledFn)f;
    }();
    return fn;
}

// rank-4 view {inner (4 or 8 doubles), rows, groups, batch} of a row-major matrix batch
bool make_map(CUtensorMap* m, const double* base, unsigned inner, long rows, long groups, long ld, int batch, long stride,
              unsigned box_rows, unsigned box_groups) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const bool bcast = (batch > 1 && stride == 0);
    cuuint64_t dims[4] = {inner, (cuuint64_t)rows, (cuuint64_t)groups, (cuuint64_t)(bcast ? 1 : batch)};
    cuuint64_t strides[3] = {(cuuint64_t)ld * 8, (cuuint64_t)inner * 8,
                             (cuuint64_t)((batch > 1 && !bcast) ? stride * 8 : (long)inner * 8)};
    cuuint32_t box[4] = {inner, box_rows, box_groups, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, const_cast<double*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// operands the rank-4 views can describe: K a multiple of 4 (an A k4 group never straddles the row end), B rows padded to
// whole 8-column groups, 16-byte aligned batch strides
bool tma_eligible(const BoGemm& g) {
    if (g.K % 4 != 0 || g.ldb < bo_round_up(g.N, 8) || g.lda < g.K) return false;
    if (g.batch > 1 && ((g.strideA & 1) || (g.strideB & 1))) return false;
    if ((long)g.lda * 8 >= (1L << 40) || (long)g.ldb * 8 >= (1L << 40)) return false;
    return encode_tiled_fn() != nullptr;
}

template <class C, int MODE>
int launch_tma_mode(const BoGemm& g, cudaStream_t stream);

template <class C>
int launch_tma(const BoGemm& g, cudaStream_t stream) {
    if (g.gram != nullptr && g.post != nullptr) return BO_ERR_INVALID;
    if (g.gram != nullptr) return launch_tma_mode<C, 1>(g, stream);
    if (g.post != nullptr && g.hpart != nullptr) return launch_tma_mode<C, 3>(g, stream);
    if (g.hpart != nullptr) return BO_ERR_INVALID;  // hfuse rides on the post mode
    if (g.post != nullptr) return launch_tma_mode<C, 2>(g, stream);
    return launch_tma_mode<C, 0>(g, stream);
}

int sm_count() {
    static int sms = [] {
        int dev = 0, n = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        return n;
    }();
    return sms;
}

template <class C, int MODE>
int launch_tma_mode(const BoGemm& g, cudaStream_t stream) {
    // LINEAR epilogues run the persistent form; the covariance epilogues stage through the ring and keep one tile per CTA
    // Measured (profiles/r02_ncu_gemm_persistent_vs_tiled.txt): the static persistent schedule LOSES 8-9% to one tile per
    // CTA (dense 4096^3: 4.25 vs 3.89 ms; C3 step 1.71 vs 1.52 ms): the block scheduler's dynamic dispatch keeps every SM
    // at 3 resident CTAs to the end (14.3 of 15 warps active), the static work lists leave 12.2.  Off unless
    // BO_GEMM_PERSISTENT=1.
    static const bool pers_on = [] { const char* e = getenv("BO_GEMM_PERSISTENT"); return e && atoi(e) != 0; }();
    const bool persistent = pers_on && g.epi == BO_EPI_LINEAR && MODE != 3;
    if (MODE == 3 && (g.hA == nullptr || g.hrows <= 0 || g.hrows > BO_HFUSE_ROWS || C::BM != BO_GRAM_BM ||
                      ((size_t)C::CONSUMERS * 3 * (C::WN / 8) * 64 + (size_t)BO_HFUSE_ROWS * (C::BM + 4)) * sizeof(double) >
                          (size_t)C::STAGES * C::STAGE_BYTES))
        return BO_ERR_INVALID;
    constexpr size_t SMEM_P = C::SMEM + (size_t)C::CONSUMERS * (C::WN / 8) * 64 * sizeof(double);
    static bool attr_set = false;
    if (!attr_set) {
        BO_CUDA_TRY(cudaFuncSetAttribute(dgemm_dmma_tma_kernel<C, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)C::SMEM));
        BO_CUDA_TRY(cudaFuncSetAttribute(dgemm_dmma_tma_persistent_kernel<C, MODE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_P));
        attr_set = true;
    }
    dim3 grid((g.N + C::BN - 1) / C::BN, (g.M + C::BM - 1) / C::BM, g.batch);
    if (grid.y > 65535u || grid.z > 65535u) return BO_ERR_UNSUPPORTED;
    CUtensorMap tmA, tmB, tmA2, tmB2;
    if (!make_map(&tmA, g.A, 4, g.M, g.K / 4, g.lda, g.batch, g.strideA, C::BM, C::BK / 4) ||
        !make_map(&tmB, g.B, 8, g.K, (g.N + 7) / 8, g.ldb, g.batch, g.strideB, C::BK, C::BN / 8))
        return BO_ERR_INVALID;
    tmA2 = tmA;
    tmB2 = tmB;
    if (g.post != nullptr) {  // the dense leading k-range [0, post_k0): same operands, k extent clipped at post_k0
        if (g.post_k0 <= 0 || g.post_k0 % 4 != 0 || g.post_k0 >= g.K || g.epi != BO_EPI_LINEAR) return BO_ERR_INVALID;
        if (!make_map(&tmA2, g.A, 4, g.M, g.post_k0 / 4, g.lda, g.batch, g.strideA, C::BM, C::BK / 4) ||
            !make_map(&tmB2, g.B, 8, g.post_k0, (g.N + 7) / 8, g.ldb, g.batch, g.strideB, C::BK, C::BN / 8))
            return BO_ERR_INVALID;
    }
    if (g.gram != nullptr && (C::BM != BO_GRAM_BM || g.epi != BO_EPI_LINEAR)) return BO_ERR_INVALID;
    ProfRec* rec = nullptr;
    if (g_prof_on) {
        if (g_prof_used == g_prof_recs.size()) {
            ProfRec r;
            BO_CUDA_TRY(cudaEventCreate(&r.a));
            BO_CUDA_TRY(cudaEventCreate(&r.b));
            g_prof_recs.push_back(r);
        }
        rec = &g_prof_recs[g_prof_used++];
        // algorithmic flops of the launch: the product itself plus, in hfuse mode, the (d+1)-row contraction it absorbs
        g_prof_flops += 2.0 * gemm_macs(g) * g.N * g.batch + (MODE == 3 ? 2.0 * g.hrows * (double)g.M * g.N * g.batch : 0.0);
        BO_CUDA_TRY(cudaEventRecord(rec->a, stream));
    }
    const int za = (g.batch > 1 && g.strideA == 0) ? 0 : 1, zb = (g.batch > 1 && g.strideB == 0) ? 0 : 1;
    if (persistent) {
        const long total = (long)grid.x * grid.y * grid.z;
        const long slots = (long)sm_count() * C::MINB;
        const unsigned ctas = (unsigned)(total < slots ? total : slots);
        dgemm_dmma_tma_persistent_kernel<C, MODE><<<ctas, C::THREADS, (MODE == 1 ? SMEM_P : C::SMEM), stream>>>(
            tmA, tmB, tmA2, tmB2, g, za, zb, (int)grid.x, (int)grid.y, (int)grid.z);
    } else {
        dgemm_dmma_tma_kernel<C, MODE><<<grid, C::THREADS, C::SMEM, stream>>>(tmA, tmB, tmA2, tmB2, g, za, zb);
    }
    if (rec) BO_CUDA_TRY(cudaEventRecord(rec->b, stream));
    g_bo_launch_count++;
    BO_CUDA_TRY(cudaGetLastError());
    return BO_OK;
}

using TCfgBig = TCfg<128, 128, 16, 64, 32, 4, 1>;    // 8 consumer warps of 64 x 32, 1 CTA/SM, 128 KB ring
using TCfgWide = TCfg<128, 64, 16, 32, 32, 4, 2>;    // 8 consumer warps of 32 x 32, 2 CTAs/SM
using TCfgMid = TCfg<64, 64, 16, 32, 32, 4, 3>;      // 4 consumer warps of 32 x 32, 3 CTAs/SM
using TCfgBig32 = TCfg<128, 128, 32, 64, 32, 3, 1>;  // BK = 32: half the barrier traffic, 192 KB ring
using TCfgTall = TCfg<256, 64, 16, 64, 32, 4, 1>;    // 8 consumer warps of 64 x 32, B tile re-read less often
using TCfgMid4 = TCfg<64, 64, 16, 32, 32, 3, 4>;     // 3 stages, 4 CTAs/SM
using TCfgMid32 = TCfg<64, 64, 32, 32, 32, 3, 2>;    // BK = 32
using TCfgRow = TCfg<64, 128, 16, 32, 32, 4, 2>;     // 8 consumer warps of 32 x 32 side by side
using TCfgCol = TCfg<128, 64, 16, 64, 32, 4, 2>;     // 4 consumer warps of 64 x 32
using TCfgMid6 = TCfg<64, 64, 16, 32, 32, 6, 2>;     // deeper ring

}  // namespace

static bool takes_tma_mid(const BoGemm& g) {
    static const int forced = [] { const char* e = getenv("BO_GEMM_CFG"); return e ? atoi(e) : -1; }();
    static const bool use_tma = [] { const char* e = getenv("BO_GEMM_TMA"); return !e || atoi(e) != 0; }();
    if (forced >= 0 || !use_tma || g.tile64) return false;
    if (g.M <= 32 && g.N >= 1024 && g.K >= 64 && g.epi == BO_EPI_LINEAR && g.tri == BO_TRI_NONE) return false;  // skinny
    const long tiles = (long)((g.M + 63) / 64) * ((g.N + 63) / 64) * g.batch;
    // K >= 256: below that the 4-stage ring never fills and the 2-stage cp.async form with 4 CTAs/SM is faster
    // (C4, K = 100, 6144 tiles: step 1.59 vs 1.67 ms)
    return tiles >= 120 && g.K >= 256 && tma_eligible(g);
}

int bo_gemm_fused_ok(const BoGemm& g) {
    if (g.M <= 0 || g.N <= 0 || g.batch <= 0) return 0;
    if ((g.lda & 1) || (g.ldb & 1) || (g.ldc & 1) || (g.strideA & 1) || (g.strideB & 1) || (g.strideC & 1)) return 0;
    return takes_tma_mid(g) ? 1 : 0;
}

int bo_gemm_launch(const BoGemm& g, cudaStream_t stream) {
    if (g.M <= 0 || g.N <= 0 || g.batch <= 0) return BO_OK;
    if ((g.lda & 1) || (g.ldb & 1) || (g.ldc & 1) || (g.D && (g.ldd & 1))) return BO_ERR_INVALID;
    if ((reinterpret_cast<uintptr_t>(g.A) & 15) || (reinterpret_cast<uintptr_t>(g.B) & 15) ||
        (reinterpret_cast<uintptr_t>(g.C) & 15))
        return BO_ERR_INVALID;
    if ((g.strideA & 1) || (g.strideB & 1) || (g.strideC & 1)) return BO_ERR_INVALID;
    static const int forced = [] { const char* e = getenv("BO_GEMM_CFG"); return e ? atoi(e) : -1; }();
    switch (forced) {
        case 0: return launch_cfg<CfgBig>(g, stream);
        case 1: return launch_cfg<CfgMid>(g, stream);
        case 2: return launch_cfg<CfgT2>(g, stream);
        case 3: return launch_cfg<CfgT3>(g, stream);
        case 4: return launch_cfg<CfgT4>(g, stream);
        case 5: return launch_cfg<CfgT5>(g, stream);
        case 6: return launch_cfg<CfgT6>(g, stream);
        case 7: return launch_cfg<CfgT7>(g, stream);
        case 8: return launch_cfg<CfgT8>(g, stream);
        case 9: return launch_cfg<CfgT9>(g, stream);
        case 10: return launch_cfg<CfgT10>(g, stream);
        case 11: return launch_cfg<CfgT11>(g, stream);
        case 12: return launch_cfg<CfgPanel>(g, stream);
        default: break;
    }
    if (forced >= 20 && forced <= 29 && tma_eligible(g)) {  // TMA-staged variants (tools/gemm_bench.py)
        switch (forced) {
            case 20: return launch_tma<TCfgBig>(g, stream);
            case 21: return launch_tma<TCfgWide>(g, stream);
            case 22: return launch_tma<TCfgMid>(g, stream);
            case 23: return launch_tma<TCfgBig32>(g, stream);
            case 24: return launch_tma<TCfgTall>(g, stream);
            case 25: return launch_tma<TCfgMid4>(g, stream);
            case 26: return launch_tma<TCfgMid32>(g, stream);
            case 27: return launch_tma<TCfgRow>(g, stream);
            case 28: return launch_tma<TCfgCol>(g, stream);
            default: return launch_tma<TCfgMid6>(g, stream);
        }
    }
    auto tiles = [&](int bm, int bn) { return (long)((g.M + bm - 1) / bm) * ((g.N + bn - 1) / bn) * g.batch; };
    // Pick the largest tile that still fills the 148 SMs; small problems trade tile efficiency for occupancy.
    // 64x64 tiles, 4 warps, double-buffered, 4 CTAs per SM: measured best on B200 (tools/gemm_bench.py) --
    // co-resident small CTAs keep the DMMA pipe fed through each other's barriers and shrink the last wave.
    if (g.tile64) return launch_cfg<CfgPanel>(g, stream);
    if (g.M <= 32 && g.N >= 1024 && g.K >= 64 && g.epi == BO_EPI_LINEAR && g.tri == BO_TRI_NONE)
        return launch_cfg<CfgSkinny>(g, stream);
    // TMA-staged 64 x 64 tiles (one producer warp + 4 DMMA warps, 3 CTAs/SM) wherever the tensor maps can describe the
    // operands: measured 36.4 TF dense against 34.8 for the cp.async form and 35.4 for cuBLAS (tools/gemm_bench.py,
    // profiles/r02_gemm_tma_sweep.txt).  BO_GEMM_TMA=0 switches back for A/B comparisons.
    if (takes_tma_mid(g)) {
        // 64 x 128 tiles (8 DMMA warps, 2 CTAs/SM) halve the per-tile fixed cost on wide problems; BO_GEMM_WIDE=0 for A/B
        static const bool wide = [] { const char* e = getenv("BO_GEMM_WIDE"); return e && atoi(e) != 0; }();
        if (wide && g.N >= 2048) return launch_tma<TCfgRow>(g, stream);
        return launch_tma<TCfgMid>(g, stream);
    }
    if (g.gram != nullptr || g.post != nullptr) return BO_ERR_INVALID;  // the caller did not ask bo_gemm_fused_ok
    if (tiles(64, 64) >= 120 || (g.M >= 48 && g.N >= 48 && tiles(32, 32) > 2048)) return launch_cfg<CfgT10>(g, stream);
    return launch_cfg<CfgSmall>(g, stream);
}
