// FP64 roofline denominators for B200 (sm_100a): pure-DFMA and DMMA.8x8x4 issue-rate
// microbenchmarks. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peaks fp64_peaks.cu
// Not part of the product path; it only measures the peaks DESIGN.md / bench.py quote.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int ILP>
__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double s) {
    double acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x * 1e-3 + i;
    double a = s, b = 1.0 - s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) acc[i] = fma(acc[i], a, b);
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int ILP>
__global__ void __launch_bounds__(256) dmma_kernel(double* out, int iters, double s) {
    double c[ILP][2];
#pragma unroll
    for (int i = 0; i < ILP; i++) { c[i][0] = 0; c[i][1] = 0; }
    double a = s * (threadIdx.x & 31), b = 1.0 - s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <typename F>
float time_ms(F f, int reps) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); f(); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("device %s sms %d clock_khz %d\n", p.name, sms, p.clockRate);
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 256));
    const int iters = 20000;
    for (int occ = 1; occ <= 8; occ *= 2) {
        int grid = sms * occ;
        {
            constexpr int ILP = 8;
            float ms = time_ms([&] { dfma_kernel<ILP><<<grid, 256>>>(out, iters, 0.999); }, 5);
            double fl = 2.0 * ILP * iters * 256.0 * grid;
            printf("DFMA  ctas/sm %d ilp %d: %.3f ms  %.2f TFLOP/s\n", occ, ILP, ms, fl / ms * 1e-9);
        }
        {
            constexpr int ILP = 8;
            float ms = time_ms([&] { dmma_kernel<ILP><<<grid, 256>>>(out, iters, 0.999); }, 5);
            double fl = 2.0 * 256 * ILP * iters * 8.0 * grid;  // 256 FMA per warp-DMMA, 8 warps
            printf("DMMA  ctas/sm %d ilp %d: %.3f ms  %.2f TFLOP/s\n", occ, ILP, ms, fl / ms * 1e-9);
        }
    }
    // sustained 3 s DMMA
    {
        constexpr int ILP = 8; int grid = sms * 4;
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0));
        int launches = 0; float ms = 0;
        while (true) {
            for (int i = 0; i < 20; i++) dmma_kernel<ILP><<<grid, 256>>>(out, iters, 0.999);
            launches += 20;
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
            if (ms > 3000) break;
        }
        double fl = 2.0 * 256 * ILP * iters * 8.0 * grid * launches;
        printf("DMMA sustained %.0f ms: %.2f TFLOP/s\n", ms, fl / ms * 1e-9);
    }
    return 0;
}
