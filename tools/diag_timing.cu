// Phase timing of potrf_diag_kernel (clock64 stamps of CTA 0): nvcc -DBO_DIAG_TIMING tools/diag_timing.cu botorch_b200/csrc/gemm.cu
#include "../botorch_b200/csrc/linalg.cu"
#include <vector>
void bo_set_last_cuda_error(cudaError_t, const char*, int) {}
int main() {
    const int n = 128, ld = 128;
    std::vector<double> h(n * ld, 0.0);
    for (int i = 0; i < n; i++)
        for (int j = 0; j <= i; j++) h[i * ld + j] = (i == j) ? 4.0 + 0.01 * i : 0.3 / (1.0 + i - j);
    double *A, *X, *B;
    int* info;
    cudaMalloc(&A, sizeof(double) * n * ld); cudaMalloc(&X, sizeof(double) * n * ld); cudaMalloc(&B, sizeof(double) * 2 * 64 * 64);
    cudaMalloc(&info, 4); cudaMemset(info, 0, 4);
    const size_t smem = sizeof(double) * 2 * 64 * 65;
    cudaFuncSetAttribute(potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const char* names[] = {"load A", "factorise", "inverse", "store L/inv/invT"};
    for (int rep = 0; rep < 3; rep++) {
        cudaMemcpy(A, h.data(), sizeof(double) * n * ld, cudaMemcpyHostToDevice);
                potrf_diag_kernel<<<1, 256, smem>>>(A + 64 * ld + 64, ld, 0, 64, 64, X + 64 * ld + 64, ld, 0, B, info);
        cudaDeviceSynchronize();
        long long st[16];
        cudaMemcpyFromSymbol(st, g_diag_stamps, sizeof(st));
        printf("rep %d:", rep);
        for (int i = 0; i < 4; i++) printf("  %s %lld", names[i], st[i + 1] - st[i]);
        printf("  total %lld cycles (%s)\n", st[4] - st[0], cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
