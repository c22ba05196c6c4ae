// Probe: what does a tiny kernel with a large dynamic shared-memory request cost when it alternates with kernels that use none?
// (nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o carveout_probe carveout_probe.cu; prints us per launch for several patterns)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_plain(double* x) { if (threadIdx.x == 0 && blockIdx.x == 0) x[0] += 1.0; }
__global__ void k_smem(double* x, int n) { extern __shared__ double s[]; if (threadIdx.x < n) s[threadIdx.x] = x[threadIdx.x]; __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0) x[0] = s[0] + 1.0; }
__global__ void k_smem2(double* x, int n) { extern __shared__ double s[]; if (threadIdx.x < n) s[threadIdx.x] = x[threadIdx.x]; __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0) x[0] = s[0] + 2.0; }
template <class F> float time_us(F f, int reps) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); for (int i = 0; i < reps; ++i) f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms * 1e3f / reps;
}
int main() {
    double* x; cudaMalloc(&x, 1 << 20); cudaMemset(x, 0, 1 << 20);
    const int big = 111 * 1024, mid = 48 * 1024;
    cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_smem2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1) {
            cudaFuncSetAttribute(k_plain, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
            std::printf("--- k_plain now prefers the maximum shared-memory carve-out ---\n");
        }
        std::printf("plain x1                 : %6.2f us per launch\n", time_us([&] { k_plain<<<4, 256>>>(x); }, 400));
        std::printf("smem 111K (1 CTA) x1     : %6.2f us per launch\n", time_us([&] { k_smem<<<1, 256, big>>>(x, 8); }, 400));
        std::printf("smem 111K (66 CTAs) x1   : %6.2f us per launch\n", time_us([&] { k_smem<<<66, 256, big>>>(x, 8); }, 400));
        std::printf("smem 48K (1 CTA) x1      : %6.2f us per launch\n", time_us([&] { k_smem<<<1, 256, mid>>>(x, 8); }, 400));
        std::printf("plain + smem 111K pair   : %6.2f us per pair\n", time_us([&] { k_plain<<<4, 256>>>(x); k_smem<<<1, 256, big>>>(x, 8); }, 400));
        std::printf("plain + smem 111K(66) pr : %6.2f us per pair\n", time_us([&] { k_plain<<<4, 256>>>(x); k_smem<<<66, 256, big>>>(x, 8); }, 400));
        std::printf("plain + smem 48K pair    : %6.2f us per pair\n", time_us([&] { k_plain<<<4, 256>>>(x); k_smem<<<1, 256, mid>>>(x, 8); }, 400));
        std::printf("smem 111K + smem2 220K pr: %6.2f us per pair\n", time_us([&] { k_smem<<<1, 256, big>>>(x, 8); k_smem2<<<1, 256, 200 * 1024>>>(x, 8); }, 400));
        std::printf("smem 111K, 1024 threads  : %6.2f us per launch\n", time_us([&] { k_smem<<<1, 1024, big>>>(x, 8); }, 400));
    }
    std::printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
