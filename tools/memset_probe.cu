// How fast is zeroing the bit arrays?  cudaMemsetAsync vs a plain 16-byte-store kernel.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/memset_probe tools/memset_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void zero16(uint4* p, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) p[i] = make_uint4(0, 0, 0, 0);
}
int main()
{
    const size_t bytes = 290ull << 20;
    void* d; cudaMalloc(&d, bytes);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float ms;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(a);
        for (int i = 0; i < 10; i++) cudaMemsetAsync(d, 0, bytes);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("cudaMemsetAsync  %.1f us  %.0f GB/s\n", ms * 100, bytes / (ms / 10 * 1e-3) / 1e9);
        for (int g = 148 * 4; g <= 148 * 32; g *= 2) {
            cudaEventRecord(a);
            for (int i = 0; i < 10; i++) zero16<<<g, 512>>>((uint4*)d, bytes / 16);
            cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
            printf("zero16 grid %5d  %.1f us  %.0f GB/s\n", g, ms * 100, bytes / (ms / 10 * 1e-3) / 1e9);
        }
    }
    return 0;
}
