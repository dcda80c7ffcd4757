// atom_probe.cu -- how fast are histogram updates on this GPU?  (design input for the list path)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/atom_probe.bin tools/atom_probe.cu
// Every thread draws pseudo-random bins (LCG) and updates a histogram ITER times:
//   mode 0: shared-memory atomicAdd (ATOMS.ADD, no return)      bins per CTA = NB
//   mode 1: global red (RED.E.ADD) into NB bins x COPIES copies (copy = blockIdx % COPIES)
//   mode 2: shared atomicAdd, only every `sparse`-th lane active (is the cost per lane or per instruction?)
//   mode 3: plain LDS + IADD + STS (not atomic; the LSU floor)
//   mode 4: half the updates to shared, half to global (do the two units overlap?)
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(1024) probe(uint32_t* g, int nb, int copies, int iters, int sparse)
{
    extern __shared__ uint32_t h[];
    for (int i = threadIdx.x; i < nb; i += blockDim.x) h[i] = 0;
    __syncthreads();
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t* gc = g + (size_t)(blockIdx.x % copies) * nb;
    const bool on = (threadIdx.x % sparse) == 0;
    for (int it = 0; it < iters; ++it) {
        s = s * 1664525u + 1013904223u;
        const uint32_t b = (uint32_t)(((uint64_t)(s >> 8) * (uint32_t)nb) >> 24);
        if (MODE == 0) atomicAdd(&h[b], 1u);
        if (MODE == 1) atomicAdd(&gc[b], 1u);
        if (MODE == 2) { if (on) atomicAdd(&h[b], 1u); }
        if (MODE == 3) h[b] = h[b] + 1u;
        if (MODE == 4) { if (it & 1) atomicAdd(&h[b], 1u); else atomicAdd(&gc[b], 1u); }
    }
    __syncthreads();
    if (MODE != 1) for (int i = threadIdx.x; i < nb; i += blockDim.x) if (h[i]) atomicAdd(&gc[i], h[i]);
}

template <int MODE>
static void run(const char* name, uint32_t* g, int nb, int copies, int blocks, int threads, int sparse)
{
    const int iters = 2048;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaMemset(g, 0, (size_t)nb * copies * 4);
        cudaEventRecord(a);
        probe<MODE><<<blocks, threads, nb * 4>>>(g, nb, copies, iters, sparse);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep) best = ms < best ? ms : best;
    }
    const double active = (double)blocks * threads / sparse * iters;
    printf("{\"probe\": \"%s\", \"bins\": %d, \"copies\": %d, \"blocks\": %d, \"threads\": %d, \"sparse\": %d, \"ms\": %.4f, "
           "\"updates_per_s\": %.4g, \"updates_per_clk_per_sm\": %.3f}\n",
           name, nb, copies, blocks, threads, sparse, best, active / (best * 1e-3),
           active / (best * 1e-3) / 148.0 / 1.965e9);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
}

int main()
{
    uint32_t* g; cudaMalloc(&g, (size_t)8192 * 4096 * 4);
    for (int nb : {1001, 5001, 32, 1}) {
        run<0>("smem_atomic", g, nb, 1, 148 * 2, 1024, 1);
        run<0>("smem_atomic", g, nb, 1, 148 * 4, 256, 1);
    }
    run<2>("smem_atomic_sparse_lanes", g, 1001, 1, 148 * 2, 1024, 4);
    run<2>("smem_atomic_sparse_lanes", g, 1001, 1, 148 * 2, 1024, 32);
    run<3>("smem_plain_rmw", g, 1001, 1, 148 * 2, 1024, 1);
    for (int copies : {1, 8, 64, 1024}) run<1>("global_red", g, 1001, copies, 148 * 2, 1024, 1);
    run<1>("global_red", g, 5001, 1, 148 * 2, 1024, 1);
    run<4>("smem_plus_global", g, 1001, 8, 148 * 2, 1024, 1);
    run<4>("smem_plus_global", g, 1001, 64, 148 * 2, 1024, 1);
    return 0;
}
