// pmb_probe.cuh -- integer-pipe calibration kernels.
//
// The roofline of the shift-window kernel is the POPC issue rate, which NVIDIA does not
// publish for sm_100 and MEASURED_PEAKS.json does not contain.  These loops measure it
// on the device: 16 independent dependency chains per thread, enough to saturate a
// pipe at 16 resident warps per SM sub-partition.  `what`:
//   0 POPC          1 LOP3          2 POPC + LOP3 interleaved (do the two pipes overlap?)
//   3 SHF (funnel)  4 IADD3         5 POPC + IMAD              6 POPC + 2 LOP3 + 1 SHF
// Reported rate counts the named instructions only (for mixes: all named ops).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace pmb {

template <int WHAT>
__global__ void __launch_bounds__(512) probe_kernel(int iters, uint32_t* sink)
{
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
    uint32_t s = threadIdx.x & 31, t = blockIdx.x | 1;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (WHAT == 0) {
                asm volatile("popc.b32 %0, %0;" : "+r"(v[i]));
            } else if (WHAT == 1) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(s), "r"(t));
            } else if (WHAT == 2) {
                uint32_t p;
                asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(v[i]));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(p), "r"(t));
            } else if (WHAT == 3) {
                asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(t), "r"(s));
            } else if (WHAT == 4) {
                asm volatile("add.u32 %0, %0, %1;" : "+r"(v[i]) : "r"(t));
            } else if (WHAT == 5) {
                uint32_t p;
                asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(v[i]));
                asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(v[i]) : "r"(p), "r"(t));
            } else {
                uint32_t p, q;
                asm volatile("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(q) : "r"(v[i]), "r"(t), "r"(s));
                asm volatile("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(p) : "r"(q), "r"(v[i]), "r"(t));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(q), "r"(p));
                asm volatile("popc.b32 %0, %1;" : "=r"(q) : "r"(p));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(v[i]) : "r"(q));
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) acc ^= v[i];
    if (acc == 0x12345678u) sink[threadIdx.x] = acc;   // keep the chains alive
}

// returns the number of counted ops per thread (0 for an unknown probe)
inline double launch_probe(int what, int blocks, int threads, int iters, uint32_t* sink, cudaStream_t st)
{
    switch (what) {
        case 0: probe_kernel<0><<<blocks, threads, 0, st>>>(iters, sink); return 16.0 * iters;
        case 1: probe_kernel<1><<<blocks, threads, 0, st>>>(iters, sink); return 16.0 * iters;
        case 2: probe_kernel<2><<<blocks, threads, 0, st>>>(iters, sink); return 32.0 * iters;
        case 3: probe_kernel<3><<<blocks, threads, 0, st>>>(iters, sink); return 16.0 * iters;
        case 4: probe_kernel<4><<<blocks, threads, 0, st>>>(iters, sink); return 16.0 * iters;
        case 5: probe_kernel<5><<<blocks, threads, 0, st>>>(iters, sink); return 32.0 * iters;
        case 6: probe_kernel<6><<<blocks, threads, 0, st>>>(iters, sink); return 80.0 * iters;
        default: return 0.0;
    }
}

}  // namespace pmb
