// Unit harness: check_intervals + mask_build (gather build) against fill_multi + derive_multi (scatter)
// on random sorted intervals, on poisoned buffers.  Prints the first differing words.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I pymasc_b200/csrc -o tools/mask_build_check.bin tools/mask_build_check.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <random>
#include <algorithm>
#include "pmb_build.cuh"
#include "pmb_events.cuh"
using namespace pmb;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

int main(int argc, char** argv)
{
    const int reps = argc > 1 ? atoi(argv[1]) : 50;
    std::mt19937_64 rng(12345);
    int bad_cases = 0;
    for (int rep = 0; rep < reps; rep++) {
        const bool big = argc > 2 && rep % 2 == 0;      // a long chromosome with a short, dense stretch of intervals
        const uint32_t L = big ? 20000000 + rng() % 230000000 : 3000 + rng() % 400000;
        uint32_t rl = 2 + rng() % 120, D = 1 + rng() % 500;
        uint32_t Lfile = 0;
        if (argc > 3 && rep == 0) { rl = 36; D = 300; Lfile = 249250621; }
        const uint32_t halo = ((2 * D + rl + 192 + 31) / 32 + 3) & ~3u;
        const uint64_t nbits = (uint64_t)(Lfile ? Lfile : L) + rl + D + 100;
        const uint32_t body = (uint32_t)(((nbits + 31) / 32 + 63) / 64 * 64);
        const uint32_t W = 2 * halo + body;
        // sorted disjoint intervals
        std::vector<uint32_t> edges;
        const uint32_t n_iv = big ? 1000 : 1 + rng() % (L / 50);
        const uint32_t win0 = big ? rng() % (L - 200000) : 0, win = big ? 150000 : L;
        for (uint32_t i = 0; i < 2 * n_iv; i++) edges.push_back(win0 + rng() % win);
        std::sort(edges.begin(), edges.end());
        std::vector<uint32_t> hb, he;
        for (uint32_t i = 0; i + 1 < edges.size(); i += 2) { hb.push_back(edges[i]); he.push_back(edges[i + 1]); }
        if (argc > 3 && rep == 0) {       // intervals from a file: "n" then n lines "begin end"
            FILE* f = fopen(argv[3], "r");
            unsigned n = 0; if (!f || fscanf(f, "%u", &n) != 1) { printf("cannot read %s\n", argv[3]); return 2; }
            hb.resize(n); he.resize(n);
            for (unsigned i = 0; i < n; i++) if (fscanf(f, "%u %u", &hb[i], &he[i]) != 2) return 2;
            fclose(f);
        }
        const uint32_t nI = (uint32_t)hb.size();
        uint32_t *d_b, *d_e, *d_M[2], *d_D[2];
        unsigned long long *d_adir, *d_long64;
        uint32_t* d_long; int* d_flags;
        CK(cudaMalloc(&d_b, nI * 4 + 16)); CK(cudaMalloc(&d_e, nI * 4 + 16));
        CK(cudaMemcpy(d_b, hb.data(), nI * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_e, he.data(), nI * 4, cudaMemcpyHostToDevice));
        for (int v = 0; v < 2; v++) {
            CK(cudaMalloc(&d_M[v], (size_t)W * 4)); CK(cudaMalloc(&d_D[v], (size_t)W * 4));
            CK(cudaMemset(d_M[v], v == 0 ? 0 : 0xA5, (size_t)W * 4));      // scatter needs zeros; gather gets poison
            CK(cudaMemset(d_D[v], 0x5A, (size_t)W * 4));
        }
        CK(cudaMalloc(&d_adir, 64)); CK(cudaMemset(d_adir, 0, 64));
        CK(cudaMalloc(&d_long64, (nI + 1) * 8)); CK(cudaMalloc(&d_long, 16)); CK(cudaMemset(d_long, 0, 16));
        CK(cudaMalloc(&d_flags, 64)); CK(cudaMemset(d_flags, 0, 64));
        Geometry g{0, (int64_t)halo * 32, (int64_t)W * 32};
        FillTask ft{d_M[0], g, d_b, d_e, 0, -(1ll << 40), (1ll << 40), nbits};
        ft.c_lo = 0; ft.c_hi = (int64_t)nbits; ft.p_lo = -(int64_t)halo * 32; ft.p_hi = ft.p_lo + (int64_t)W * 32;
        ft.lag0 = rl - 1; ft.lag1 = (int64_t)rl - 2; ft.agather = d_adir + 2;
        uint64_t fp[2] = {0, nI};
        DeriveTask dt[2];
        for (int v = 0; v < 2; v++) {
            DeriveTask d{};
            d.M = d_M[v]; d.Er = d_D[v]; d.Or = d_D[v] + W / 2; d.adirect = d_adir + 2 * v;
            d.a_lo = (long long)halo * 32; d.a_hi = d.a_lo + (long long)nbits;
            d.lag0 = rl - 1; d.lag1 = (long long)rl - 2; d.array_words = W;
            d.n_intervals = nI; d.begins = d_b; d.ends = d_e; d.p_off = -(long long)halo * 32;
            dt[v] = d;
        }
        const uint64_t ctas = (W / 4 + DERIVE_QUADS_PER_CTA - 1) / DERIVE_QUADS_PER_CTA;
        uint64_t dp[2] = {0, ctas};
        FillTask* d_ft; uint64_t *d_fp, *d_dp; DeriveTask* d_dt;
        CK(cudaMalloc(&d_ft, sizeof ft)); CK(cudaMemcpy(d_ft, &ft, sizeof ft, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_fp, 16)); CK(cudaMemcpy(d_fp, fp, 16, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_dp, 16)); CK(cudaMemcpy(d_dp, dp, 16, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_dt, sizeof dt)); CK(cudaMemcpy(d_dt, dt, sizeof dt, cudaMemcpyHostToDevice));
        // scatter reference
        fill_multi_kernel<<<256, 256>>>(d_ft, d_fp, 1, d_flags, d_long64, d_long, nullptr);
        fill_long_multi_kernel<<<64, 256>>>(d_ft, d_long64, d_long, nullptr);
        derive_multi_kernel<<<(unsigned)ctas, 256>>>(d_dt, d_dp, 1, 2, nullptr);
        // gather
        check_intervals_multi_kernel<<<64, 256>>>(d_ft, d_fp, 1, 2, d_flags, d_flags + 1);
        const uint32_t K = 0;
        mask_build_multi_kernel<<<(unsigned)ctas, 256>>>(d_dt + 1, d_dp, 1, d_flags + 1);
        CK(cudaDeviceSynchronize());
        std::vector<uint32_t> M0(W), M1(W), D0(W), D1(W);
        unsigned long long adir[4]; int flags[2];
        CK(cudaMemcpy(M0.data(), d_M[0], (size_t)W * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(M1.data(), d_M[1], (size_t)W * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(D0.data(), d_D[0], (size_t)W * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(D1.data(), d_D[1], (size_t)W * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(adir, d_adir, 32, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(flags, d_flags, 8, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (uint32_t i = 0; i < W && bad < 5; i++)
            if (M0[i] != M1[i]) { printf("  M[%u] scatter %08x gather %08x\n", i, M0[i], M1[i]); bad++; }
        for (uint32_t i = 0; i < W && bad < 10; i++)
            if (D0[i] != D1[i]) { printf("  ErOr[%u] scatter %08x gather %08x\n", i, D0[i], D1[i]); bad++; }
        if (adir[0] != adir[2] || adir[1] != adir[3]) { printf("  adirect %llu %llu vs %llu %llu\n", adir[0], adir[1], adir[2], adir[3]); bad++; }
        if (flags[0] || flags[1]) { printf("  flags %d %d\n", flags[0], flags[1]); bad++; }
        if (bad) { bad_cases++; printf("case %d: L %u rl %u D %u W %u nI %u K %u\n", rep, L, rl, D, W, nI, K); }
        cudaFree(d_b); cudaFree(d_e); cudaFree(d_M[0]); cudaFree(d_M[1]); cudaFree(d_D[0]); cudaFree(d_D[1]);
        cudaFree(d_adir); cudaFree(d_long64); cudaFree(d_long); cudaFree(d_flags);
        cudaFree(d_ft); cudaFree(d_fp); cudaFree(d_dp); cudaFree(d_dt);
    }
    printf("%d / %d cases differ\n", bad_cases, reps);
    return bad_cases != 0;
}
