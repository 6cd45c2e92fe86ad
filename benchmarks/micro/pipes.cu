// pipes.cu -- which issue pipes the integer instructions of the JSON fast lane use on sm_100a.
// Each kernel runs a long unrolled chain of independent instructions of one or two kinds per warp
// (8 independent accumulators, 16 warps per SM sub-partition) and reports warp-instructions per
// cycle per SM sub-partition.  Two kinds that share a pipe add up to the single-kind rate; two
// kinds on different pipes approach the sum of their rates (up to the issue limit of 1).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define REP 64
#define ACC 8

template <int MODE>
__global__ void __launch_bounds__(512) k(uint32_t* out, uint32_t seed, long long* cycles) {
    uint32_t a[ACC], b = seed | 1u, c = seed * 3u + 7u;
#pragma unroll
    for (int i = 0; i < ACC; ++i) a[i] = seed + i * 977u + threadIdx.x;
    const long long t0 = clock64();
    for (int it = 0; it < 64; ++it) {
#pragma unroll
        for (int r = 0; r < REP / ACC; ++r) {
#pragma unroll
            for (int i = 0; i < ACC; ++i) {
                if (MODE == 0 || MODE == 10 || MODE == 12 || MODE == 14 || MODE == 16 || MODE == 18 || MODE == 20)   // LOP3
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 1 || MODE == 10 || MODE == 11 || MODE == 13)                                            // IMAD
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 2 || MODE == 11 || MODE == 12)                                                          // IDP4A
                    asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 3 || MODE == 13 || MODE == 14)                                                          // IMAD.HI
                    asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 4 || MODE == 15 || MODE == 16)                                                          // SHF (funnel)
                    asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 5 || MODE == 17 || MODE == 18)                                                          // PRMT
                    asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 6 || MODE == 19 || MODE == 20) {                                                        // IMAD.WIDE
                    unsigned long long w;
                    asm volatile("mad.wide.u32 %0, %1, %2, %3;" : "=l"(w) : "r"(a[i]), "r"(b), "l"((unsigned long long)c));
                    a[i] = (uint32_t)(w >> 32) ^ (uint32_t)w;                                                       // (one LOP3 with it)
                }
                if (MODE == 7)                                                                                      // IADD3
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
                if (MODE == 15 || MODE == 17 || MODE == 19)                                                         // + IMAD
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            }
        }
    }
    const long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < ACC; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE>
static void run(const char* name, int per_iter) {
    uint32_t* out; long long* cyc; long long h = 0;
    cudaMalloc(&out, 148 * 2048 * 4); cudaMalloc(&cyc, 8);
    k<MODE><<<148 * 4, 512>>>(out, 12345u, cyc);          // 4 blocks of 16 warps per SM = 16 warps per sub-partition
    k<MODE><<<148 * 4, 512>>>(out, 12345u, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double inst = 64.0 * REP * per_iter * 16;       // warp-instructions per sub-partition
    printf("%-28s %.3f warp-inst/cycle/SMSP (%s)\n", name, inst / (double)h, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("LOP3", 1); run<1>("IMAD", 1); run<2>("IDP4A", 1); run<3>("IMAD.HI", 1); run<4>("SHF", 1);
    run<5>("PRMT", 1); run<6>("IMAD.WIDE + LOP3", 2); run<7>("IADD", 1);
    run<10>("LOP3 + IMAD", 2); run<11>("IMAD + IDP4A", 2); run<12>("LOP3 + IDP4A", 2);
    run<13>("IMAD + IMAD.HI", 2); run<14>("LOP3 + IMAD.HI", 2); run<15>("SHF + IMAD", 2); run<16>("LOP3 + SHF", 2);
    run<17>("PRMT + IMAD", 2); run<18>("LOP3 + PRMT", 2); run<19>("IMAD.WIDE + LOP3 + IMAD", 3); run<20>("LOP3 + IMAD.WIDE + LOP3", 3);
    return 0;
}
