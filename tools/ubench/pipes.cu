// Developer micro-benchmark: issue rates of the integer instructions the streaming kernels
// lean on (per SM, warp-instructions per cycle), alone and mixed.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(unsigned* out, unsigned seed, int iters) {
    unsigned a[8], b = seed + threadIdx.x, c = seed * 3u + 1u;
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * (i + 1) + seed;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = a[i] + b - c;                        // IADD3
            if (MODE == 1) a[i] = a[i] * b + c;                        // IMAD
            if (MODE == 2) a[i] = __dp2a_lo(b, 0x0101u, a[i]);         // IDP.2A
            if (MODE == 3) a[i] = (i & 1) ? a[i] + b - c : __dp2a_lo(b, 0x0101u, a[i]);  // mix
            if (MODE == 4) a[i] = min(a[i], b) ^ c;                    // VIMNMX + LOP3
            if (MODE == 5) a[i] = (i & 1) ? a[i] * b + c : a[i] + b - c;  // IMAD + IADD3
            if (MODE == 6) a[i] = __vminu2(a[i], b) + c;               // VIMNMX.U16x2 + IADD
            if (MODE == 7) a[i] = __funnelshift_r(a[i], b, 16);        // SHF
            if (MODE == 8) a[i] = __byte_perm(a[i], b, 0x5410);        // PRMT
            if (MODE == 9) { float f = __uint_as_float(a[i]); asm("sqrt.approx.ftz.f32 %0, %0;" : "+f"(f)); a[i] = __float_as_uint(f); }
            if (MODE == 10) a[i] = __float2int_rz(__uint_as_float(a[i])) ;  // F2I
            if (MODE == 11) a[i] = __float_as_uint((float)a[i]);        // I2F
            if (MODE == 12) a[i] = __shfl_up_sync(0xffffffffu, a[i], 1);  // SHFL
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE>
void run(const char* name, int ops_per) {
    unsigned* d; cudaMalloc(&d, 148 * 1024 * 4 * 4);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int iters = 4096;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int warps : {4, 8, 16, 32}) {
        k<MODE><<<148, warps * 32>>>(d, 7, iters);
        cudaEventRecord(e0);
        k<MODE><<<148, warps * 32>>>(d, 7, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double winst = (double)warps * iters * 8 * ops_per;
        printf("%-22s warps/SM %2d  %.3f ms  %.2f warp-inst/ns/SM (x clock GHz = per cycle)\n", name, warps, ms, winst / (ms * 1e6));
    }
    cudaFree(d);
}
int main() {
    run<0>("IADD3", 1); run<1>("IMAD", 1); run<2>("IDP.2A", 1); run<3>("IADD3+IDP mix", 1);
    run<4>("VIMNMX+LOP3", 2); run<5>("IMAD+IADD3 mix", 1); run<6>("VMIN.U16x2+IADD", 2);
    run<7>("SHF", 1); run<8>("PRMT", 1); run<9>("MUFU.SQRT", 1); run<10>("F2I", 1); run<11>("I2F", 1);
    run<12>("SHFL", 1);
    return 0;
}
