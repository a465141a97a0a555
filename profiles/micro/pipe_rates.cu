// Issue-rate probe for the integer instructions the transform kernel leans on (sm_100a):
// cycles per warp instruction per SM sub-partition for IMAD, IDP.2A, IDP.4A, LOP3 and pairwise mixes.
// One CTA of 4 x W warps per SM (W warps per sub-partition), 8 independent chains per thread.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 2048

template <int MODE>
__global__ void probe(int *out, long long *cyc, int seed)
{
    int v[CHAINS], u[CHAINS];
    for (int i = 0; i < CHAINS; i++) { v[i] = seed + i + threadIdx.x; u[i] = seed * 3 + i; }
    const int k = seed | 0x01010101;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) {
            if (MODE == 0) v[i] = v[i] * k + u[i];                                                     // IMAD
            if (MODE == 1) asm volatile("dp2a.lo.s32.u32 %0, %1, %2, %0;" : "+r"(v[i]) : "r"(k), "r"(u[i]));   // IDP.2A
            if (MODE == 2) asm volatile("dp4a.s32.u32 %0, %1, %2, %0;" : "+r"(v[i]) : "r"(k), "r"(u[i]));      // IDP.4A
            if (MODE == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(k), "r"(u[i]));    // LOP3 (alu)
            if (MODE == 4) {    // IDP.2A + LOP3
                asm volatile("dp2a.lo.s32.u32 %0, %1, %2, %0;" : "+r"(v[i]) : "r"(k), "r"(u[i]));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(k), "r"(seed));
            }
            if (MODE == 5) {    // IMAD + LOP3
                v[i] = v[i] * k + seed;
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(k), "r"(seed));
            }
            if (MODE == 6) {    // IDP.2A + IMAD
                asm volatile("dp2a.lo.s32.u32 %0, %1, %2, %0;" : "+r"(v[i]) : "r"(k), "r"(seed));
                u[i] = u[i] * k + seed;
            }
            if (MODE == 7) {    // IDP.2A + FFMA-class (float fma on the same registers)
                asm volatile("dp2a.lo.s32.u32 %0, %1, %2, %0;" : "+r"(v[i]) : "r"(k), "r"(seed));
                asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+r"(u[i]) : "r"(k));
            }
        }
    }
    long long t1 = clock64();
    int acc = 0;
    for (int i = 0; i < CHAINS; i++) acc += v[i] ^ u[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int per_iter, int warps_per_smsp)
{
    int *out; long long *cyc;
    const int threads = 128 * warps_per_smsp, blocks = 148;
    cudaMalloc(&out, sizeof(int) * threads * blocks);
    cudaMalloc(&cyc, sizeof(long long) * blocks);
    probe<MODE><<<blocks, threads>>>(out, cyc, 7);
    probe<MODE><<<blocks, threads>>>(out, cyc, 7);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < blocks; i++) c += (double)h[i];
    c /= blocks;
    const double inst = (double)ITERS * CHAINS * per_iter * warps_per_smsp;     // warp instructions per sub-partition
    printf("%-22s warps/smsp %d  cycles/warp-inst/smsp %.3f\n", name, warps_per_smsp, c / inst);
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    for (int w = 1; w <= 4; w *= 2) {
        run<0>("IMAD", 1, w);
        run<1>("IDP.2A", 1, w);
        run<2>("IDP.4A", 1, w);
        run<3>("LOP3", 1, w);
        run<4>("IDP.2A+LOP3", 2, w);
        run<5>("IMAD+LOP3", 2, w);
        run<6>("IDP.2A+IMAD", 2, w);
        run<7>("IDP.2A+FFMA", 2, w);
    }
    return 0;
}
