// Issue-rate microbenchmark for the integer ops the scan kernel is made of (sm_100a).
// Prints warp-instructions per clock per SM for each op.  nvcc -arch=sm_100a -O3 pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITER 4096
template <int OP> __global__ void k(uint32_t *out, uint32_t seed, long long *clk)
{
    __shared__ uint32_t sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = i * 2654435761u;
    __syncthreads();
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + threadIdx.x * 8 + i;
    uint32_t b = seed ^ 0x12345, c = seed | 7;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) a[i] = (a[i] ^ b) & c;                                   // LOP3
            if (OP == 1) a[i] = __funnelshift_r(a[i], b, c);                      // SHF
            if (OP == 2) a[i] = a[i] + b + c;                                     // IADD3
            if (OP == 3) a[i] = a[i] * b + c;                                     // IMAD
            if (OP == 4) a[i] = __dp4a(a[i], b, c);                               // IDP4A
            if (OP == 5) a[i] = __popc(a[i]) + b;                                 // POPC (+IADD)
            if (OP == 6) a[i] = sm[(a[i] >> 5) & 2047];                           // LDS random (+2 alu)
            if (OP == 7) a[i] = sm[((a[i] >> 5) & 63) * 32 + (threadIdx.x & 31)]; // LDS conflict-free
            if (OP == 8) a[i] = __shfl_sync(0xffffffffu, a[i], b & 31);           // SHFL
            if (OP == 9) a[i] = __byte_perm(a[i], b, c);                          // PRMT
            if (OP == 10) { a[i] = (a[i] ^ b) & c; a[(i + 1) & 7] = a[(i + 1) & 7] * b + c; } // LOP3 + IMAD mix
            if (OP == 11) a[i] = __vcmpeq4(a[i], b) + c;
            if (OP == 12) a[i] = __ffs(a[i]) + b;
            if (OP == 13) a[i] = __vcmpgeu4(a[i], b) + c;
        }
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
template <int OP> void run(const char *name, int instr_per_op)
{
    uint32_t *out; long long *clk, h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 8);
    k<OP><<<148, 1024>>>(out, 3, clk); cudaDeviceSynchronize();
    k<OP><<<148, 1024>>>(out, 3, clk); cudaDeviceSynchronize();
    cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    double ops = 32.0 * ITER * 8;   // warp-ops per SM (32 warps)
    printf("%-22s %8.3f warp-ops/clk/SM  (%d SASS instr per op => %.2f instr/clk/SM)\n", name, ops / h, instr_per_op,
           ops * instr_per_op / h);
    cudaFree(out); cudaFree(clk);
}
int main()
{
    run<0>("LOP3", 1); run<1>("SHF", 1); run<2>("IADD3", 1); run<3>("IMAD", 1); run<4>("IDP4A", 1);
    run<5>("POPC+IADD", 2); run<6>("LDS random +2alu", 3); run<7>("LDS conflict-free", 4); run<8>("SHFL", 1);
    run<9>("PRMT", 1); run<10>("LOP3+IMAD pair", 2); run<11>("vcmpeq4+add", 1); run<12>("FFS+add", 1); run<13>("vcmpgeu4+add", 1);
    return 0;
}
