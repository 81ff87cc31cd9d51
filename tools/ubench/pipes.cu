// pipes.cu -- per-SM issue rates of the instruction classes the demodulator's per-sample chain uses
// (DFMA, 64-bit conversions, packed FP32, integer), measured with clock64 on one resident grid.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu ; run: ./pipes
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ILP 8
#define ITERS 2048

template <int OP>
__global__ void k(double* out, long long* cyc, double seed_d, float seed_f, int seed_i)
{
    double d[ILP]; float f[ILP]; int n[ILP]; unsigned long long p[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) { d[j] = seed_d + j + threadIdx.x; f[j] = seed_f + j + threadIdx.x; n[j] = seed_i + j + threadIdx.x; p[j] = (unsigned long long)n[j] * 0x100000001ull; }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            if (OP == 0) d[j] = __fma_rn(d[j], 1.0000001, 0.5);
            if (OP == 1) { f[j] = (float)d[j]; d[j] = d[j] + 1.0; }                 // D2F + DADD
            if (OP == 2) { d[j] = (double)f[j]; f[j] = f[j] + 1.0f; }               // F2D + FADD
            if (OP == 3) { n[j] = (int)d[j]; d[j] = d[j] + 1.0; }                   // D2I + DADD
            if (OP == 4) { d[j] = (double)n[j]; n[j] = n[j] + 1; }                  // I2D + IADD
            if (OP == 5) f[j] = __fmaf_rn(f[j], 1.0000001f, 0.5f);
            if (OP == 6) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[j]) : "l"(0x3f8000013f800001ull));
            if (OP == 7) n[j] = (n[j] ^ (n[j] >> 3)) + 0x9e3779b9;                  // 2-3 ALU ops
            if (OP == 8) d[j] = d[j] + 1.0;                                          // DADD alone
            if (OP == 9) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[j]) : "l"(0x3f8000013f800001ull));
            if (OP == 10) n[j] = __float2int_rn(f[j]) , f[j] = f[j] + 1.0f;          // F2I + FADD
            if (OP == 11) { d[j] = __dmul_rn(d[j], 1.0000001); }
        }
    }
    long long t1 = clock64();
    double acc = 0;
#pragma unroll
    for (int j = 0; j < ILP; ++j) acc += d[j] + f[j] + n[j] + (double)p[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int ops_per_iter)
{
    int sm; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
    const int threads = 1024, blocks = sm;   // 32 warps per SM
    double* out; long long* cyc;
    cudaMalloc(&out, sizeof(double) * threads * blocks); cudaMalloc(&cyc, sizeof(long long) * blocks);
    k<OP><<<blocks, threads>>>(out, cyc, 1.0, 1.0f, 1);
    k<OP><<<blocks, threads>>>(out, cyc, 1.0, 1.0f, 1);
    cudaDeviceSynchronize();
    long long h[1024]; cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < blocks; ++i) avg += h[i]; avg /= blocks;
    double per_clk = (double)threads * ILP * ITERS * ops_per_iter / avg;
    printf("%-28s %8.1f thread-ops/clk/SM  (%s)\n", name, per_clk, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<0>("DFMA", 1);
    run<11>("DMUL", 1);
    run<8>("DADD", 1);
    run<1>("D2F (+DADD)", 1);
    run<2>("F2D (+FADD)", 1);
    run<3>("D2I (+DADD)", 1);
    run<4>("I2D (+IADD)", 1);
    run<5>("FFMA", 1);
    run<6>("FMUL2 (f32x2)", 1);
    run<9>("FADD2 (f32x2)", 1);
    run<7>("INT xor/shr/add (3 ops)", 3);
    run<10>("F2I (+FADD)", 1);
    return 0;
}
