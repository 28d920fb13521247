// ubench_issue.cu -- does FFMA2 (fma.rn.f32x2) leave issue slots free for other pipes?  Streams of independent
// FFMA2 / FFMA mixed with independent integer (ALU-pipe) ops; cycles per loop iteration per SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_issue ubench_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long p2;
__device__ __forceinline__ p2 fma2(p2 a, p2 b, p2 c) { p2 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ unsigned alu(unsigned a, unsigned b) { unsigned r; asm volatile("lop3.b32 %0, %1, %2, %1, 0x96;" : "=r"(r) : "r"(a), "r"(b)); return r; }

// NF2 packed FMAs, NF1 scalar FMAs, NA ALU ops per iteration, all on 8 independent chains
template <int NF2, int NF1, int NA>
__global__ void __launch_bounds__(128) kern(float* out, long long* cyc, int iters)
{
    p2 c2[8]; float c1[8]; unsigned u[8];
    p2 a2 = 0x3f8000003f800000ull + threadIdx.x, b2 = 0x3f0000003f000000ull;
#pragma unroll
    for (int i = 0; i < 8; ++i) { c2[i] = i; c1[i] = i; u[i] = i + threadIdx.x; }
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (i < NF2) c2[i] = fma2(a2, b2, c2[i]);
                if (i < NF1) c1[i] = fma1(1.0001f, 0.5f, c1[i]);
                if (i < NA) u[i] = alu(u[i], 0x9e3779b9u);
                if (i + 8 < NA) u[i] = alu(u[i], 0x7f4a7c15u);
            }
        }
    }
    long long t1 = clock64();
    float r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r += c1[i] + (float)(c2[i] & 0xffff) + (float)u[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int NF2, int NF1, int NA> void run(const char* name, int warps_per_smsp)
{
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out; long long* cyc; cudaMalloc(&out, sms * 16 * 128 * 4); cudaMallocManaged(&cyc, 8);
    const int iters = 2000;
    kern<NF2, NF1, NA><<<sms * warps_per_smsp, 128>>>(out, cyc, 10);
    kern<NF2, NF1, NA><<<sms * warps_per_smsp, 128>>>(out, cyc, iters);
    cudaDeviceSynchronize();
    const double per_iter = (double)*cyc / iters / 4.0;      // cycles per unrolled group, one warp's view
    printf("%-34s warps/SMSP=%d: %6.2f cycles per group of (%d FFMA2 + %d FFMA + %d ALU) -> %5.2f cycles per warp-group per SMSP\n",
           name, warps_per_smsp, per_iter, NF2, NF1, NA, per_iter / warps_per_smsp);
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    for (int w : {1, 2, 4}) {
        run<8, 0, 0>("8 FFMA2", w);
        run<8, 0, 8>("8 FFMA2 + 8 ALU", w);
        run<8, 0, 16>("8 FFMA2 + 16 ALU", w);
        run<0, 8, 0>("8 FFMA", w);
        run<0, 8, 8>("8 FFMA + 8 ALU", w);
        run<0, 0, 8>("8 ALU", w);
        run<0, 0, 16>("16 ALU", w);
        run<4, 8, 0>("4 FFMA2 + 8 FFMA", w);
    }
    return 0;
}
