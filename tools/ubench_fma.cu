// ubench_fma.cu -- FP pipe microbenchmark for design decisions (sm_100a): issue rate of FFMA (3 distinct
// registers), FFMA2 (fma.rn.f32x2), DFMA, DADD+DMUL, and dependent-chain latency of each.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fma ubench_fma.cu ; run on the B200.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE, int CH>
__global__ void __launch_bounds__(256) kern(float* out, int iters, float seed)
{
    // CH independent chains per thread
    float a[CH], b[CH], c[CH];
    float2 a2[CH], b2[CH], c2[CH];
    double da[CH], db[CH], dc[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        a[i] = seed + i + threadIdx.x; b[i] = seed * 0.5f + i; c[i] = i;
        a2[i] = make_float2(a[i], a[i] + 1); b2[i] = make_float2(b[i], b[i] + 1); c2[i] = make_float2(c[i], c[i] + 1);
        da[i] = a[i]; db[i] = b[i]; dc[i] = c[i];
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                if (MODE == 0) c[i] = fmaf(a[i], b[i], c[i]);
                if (MODE == 1) c2[i] = __ffma2_rn(a2[i], b2[i], c2[i]);
                if (MODE == 2) dc[i] = fma(da[i], db[i], dc[i]);
                if (MODE == 3) dc[i] = __dadd_rn(__dmul_rn(da[i], db[i]), dc[i]);
                if (MODE == 4) c[i] = fmaf(a[0], b[0], c[i]);      // shared multiplicands (reuse)
                if (MODE == 5) c2[i] = __ffma2_rn(a2[0], b2[i], c2[i]);
            }
        }
    }
    float r = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) r += c[i] + c2[i].x + c2[i].y + (float)dc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE, int CH> void run(const char* name, int flops_per, int blocks_per_sm, int threads)
{
    int dev = 0, sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    float* out; cudaMalloc(&out, (size_t)sms * blocks_per_sm * threads * 4);
    const int iters = 4096;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<MODE, CH><<<sms * blocks_per_sm, threads>>>(out, 64, 1.0f);
    cudaEventRecord(e0);
    kern<MODE, CH><<<sms * blocks_per_sm, threads>>>(out, iters, 1.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)sms * blocks_per_sm * threads * iters * 8.0 * CH * flops_per;   // FMAs
    const double per_clk_sm = ops / (ms * 1e-3) / (khz * 1e3) / sms;
    printf("%-28s CH=%2d warps/SM=%3d  %8.3f ms  %7.1f FMA/clk/SM (at %d MHz nominal)  %7.2f TFMA/s\n", name, CH,
           blocks_per_sm * threads / 32, ms, per_clk_sm, khz / 1000, ops / (ms * 1e-3) / 1e12);
    cudaFree(out);
}

int main()
{
    for (int w : {4, 8, 16, 32}) {
        const int thr = 128, bps = w * 32 / thr;
        run<0, 8>("FFMA 3 distinct regs", 1, bps, thr);
        run<4, 8>("FFMA shared a,b", 1, bps, thr);
        run<1, 8>("FFMA2 distinct", 2, bps, thr);
        run<5, 8>("FFMA2 shared a", 2, bps, thr);
        run<2, 8>("DFMA", 1, bps, thr);
        run<3, 8>("DMUL+DADD (counted as 1)", 1, bps, thr);
    }
    // dependent-chain latency: 1 chain, 1 warp per SMSP
    run<0, 1>("FFMA chain", 1, 1, 128);
    run<1, 1>("FFMA2 chain", 2, 1, 128);
    run<2, 1>("DFMA chain", 1, 1, 128);
    run<3, 1>("DMUL+DADD chain", 1, 1, 128);
    run<0, 2>("FFMA 2 chains", 1, 1, 128);
    run<1, 2>("FFMA2 2 chains", 2, 1, 128);
    run<0, 4>("FFMA 4 chains", 1, 1, 128);
    run<1, 4>("FFMA2 4 chains", 2, 1, 128);
    return 0;
}
