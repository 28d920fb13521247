// za_device.cuh -- sm_100a device helpers shared by the plugin kernels:
// rounding-pinned arithmetic, TMA (cp.async.bulk.tensor) + mbarrier wrappers, swizzled tile access.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace za {

constexpr int kL = 32;   // frames per thread-chunk = one 128-byte row of a SWIZZLE_128B tile

// ---- arithmetic ------------------------------------------------------------------------------
// fp64 mode must reproduce the reference's operation order exactly: the AOT compiler emits
// separate fmul/fadd/fsub for generic x86-64 (no FMA, no fast-math; dsp_jsfx_aot.py:6211-6228), so
// double ops go through the _rn intrinsics, which ptxas never contracts.  fp32 mode only has to
// null below -100 dBFS, so float ops stay contractible.
__device__ __forceinline__ float mul(float a, float b) { return a * b; }
__device__ __forceinline__ float add(float a, float b) { return a + b; }
__device__ __forceinline__ float sub(float a, float b) { return a - b; }
__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
// explicitly fused (scan/carry math, where only accuracy matters)
__device__ __forceinline__ float fma_(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_(double a, double b, double c) { return fma(a, b, c); }

template <typename R> struct Vec2;
template <> struct Vec2<float> { using type = float2; };
template <> struct Vec2<double> { using type = double2; };
template <typename R> struct Quad { R x, y, z, w; };

__device__ __forceinline__ Quad<float> ld4(const float* p)
{
    const float4 v = *reinterpret_cast<const float4*>(p);
    return {v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ Quad<double> ld4(const double* p)
{
    const double2 a = *reinterpret_cast<const double2*>(p);
    const double2 b = *reinterpret_cast<const double2*>(p + 2);
    return {a.x, a.y, b.x, b.y};
}

__device__ __forceinline__ float shfl_up(float v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
__device__ __forceinline__ double shfl_up(double v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
__device__ __forceinline__ float shfl_idx(float v, int l) { return __shfl_sync(0xffffffffu, v, l); }
__device__ __forceinline__ double shfl_idx(double v, int l) { return __shfl_sync(0xffffffffu, v, l); }

// ---- shared-memory addresses, mbarrier, TMA ---------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_barrier_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// 3-D tiled load: coordinates (c0 = frame-in-chunk, c1 = chunk, c2 = row)
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void tma_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* map)
{
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// ---- swizzled tile rows ------------------------------------------------------------------------
// A tile is T rows x 32 floats (128 B per row), 1024-byte aligned, laid out as TMA SWIZZLE_128B
// writes it: the 16-byte piece j of row r sits at piece (j ^ (r & 7)).  Thread r reading piece j
// of its own row is therefore bank-conflict free across a quarter-warp.
__device__ __forceinline__ int swz_piece(int row, int piece) { return row * kL + ((piece ^ (row & 7)) << 2); }
__device__ __forceinline__ int swz_elem(int row, int col) { return row * kL + ((((col >> 2) ^ (row & 7))) << 2) + (col & 3); }

template <typename R>
__device__ __forceinline__ void tile_row_load(const float* tile, int row, R (&x)[kL])
{
#pragma unroll
    for (int j = 0; j < kL / 4; ++j) {
        const float4 v = *reinterpret_cast<const float4*>(tile + swz_piece(row, j));
        x[4 * j + 0] = (R)v.x; x[4 * j + 1] = (R)v.y; x[4 * j + 2] = (R)v.z; x[4 * j + 3] = (R)v.w;
    }
}
template <typename R>
__device__ __forceinline__ void tile_row_store(float* tile, int row, const R (&x)[kL])
{
#pragma unroll
    for (int j = 0; j < kL / 4; ++j) {
        float4 v;
        v.x = (float)x[4 * j + 0]; v.y = (float)x[4 * j + 1]; v.z = (float)x[4 * j + 2]; v.w = (float)x[4 * j + 3];
        *reinterpret_cast<float4*>(tile + swz_piece(row, j)) = v;
    }
}

}  // namespace za
