// k_modtilt_scan.cu -- Faust leaf "ModTilt" as five blocked linear-recurrence scans over time (modtilt_scan_core.h):
// ONE WARP PER STREAM, lane l owns 8 consecutive frames of a 256-frame tile, chunk end states cross lanes by shuffle,
// the tile-to-tile carry is the exact sequential state.  4096 streams are 4096 warps = 28 per SM on 148 SMs (7 per
// scheduler, one wave) where the stream-per-thread kernel (k_modtilt.cu) has 128 full warps' worth of work for 592
// schedulers: the time axis supplies the parallelism the batch does not have.
//   * CTA = 4 independent warps (no block-wide barrier); __launch_bounds__(128, 7) -> at most 72 registers, 7 CTAs per SM;
//   * audio: 16-byte cp.async (LDGSTS, no register staging) of the warp's two 1 KB channel rows per tile, double
//     buffered; the 16-byte units are XOR-swizzled (u ^ ((u >> 3) & 1)) so that both the per-lane chunk reads
//     (lane l: units 2l, 2l + 1) and the row-order stores are bank-conflict free; outputs go back through the same
//     tile and leave as full 128-byte lines.  in == out is legal (tile t + 1 is fetched before tile t is stored; they
//     are different frames);
//   * a host block is a whole number of tiles (block_len % 256 == 0, checked by the host), so coefficients are constant
//     inside a tile and parameter changes land on the exact frame; other shapes run on k_modtilt.cu.
// The stream-per-thread kernel stays as the tail / odd-shape path and as the cross-check in the tests.
#include "modtilt_scan_core.h"

namespace za {

namespace {

using namespace mtscan;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ int swz(int u) { return u ^ ((u >> 3) & 1); }

struct TileIO {
    float* buf;                 // [2 channels][256 floats], 16-byte units swizzled
    int lane;
    __device__ __forceinline__ void load(float (&xl)[kC], float (&xr)[kC]) const
    {
#pragma unroll
        for (int j = 0; j < kC / 4; ++j) {
            const int u = swz(lane * (kC / 4) + j);
            const float4 l = *reinterpret_cast<const float4*>(buf + u * 4);
            const float4 r = *reinterpret_cast<const float4*>(buf + kTile + u * 4);
            xl[4 * j] = l.x; xl[4 * j + 1] = l.y; xl[4 * j + 2] = l.z; xl[4 * j + 3] = l.w;
            xr[4 * j] = r.x; xr[4 * j + 1] = r.y; xr[4 * j + 2] = r.z; xr[4 * j + 3] = r.w;
        }
    }
    __device__ __forceinline__ void store(const float (&xl)[kC], const float (&xr)[kC]) const
    {
#pragma unroll
        for (int j = 0; j < kC / 4; ++j) {
            const int u = swz(lane * (kC / 4) + j);
            *reinterpret_cast<float4*>(buf + u * 4) = make_float4(xl[4 * j], xl[4 * j + 1], xl[4 * j + 2], xl[4 * j + 3]);
            *reinterpret_cast<float4*>(buf + kTile + u * 4) = make_float4(xr[4 * j], xr[4 * j + 1], xr[4 * j + 2], xr[4 * j + 3]);
        }
    }
};

constexpr int kWarps = 4;

template <bool FAST>
__global__ void __launch_bounds__(32 * kWarps, 7) modtilt_scan_kernel(const MTScanArgs A)
{
    __shared__ __align__(16) float sm[kWarps][2][2 * kTile];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * kWarps + warp;
    if (s >= A.a.n_streams) return;                      // warps are independent: no block-wide barrier below
    const MTArgs& a = A.a;
    const size_t ns = (size_t)a.n_streams;
    const float* in0 = a.in + (size_t)s * 2 * a.row_stride;
    float* out0 = a.out + (size_t)s * 2 * a.row_stride;
    const int64_t n_tiles = a.n_frames / kTile;
    const int tpb = a.block_len / kTile;                 // tiles per host block

    Carry cy{a.state[s], a.state[ns + s], a.state[2 * ns + s], a.state[3 * ns + s], a.state[4 * ns + s]};
    const float q_env = A.qfix[lane], q_base = A.qfix[32 + lane], q_ratio = A.qfix[64 + lane], q_trim = A.qfix[96 + lane];
    float q_piv = 0.f;
    MTCoef c{};
    int cb = -1;

    auto issue = [&](int64_t t) {
        float* buf = sm[warp][t & 1];
#pragma unroll
        for (int ch = 0; ch < 2; ++ch)
#pragma unroll
            for (int j = 0; j < kTile / 128; ++j) {
                const int u = lane + 32 * j;
                cp_async16(buf + ch * kTile + swz(u) * 4, in0 + (size_t)ch * a.row_stride + t * kTile + u * 4);
            }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    if (n_tiles > 0) issue(0);
    int in_block = 0;
    int64_t blk = 0;
    for (int64_t t = 0; t < n_tiles; ++t) {
        if (t + 1 < n_tiles) {
            issue(t + 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        if (in_block == 0) {                             // first tile of a host block: its coefficient row (the last row persists)
            const int b = blk < a.n_coef_blocks ? (int)blk : a.n_coef_blocks - 1;
            if (b != cb) {
                cb = b;
                const float4* p = reinterpret_cast<const float4*>(a.coef + (size_t)b * ns + s);
                const float4 u = __ldg(p), v = __ldg(p + 1);
                c.a_piv = u.x; c.na_piv = u.y; c.g_lo = u.z; c.g_hi = u.w; c.mix = v.x; c.nmix = v.y;
                q_piv = __ldg(A.qpiv + (size_t)__float_as_int(v.z) * 32 + lane);
            }
        }
        if (++in_block == tpb) { in_block = 0; ++blk; }
        TileIO io{sm[warp][t & 1], lane};
        tile<FAST>(cy, A.K, q_env, q_base, q_ratio, q_trim, q_piv, c, io);
        __syncwarp();
        {
            const float* buf = sm[warp][t & 1];
#pragma unroll
            for (int ch = 0; ch < 2; ++ch)
#pragma unroll
                for (int j = 0; j < kTile / 128; ++j) {
                    const int u = lane + 32 * j;
                    const float4 v = *reinterpret_cast<const float4*>(buf + ch * kTile + swz(u) * 4);
                    *reinterpret_cast<float4*>(out0 + (size_t)ch * a.row_stride + t * kTile + u * 4) = v;
                }
        }
        __syncwarp();
    }
    if (lane == 0) {
        a.state[s] = cy.env2; a.state[ns + s] = cy.base; a.state[2 * ns + s] = cy.lp;
        a.state[3 * ns + s] = cy.rs; a.state[4 * ns + s] = cy.mean;
    }
}

}  // namespace

bool modtilt_scan_ok(const MTArgs& a)
{
    return a.block_len % mtscan::kTile == 0 && a.n_frames % a.block_len == 0 && a.row_stride % 4 == 0 &&
           ((uintptr_t)a.in & 15) == 0 && ((uintptr_t)a.out & 15) == 0;
}

bool modtilt_scan_launch(const MTScanArgs& A, cudaStream_t stream, bool exact_maps)
{
    if (A.a.n_streams <= 0 || A.a.n_frames <= 0) return true;
    const unsigned grid = (unsigned)((A.a.n_streams + kWarps - 1) / kWarps);
    if (exact_maps) modtilt_scan_kernel<false><<<grid, 32 * kWarps, 0, stream>>>(A);
    else modtilt_scan_kernel<true><<<grid, 32 * kWarps, 0, stream>>>(A);
    return cuda_ok(cudaGetLastError(), "modtilt scan launch");
}

}  // namespace za
