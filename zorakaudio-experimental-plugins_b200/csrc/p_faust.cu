// p_faust.cu -- the sequential Faust leaves RED, ClickBeGoneSG and VAR behind the C ABI: one generic stream-per-thread
// kernel over the leaf structs of faust_leaves.cuh plus its host half.  Path replaced: FaustJuceProcessor::prepareToPlay /
// processBlock (src/FaustJuceProcessor.cpp:430-440, 463-482) around mydsp::init / mydsp::compute
// (src/faust_support_min.h:41-53), for a batch of independent instances.  Float32 only (the reference builds Faust leaves
// without -double); compiled with -fmad=false (za_gpu/build.py) so the .dsp's operation order is not contracted.
//   * kernel: one warp per CTA, lane = stream, state in registers / local memory for the whole call; audio moves as
//     [channels x 32 streams] x 32-frame tiles through shared memory so HBM sees 128-byte row segments (in == out legal);
//     the next tile arrives by cp.async while the current one is computed (with one warp per SM a plain load loop paid a
//     full memory latency per row: 192 rows per tile for the six-channel RED);
//   * slider-rate values (Faust's fSlow expressions) are evaluated on the HOST with the host libm once per (host block,
//     stream) whose zone values changed, uploaded as rows;
//   * zones: the quantised value narrowed to float (FaustJuceProcessor.cpp:467-471).
// These leaves are sequential by nature (switching followers, max-hold envelopes, hysteresis): the kernel is
// latency-bound at 4096 streams like every stream-per-thread kernel here; parity target = the float restatement
// (the test-side C port of the .dsp; PARITY UNPINNED).
#include <cstddef>
#include <cstring>

#include "za_host.h"
#include "faust_leaves.cuh"

namespace za {

namespace {

struct FaustArgs {
    const float* in; float* out;
    int64_t row_stride, n_frames;
    int n_streams, block_len, n_rows;
    const void* slow;           // [n_rows][n_streams] Leaf::Slow
    unsigned* state;            // [words][n_streams]
    int lanes;                  // generic kernel: streams per warp (8 / 16 / 32)
    int force_tiled;            // option scalar_f32 = 2: the tiled kernel whatever the batch size
};

constexpr int kP = 33;

template <class L>
__global__ void __launch_bounds__(128) faust_leaf_kernel(const FaustArgs a, const typename L::Konst k)
{
    constexpr int kCh = L::kIn > L::kOut ? L::kIn : L::kOut;
    constexpr int kWords = (int)(sizeof(typename L::State) / 4);
    // Four independent warps per CTA (they never synchronise with each other), SL = a.lanes streams per warp: a stream's
    // frame() is one long dependent chain, so a warp with 8 busy lanes runs as fast as one with 32 and a few thousand
    // streams are spread over 4x as many warps; all 32 lanes move the audio tiles.
    const int SL = a.lanes, kTile = kCh * SL * kP;
    extern __shared__ __align__(16) float faust_tiles_cta[];    // per warp two tiles: the next one arrives by cp.async while this one is computed
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, s0 = (blockIdx.x * (blockDim.x >> 5) + warp) * SL, s = s0 + lane;
    if (s0 >= a.n_streams) return;
    float* const faust_tiles = faust_tiles_cta + warp * 2 * kTile;
    const bool live = lane < SL && s < a.n_streams;
    const int n_w = min(SL, a.n_streams - s0);
    const size_t ns = (size_t)a.n_streams;
    typename L::State st;
    unsigned* sw = reinterpret_cast<unsigned*>(&st);
    // small states: fully unrolled word copies, so every access to `st` has a constant index and it lives in registers
    // (a rolled loop takes its address and pins the whole struct to local memory)
    constexpr int kUnroll = kWords <= 48 ? kWords : 1;
#pragma unroll kUnroll
    for (int w = 0; w < kWords; ++w) sw[w] = live ? a.state[(size_t)w * ns + s] : 0u;
    const typename L::Slow* rows = reinterpret_cast<const typename L::Slow*>(a.slow);
    typename L::Slow c = rows[live ? s : 0];
    int blk = 0;
    int64_t blk_end = a.block_len;
    // row q of the CTA: stream q / kIn, channel q % kIn; lane = frame.  One 4-byte cp.async per element (the tile pitch of 33
    // keeps lane = stream accesses conflict-free, so rows are not 16-byte aligned); frames past the call are zero-filled.
    auto fetch = [&](int64_t f0, float* tile) {
        const int nv = (int)min((int64_t)32, a.n_frames - f0);
        const int bytes = lane < nv ? 4 : 0;
#pragma unroll 4
        for (int q = 0; q < n_w * L::kIn; ++q) {
            const int si = q / L::kIn, ch = q - si * L::kIn;
            const float* src = a.in + ((size_t)(s0 + si) * L::kIn + ch) * a.row_stride + (bytes ? f0 + lane : 0);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(tile + (ch * SL + si) * kP + lane)),
                         "l"(src), "r"(bytes) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (a.n_frames > 0) fetch(0, faust_tiles);
    int buf = 0;
    for (int64_t f0 = 0; f0 < a.n_frames; f0 += 32, buf ^= 1) {
        const int nv = (int)min((int64_t)32, a.n_frames - f0);
        float* tile = faust_tiles + buf * kTile;
        if (f0 + 32 < a.n_frames) {
            fetch(f0 + 32, faust_tiles + (buf ^ 1) * kTile);     // in == out is legal: those frames are stored two iterations from now
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        if (live) {
            for (int f = 0; f < nv; ++f) {
                if (f0 + f >= blk_end) {                        // a new host block: its slider-rate row (the last row persists)
                    ++blk; blk_end += a.block_len;
                    c = rows[(size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns + s];
                }
                float iv[L::kIn], ov[L::kOut];
#pragma unroll
                for (int ch = 0; ch < L::kIn; ++ch) iv[ch] = tile[(ch * SL + lane) * kP + f];
                L::frame(st, c, k, iv, ov);
#pragma unroll
                for (int ch = 0; ch < L::kOut; ++ch) tile[(ch * SL + lane) * kP + f] = ov[ch];
            }
        }
        __syncwarp();
#pragma unroll 4
        for (int q = 0; q < n_w * L::kOut; ++q) {
            const int si = q / L::kOut, ch = q - si * L::kOut;
            if (lane < nv)
                a.out[((size_t)(s0 + si) * L::kOut + ch) * a.row_stride + f0 + lane] = tile[(ch * SL + si) * kP + lane];
        }
        __syncwarp();
    }
    if (live)
#pragma unroll kUnroll
        for (int w = 0; w < kWords; ++w) a.state[(size_t)w * ns + s] = sw[w];
}

template <class L> bool launch_generic(FaustArgs a, const typename L::Konst& k, cudaStream_t st)
{
    constexpr int kCh = L::kIn > L::kOut ? L::kIn : L::kOut;
    int SL = a.lanes;
    if (SL != 8 && SL != 16) SL = 32;                // option "lanes_per_warp"; narrower warps measured +7 % solo, -15 % in the cfg5 mix (jsfx_rt.cuh)
    a.lanes = SL;
    const int W = 1;                                 // one warp per CTA (the round-1 shape; see the measurement note in jsfx_rt.cuh)
    const size_t smem = (size_t)W * 2 * kCh * SL * kP * sizeof(float);
    if (smem > 48 * 1024 && !cuda_ok(cudaFuncSetAttribute(faust_leaf_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "Faust leaf smem attr")) return false;
    faust_leaf_kernel<L><<<(unsigned)((a.n_streams + SL * W - 1) / (SL * W)), 32 * W, smem, st>>>(a, k);
    return cuda_ok(cudaGetLastError(), "faust_leaf_kernel");
}

// ---- ClickBeGoneSG, tiled ---------------------------------------------------------------------------------------------------
// The leaf's cost is its feed-forward half: two Savitzky-Golay predictors per channel on delayed taps (72 .. 104 multiply-adds
// per frame), which the generic kernel above evaluates lane = stream out of a dynamically indexed 32-entry delay line in
// local memory (one warp per SM at 4096 streams, ~12 000 cycles per frame).  Here one CTA = 32 streams, 8 FIR warps + 1
// recurrence warp, software-pipelined over 32-frame tiles with one __syncthreads per tile:
//   FIR warp w (lane = frame), streams 4w .. 4w + 3:  store the outputs of tile i - 2 | cp.async the inputs of tile i + 1 into a
//     128-frame shared-memory ring | predictors of tile i with compile-time taps (Cbg::predict, mode from the frame's row);
//   recurrence warp (lane = stream): the detector high-pass, max-hold envelope, baseline, trigger hold and mix of tile i - 1
//     (Cbg::tail), frame after frame.
// Both halves are the SAME functions the generic kernel calls, in the same operation order, so the outputs and the state rows
// (delay-line ring included: same entries at the same indices) are bit-identical to the generic kernel's.
namespace cbg {
constexpr int kFirWarps = 8, kThreads = (kFirWarps + 1) * 32;
constexpr int kRing = 128, kRingP = 2 * kRing + 1;      // [stream][ch][128 frames] + 1: conflict-free for lane = frame and lane = stream
constexpr int kTP = 33;                                 // [stream][frame] tiles
struct Smem {
    float ring[32 * kRingP];
    float fir[2][5][32 * kTP];                          // xC_L, xC_R, pred_L, pred_R, e_norm of a tile
    float out[2][2][32 * kTP];
};
struct RingX { const float* row; int n; __device__ __forceinline__ float operator()(int d) const { return row[(n - d) & (kRing - 1)]; } };
}  // namespace cbg

__global__ void __launch_bounds__(cbg::kThreads) cbg_tiled_kernel(const FaustArgs a)
{
    using L = faust::Cbg;
    using namespace cbg;
    extern __shared__ __align__(16) unsigned char cbg_smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(cbg_smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, s0 = blockIdx.x * 32;
    const size_t ns = (size_t)a.n_streams;
    const int n_cta = min(32, a.n_streams - s0);
    const L::Slow* rows = reinterpret_cast<const L::Slow*>(a.slow);
    const int64_t n = a.n_frames;
    const int n_tiles = (int)((n + 31) / 32);
    // State words (faust::Cbg::State): hL[32] 0..31, hR[32] 32..63, pos 64, hpL 65, hpR 66, env 67, base 68, holdEnv 69
    if (warp < kFirWarps) {
        int pos0[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {                    // frames -1 - d (d = lane) of the call = delay-line entry (pos - d) & 31
            const int si = warp * 4 + q;
            pos0[q] = 0;
            if (si < n_cta) {
                const size_t s = (size_t)(s0 + si);
                pos0[q] = (int)a.state[64 * ns + s] & 31;
                float* r = sm.ring + si * kRingP;
                r[(-1 - lane) & (kRing - 1)] = __uint_as_float(a.state[(size_t)((pos0[q] - lane) & 31) * ns + s]);
                r[kRing + ((-1 - lane) & (kRing - 1))] = __uint_as_float(a.state[(size_t)(32 + ((pos0[q] - lane) & 31)) * ns + s]);
            }
        }
        __syncwarp();
        // inputs of tile t into the ring by cp.async (4 bytes per lane and row; frames past the call are zero-filled): issued one
        // tile ahead, so the FIR warps do not sit out a DRAM latency per tile (ncu before: 35 % of the samples on these stores)
        auto fetch = [&](int t) {
            if (t < n_tiles) {
                const int64_t nn = (int64_t)t * 32 + lane;
                const int bytes = nn < n ? 4 : 0;
                const int rp = (int)(nn & (kRing - 1));
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int si = warp * 4 + q;
                    if (si < n_cta) {
                        const float* x = a.in + (size_t)(s0 + si) * 2 * a.row_stride + (bytes ? nn : 0);
                        float* r = sm.ring + si * kRingP;
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(r + rp)), "l"(x), "r"(bytes) : "memory");
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(r + kRing + rp)), "l"(x + a.row_stride), "r"(bytes) : "memory");
                    }
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        fetch(0);
        int mode_q[4] = {0, 0, 0, 0};
        bool was_ragged = false;
        for (int i = 0; i < n_tiles + 2; ++i) {
            asm volatile("cp.async.wait_group 0;" ::: "memory");       // tile i is in the ring
            __syncwarp();
            if (i >= 2) {                                // outputs of tile i - 2
                const int t = i - 2;
                const int64_t nn = (int64_t)t * 32 + lane;
                if (nn < n) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int si = warp * 4 + q;
                        if (si < n_cta) {
                            float* o = a.out + (size_t)(s0 + si) * 2 * a.row_stride + nn;
                            o[0] = sm.out[t & 1][0][si * kTP + lane];
                            o[a.row_stride] = sm.out[t & 1][1][si * kTP + lane];
                        }
                    }
                }
            }
            fetch(i + 1);                                // in == out is legal: tile i + 1 is stored three iterations from now
            if (i < n_tiles) {
                const int64_t f0 = (int64_t)i * 32, nn = f0 + lane;
                const int rp = (int)(nn & (kRing - 1));
                // slider-rate row of this lane's frame (the last row persists)
                const int64_t r0 = f0 % a.block_len;
                int64_t blk = f0 / a.block_len;
                if (r0 + 32 > a.block_len) blk = nn / a.block_len;
                const size_t row = (size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns;
                const bool ragged = r0 + 32 > a.block_len;                          // the tile spans host blocks: row per lane
                const bool reread = i == 0 || r0 == 0 || ragged || was_ragged;      // else the row can only change where a host block starts
                was_ragged = ragged;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int si = warp * 4 + q;
                    if (si < n_cta) {
                        const float* r = sm.ring + si * kRingP;
                        const RingX xl{r, rp}, xr{r + kRing, rp};
                        if (reread) mode_q[q] = rows[row + s0 + si].mode;
                        const int mode = mode_q[q];
                        L::Pred p;
                        if (mode == 0) p = L::predict(xl, xr, 0);
                        else if (mode == 1) p = L::predict(xl, xr, 1);
                        else p = L::predict(xl, xr, 2);
                        float (*fo)[32 * kTP] = sm.fir[i & 1];
                        const int o = si * kTP + lane;
                        fo[0][o] = p.xC_L; fo[1][o] = p.xC_R; fo[2][o] = p.pred_L; fo[3][o] = p.pred_R; fo[4][o] = p.e_norm;
                    }
                }
            }
            __syncthreads();
        }
        // the delay line as the generic kernel leaves it: pos advanced by n, entry (pos - d) & 31 = frame n - 1 - d
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int si = warp * 4 + q;
            if (si < n_cta) {
                const size_t s = (size_t)(s0 + si);
                const int pos = (int)((pos0[q] + n) & 31);
                const float* r = sm.ring + si * kRingP;
                const int rp = (int)((n - 1 - lane) & (kRing - 1));
                a.state[(size_t)((pos - lane) & 31) * ns + s] = __float_as_uint(r[rp]);
                a.state[(size_t)(32 + ((pos - lane) & 31)) * ns + s] = __float_as_uint(r[kRing + rp]);
                if (lane == 0) a.state[64 * ns + s] = (unsigned)pos;
            }
        }
    } else {
        const bool live = lane < n_cta;
        const size_t s = (size_t)(s0 + (live ? lane : 0));
        L::Rec m{__uint_as_float(a.state[65 * ns + s]), __uint_as_float(a.state[66 * ns + s]), __uint_as_float(a.state[67 * ns + s]),
                 __uint_as_float(a.state[68 * ns + s]), __uint_as_float(a.state[69 * ns + s])};
        L::Slow c = rows[s];
        int blk = 0;
        int64_t blk_end = a.block_len;
        const float* r = sm.ring + (live ? lane : 0) * kRingP;
        for (int i = 0; i < n_tiles + 2; ++i) {
            if (live && i >= 1 && i - 1 < n_tiles) {
                const int t = i - 1;
                const int64_t f0 = (int64_t)t * 32;
                const int nv = (int)min((int64_t)32, n - f0);
                float (*fi)[32 * kTP] = sm.fir[t & 1];
                float (*fo)[32 * kTP] = sm.out[t & 1];
                if (f0 >= blk_end) {                     // the tile starts a new host block (block_len % 32 == 0 is the common case)
                    ++blk; blk_end += a.block_len;
                    c = rows[(size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns + s];
                }
                if (nv == 32 && f0 + 32 <= blk_end) {
                    // whole tile under one slider row: unrolled, so that the loads and the divisions of neighbouring frames overlap
                    // and only the four carried chains (hp, env, base, holdEnv) stay serial
                    const int rb = (int)(f0 & (kRing - 1));
                    float pl = r[(rb - 1) & (kRing - 1)], pr = r[kRing + ((rb - 1) & (kRing - 1))];
#pragma unroll 8
                    for (int f = 0; f < 32; ++f) {
                        const int o = lane * kTP + f;
                        const L::Pred p{fi[0][o], fi[1][o], fi[2][o], fi[3][o], fi[4][o]};
                        const float xl = r[rb + f], xr = r[kRing + rb + f];
                        float ov[2];
                        L::tail(m, c, xl, xr, pl, pr, p, ov);
                        pl = xl; pr = xr;
                        fo[0][o] = ov[0]; fo[1][o] = ov[1];
                    }
                } else
                for (int f = 0; f < nv; ++f) {
                    const int64_t nn = f0 + f;
                    if (nn >= blk_end) {                 // a new host block: its slider-rate row (the last row persists)
                        ++blk; blk_end += a.block_len;
                        c = rows[(size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns + s];
                    }
                    const int rp = (int)(nn & (kRing - 1)), rq = (int)((nn - 1) & (kRing - 1));
                    const int o = lane * kTP + f;
                    const L::Pred p{fi[0][o], fi[1][o], fi[2][o], fi[3][o], fi[4][o]};
                    float ov[2];
                    L::tail(m, c, r[rp], r[kRing + rp], r[rq], r[kRing + rq], p, ov);
                    fo[0][o] = ov[0]; fo[1][o] = ov[1];
                }
            }
            __syncthreads();
        }
        if (live) {
            a.state[65 * ns + s] = __float_as_uint(m.hpL); a.state[66 * ns + s] = __float_as_uint(m.hpR);
            a.state[67 * ns + s] = __float_as_uint(m.env); a.state[68 * ns + s] = __float_as_uint(m.base);
            a.state[69 * ns + s] = __float_as_uint(m.holdEnv);
        }
    }
}
static_assert(sizeof(faust::Cbg::State) == 70 * 4, "cbg_tiled_kernel addresses the state words of faust::Cbg::State by number");

// ---- RED, tiled ---------------------------------------------------------------------------------------------------------------
// RED's frame is a chain of two envelope followers, a memoryless level detector (two sqrt, a division, log10, a knee), six more
// followers and a memoryless dB -> gain (powf): the generic kernel above walks that chain of ~300 dependent instructions once per
// frame with one warp per SM (48.8 ms per 5041 streams x 1 s).  Here one CTA = 32 streams = 8 map warps (lane = frame, four
// streams each) + 2 recurrence warps (lane = stream), software-pipelined over 32-frame tiles with one __syncthreads per tile;
// at iteration i:
//   map warps    powers of tile i (inputs fetched one iteration ahead, in registers)        -> P
//   warp R1      the two envelope followers over tile i - 1, frame after frame              P -> E
//   map warps    detector of tile i - 2 (Red::detect, slider-rate row of the frame)         E -> D
//   warp R2      the six followers over tile i - 3 (Red::reduction_db)                      D -> G
//   map warps    gain of tile i - 4 (Red::gain), inputs of that tile re-read (L2), six output rows stored.
// All five are the functions Red::frame is made of, so outputs and state rows are bit-identical to the generic kernel's.
namespace red {
constexpr int kMapWarps = 8, kThreads = (kMapWarps + 2) * 32, kTP = 33, kPlane = 32 * kTP, kLag = 4;
struct Smem { float P[2][2][kPlane], E[2][2][kPlane], D[2][3][kPlane], G[2][kPlane]; };
}  // namespace red

__global__ void __launch_bounds__(red::kThreads) red_tiled_kernel(const FaustArgs a, const faust::Red::Konst k)
{
    using L = faust::Red;
    using namespace red;
    extern __shared__ __align__(16) unsigned char red_smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(red_smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, s0 = blockIdx.x * 32;
    const size_t ns = (size_t)a.n_streams;
    const int n_cta = min(32, a.n_streams - s0);
    const L::Slow* rows = reinterpret_cast<const L::Slow*>(a.slow);
    const int64_t n = a.n_frames;
    const int n_tiles = (int)((n + 31) / 32);
    // State words (faust::Red::State): wet_env2 0, ref_env2 1 (warp R1); offA_s 2, tgt_db 3, dryA_s 4, tgt_hold_s 5, gr_norm 6, gr_fast 7 (R2)
    if (warp < kMapWarps) {
        float nx[4][4];                                   // wetL, wetR, refL, refR of the next tile, per stream of this warp
        auto fetch = [&](int t) {
            const int64_t nn = (int64_t)t * 32 + lane;
            const bool ok = t < n_tiles && nn < n;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int si = warp * 4 + q;
                const float* x = a.in + (size_t)(s0 + si) * 6 * a.row_stride + nn;
                const bool go = ok && si < n_cta;
                nx[q][0] = go ? x[0] : 0.0f; nx[q][1] = go ? x[a.row_stride] : 0.0f;
                nx[q][2] = go ? x[4 * a.row_stride] : 0.0f; nx[q][3] = go ? x[5 * a.row_stride] : 0.0f;
            }
        };
        fetch(0);
        L::Slow cq[4];
        bool was_ragged = false;
        for (int i = 0; i < n_tiles + kLag; ++i) {
            if (i < n_tiles) {                           // powers of tile i
                const int o = lane;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int si = warp * 4 + q;
                    float wp, rp;
                    L::powers(nx[q][0], nx[q][1], nx[q][2], nx[q][3], wp, rp);
                    sm.P[i & 1][0][si * kTP + o] = wp; sm.P[i & 1][1][si * kTP + o] = rp;
                }
            }
            fetch(i + 1);                                // in == out is legal: tile i + 1 is stored five iterations from now
            if (i >= 2 && i - 2 < n_tiles) {             // detector of tile i - 2
                const int t = i - 2;
                const int64_t f0 = (int64_t)t * 32, nn = f0 + lane;
                const int64_t r0 = f0 % a.block_len;
                int64_t blk = f0 / a.block_len;
                if (r0 + 32 > a.block_len) blk = nn / a.block_len;
                const size_t row = (size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns;
                const bool ragged = r0 + 32 > a.block_len;                            // the tile spans host blocks: row per lane
                const bool reread = t == 0 || r0 == 0 || ragged || was_ragged;        // else the row can only change where a host block starts
                was_ragged = ragged;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int si = warp * 4 + q;
                    if (si < n_cta) {
                        if (reread) cq[q] = rows[row + s0 + si];
                        const L::Det d = L::detect(cq[q], k, sm.E[t & 1][0][si * kTP + lane], sm.E[t & 1][1][si * kTP + lane]);
                        sm.D[t & 1][0][si * kTP + lane] = d.dryA; sm.D[t & 1][1][si * kTP + lane] = d.offA; sm.D[t & 1][2][si * kTP + lane] = d.tgt0;
                    }
                }
            }
            if (i >= kLag) {                             // gain and outputs of tile i - 4
                const int t = i - kLag;
                const int64_t nn = (int64_t)t * 32 + lane;
                if (nn < n) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int si = warp * 4 + q;
                        if (si < n_cta) {
                            const float* x = a.in + (size_t)(s0 + si) * 6 * a.row_stride + nn;
                            float* y = a.out + (size_t)(s0 + si) * 6 * a.row_stride + nn;
                            float v[6];
#pragma unroll
                            for (int ch = 0; ch < 6; ++ch) v[ch] = x[ch * a.row_stride];
                            const float g = L::gain(sm.G[t & 1][si * kTP + lane]);
                            y[0] = v[0] * g; y[a.row_stride] = v[1] * g;
#pragma unroll
                            for (int ch = 2; ch < 6; ++ch) y[ch * a.row_stride] = v[ch];
                        }
                    }
                }
            }
            __syncthreads();
        }
    } else {
        const bool r1 = warp == kMapWarps;               // R1: envelopes, one tile behind the powers; R2: followers, three behind
        const int lag = r1 ? 1 : 3;
        const bool live = lane < n_cta;
        const size_t s = (size_t)(s0 + (live ? lane : 0));
        L::State m{};
        if (r1) { m.wet_env2 = __uint_as_float(a.state[0 * ns + s]); m.ref_env2 = __uint_as_float(a.state[1 * ns + s]); }
        else {
            m.offA_s = __uint_as_float(a.state[2 * ns + s]); m.tgt_db = __uint_as_float(a.state[3 * ns + s]); m.dryA_s = __uint_as_float(a.state[4 * ns + s]);
            m.tgt_hold_s = __uint_as_float(a.state[5 * ns + s]); m.gr_norm = __uint_as_float(a.state[6 * ns + s]); m.gr_fast = __uint_as_float(a.state[7 * ns + s]);
        }
        L::Slow c = rows[s];
        int blk = 0;
        int64_t blk_end = a.block_len;
        for (int i = 0; i < n_tiles + kLag; ++i) {
            const int t = i - lag;
            if (live && t >= 0 && t < n_tiles) {
                const int64_t f0 = (int64_t)t * 32;
                const int nv = (int)min((int64_t)32, n - f0);
                const int o = lane * kTP;
                if (r1) {
#pragma unroll 8
                    for (int f = 0; f < nv; ++f) {
                        L::envelopes(m, k, sm.P[t & 1][0][o + f], sm.P[t & 1][1][o + f]);
                        sm.E[t & 1][0][o + f] = m.wet_env2; sm.E[t & 1][1][o + f] = m.ref_env2;
                    }
                } else {
#pragma unroll 4
                    for (int f = 0; f < nv; ++f) {
                        if (f0 + f >= blk_end) {         // a new host block: its slider-rate row (the last row persists)
                            const int64_t b = (f0 + f) / a.block_len;
                            blk = (int)b; blk_end = (b + 1) * a.block_len;
                            c = rows[(size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns + s];
                        }
                        const L::Det d{sm.D[t & 1][0][o + f], sm.D[t & 1][1][o + f], sm.D[t & 1][2][o + f]};
                        sm.G[t & 1][o + f] = L::reduction_db(m, c, k, d);
                    }
                }
            }
            __syncthreads();
        }
        if (live) {
            if (r1) { a.state[0 * ns + s] = __float_as_uint(m.wet_env2); a.state[1 * ns + s] = __float_as_uint(m.ref_env2); }
            else {
                a.state[2 * ns + s] = __float_as_uint(m.offA_s); a.state[3 * ns + s] = __float_as_uint(m.tgt_db); a.state[4 * ns + s] = __float_as_uint(m.dryA_s);
                a.state[5 * ns + s] = __float_as_uint(m.tgt_hold_s); a.state[6 * ns + s] = __float_as_uint(m.gr_norm); a.state[7 * ns + s] = __float_as_uint(m.gr_fast);
            }
        }
    }
}
static_assert(sizeof(faust::Red::State) == 8 * 4, "red_tiled_kernel addresses the state words of faust::Red::State by number");

// ---- VAR, tiled ---------------------------------------------------------------------------------------------------------------
// VAR's frame: a detector chain (two biquads, five one-poles, a 3-tap Laplacian), memoryless control values (powf, three
// divisions), a switching envelope, a high-pass biquad pair per channel plus a noise band, and a memoryless mix (powf, two
// divisions).  One CTA = 16 streams (so that two CTAs share an SM) = 4 map warps (lane = frame, four streams each) + 3
// recurrence warps (lane = stream), 32-frame tiles, one __syncthreads per tile; at iteration i:
//   map   inputs of tile i (fetched one iteration ahead, in registers) -> IN               R1  Var::detector over tile i - 1   IN -> A
//   map   Var::control of tile i - 2 (slider-rate row of the frame)     A -> Bc, Bq         R2  Var::envelope over tile i - 3   Bc -> C
//   map   Var::mix of tile i - 4, inputs re-read (L2), outputs stored   Bq, C, D -> out     R3  Var::bands over tile i - 3      IN -> D
// The five functions are what Var::frame is made of: outputs and the 27 state words are bit-identical to the generic kernel's.
namespace var {
constexpr int kStreams = 16, kMapWarps = 4, kThreads = (kMapWarps + 3) * 32, kTP = 33, kPlane = kStreams * kTP, kLag = 4;
struct Smem { float IN[4][2][kPlane], A[2][7][kPlane], Bc[2][kPlane], Bq[3][4][kPlane], C[2][kPlane], D[2][3][kPlane]; };
}  // namespace var

__global__ void __launch_bounds__(var::kThreads, 2) var_tiled_kernel(const FaustArgs a, const faust::Var::Konst k)
{
    using L = faust::Var;
    using namespace var;
    extern __shared__ __align__(16) unsigned char var_smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(var_smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, s0 = blockIdx.x * kStreams;
    const size_t ns = (size_t)a.n_streams;
    const int n_cta = min(kStreams, a.n_streams - s0);
    const L::Slow* rows = reinterpret_cast<const L::Slow*>(a.slow);
    const int64_t n = a.n_frames;
    const int n_tiles = (int)((n + 31) / 32);
    auto ldw = [&](int w, size_t s) { return __uint_as_float(a.state[(size_t)w * ns + s]); };
    auto stw = [&](int w, size_t s, float v) { a.state[(size_t)w * ns + s] = __float_as_uint(v); };
    if (warp < kMapWarps) {
        float nx[4][2];
        auto fetch = [&](int t) {
            const int64_t nn = (int64_t)t * 32 + lane;
            const bool ok = t < n_tiles && nn < n;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int si = warp * 4 + q;
                const float* x = a.in + (size_t)(s0 + si) * 2 * a.row_stride + nn;
                const bool go = ok && si < n_cta;
                nx[q][0] = go ? x[0] : 0.0f; nx[q][1] = go ? x[a.row_stride] : 0.0f;
            }
        };
        fetch(0);
        L::Slow cq[4];
        bool was_ragged = false;
        for (int i = 0; i < n_tiles + kLag; ++i) {
            if (i < n_tiles) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int o = (warp * 4 + q) * kTP + lane;
                    sm.IN[i & 3][0][o] = nx[q][0]; sm.IN[i & 3][1][o] = nx[q][1];
                }
            }
            fetch(i + 1);                                // in == out is legal: tile i + 1 is stored five iterations from now
            if (i >= 2 && i - 2 < n_tiles) {             // control values of tile i - 2
                const int t = i - 2;
                const int64_t f0 = (int64_t)t * 32, nn = f0 + lane;
                const int64_t r0 = f0 % a.block_len;
                int64_t blk = f0 / a.block_len;
                const bool ragged = r0 + 32 > a.block_len;                            // the tile spans host blocks: row per lane
                if (ragged) blk = nn / a.block_len;
                const size_t row = (size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns;
                const bool reread = t == 0 || r0 == 0 || ragged || was_ragged;
                was_ragged = ragged;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int si = warp * 4 + q, o = si * kTP + lane;
                    if (si < n_cta) {
                        if (reread) cq[q] = rows[row + s0 + si];
                        const float (*A)[kPlane] = sm.A[t & 1];
                        const L::DetOut d{A[0][o], A[1][o], A[2][o], A[3][o], A[4][o], A[5][o], A[6][o]};
                        const L::Ctl c = L::control(cq[q], d);
                        sm.Bc[t & 1][o] = c.curvN;
                        float (*B)[kPlane] = sm.Bq[t % 3];
                        B[0][o] = c.thrN; B[1][o] = c.gate; B[2][o] = c.mE1; B[3][o] = c.airMix;
                    }
                }
            }
            if (i >= kLag) {                             // mix and outputs of tile i - 4
                const int t = i - kLag;
                const int64_t nn = (int64_t)t * 32 + lane;
                if (nn < n) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int si = warp * 4 + q, o = si * kTP + lane;
                        if (si < n_cta) {
                            const float* x = a.in + (size_t)(s0 + si) * 2 * a.row_stride + nn;
                            float* y = a.out + (size_t)(s0 + si) * 2 * a.row_stride + nn;
                            const float inL = x[0], inR = x[a.row_stride];
                            const float (*B)[kPlane] = sm.Bq[t % 3];
                            const L::Ctl c{0.0f, B[0][o], B[1][o], B[2][o], B[3][o]};
                            const L::Hf h{sm.D[t & 1][0][o], sm.D[t & 1][1][o], sm.D[t & 1][2][o]};
                            float ov[2];
                            L::mix(k, c, sm.C[t & 1][o], h, inL, inR, ov);
                            y[0] = ov[0]; y[a.row_stride] = ov[1];
                        }
                    }
                }
            }
            __syncthreads();
        }
    } else {
        const int role = warp - kMapWarps;               // 0: detector (R1), 1: envelope (R2), 2: bands (R3)
        const int lag = role == 0 ? 1 : 3;
        const bool live = lane < n_cta;
        const size_t s = (size_t)(s0 + (live ? lane : 0));
        // State words (faust::Var::State): amount 0, sens 1, detL 2-3, detR 4-5, hf1L 6-7, hf2L 8-9, hf1R 10-11, hf2R 12-13, airZ 14-15,
        // hfLvl 16, sm1L 17, sm2L 18, sm1R 19, sm2R 20, s1L 21, s2L 22, s1R 23, s2R 24, env 25, noise 26
        L::State m{};
        if (role == 0) {
            m.amount = ldw(0, s); m.sens = ldw(1, s); m.detL = {ldw(2, s), ldw(3, s)}; m.detR = {ldw(4, s), ldw(5, s)};
            m.hfLvl = ldw(16, s); m.sm1L = ldw(17, s); m.sm2L = ldw(18, s); m.sm1R = ldw(19, s); m.sm2R = ldw(20, s);
            m.s1L = ldw(21, s); m.s2L = ldw(22, s); m.s1R = ldw(23, s); m.s2R = ldw(24, s);
        } else if (role == 1) {
            m.env = ldw(25, s);
        } else {
            m.hf1L = {ldw(6, s), ldw(7, s)}; m.hf2L = {ldw(8, s), ldw(9, s)}; m.hf1R = {ldw(10, s), ldw(11, s)}; m.hf2R = {ldw(12, s), ldw(13, s)};
            m.airZ = {ldw(14, s), ldw(15, s)}; m.noise = (int)a.state[26 * ns + s];
        }
        L::Slow c = rows[s];
        int64_t blk_end = a.block_len;
        for (int i = 0; i < n_tiles + kLag; ++i) {
            const int t = i - lag;
            if (live && t >= 0 && t < n_tiles) {
                const int64_t f0 = (int64_t)t * 32;
                const int nv = (int)min((int64_t)32, n - f0);
                const int o = lane * kTP;
                if (role == 0) {
#pragma unroll 4
                    for (int f = 0; f < nv; ++f) {
                        if (f0 + f >= blk_end) {         // a new host block: its slider-rate row (the last row persists)
                            const int64_t b = (f0 + f) / a.block_len;
                            blk_end = (b + 1) * a.block_len;
                            c = rows[(size_t)(b < a.n_rows ? b : a.n_rows - 1) * ns + s];
                        }
                        const L::DetOut d = L::detector(m, c, k, sm.IN[t & 3][0][o + f], sm.IN[t & 3][1][o + f]);
                        float (*A)[kPlane] = sm.A[t & 1];
                        A[0][o + f] = d.amount; A[1][o + f] = d.sens; A[2][o + f] = d.hfLvl; A[3][o + f] = d.lapL; A[4][o + f] = d.denomL;
                        A[5][o + f] = d.lapR; A[6][o + f] = d.denomR;
                    }
                } else if (role == 1) {
#pragma unroll 8
                    for (int f = 0; f < nv; ++f) sm.C[t & 1][o + f] = L::envelope(m, k, sm.Bc[t & 1][o + f]);
                } else {
#pragma unroll 4
                    for (int f = 0; f < nv; ++f) {
                        const L::Hf h = L::bands(m, k, sm.IN[t & 3][0][o + f], sm.IN[t & 3][1][o + f]);
                        sm.D[t & 1][0][o + f] = h.hfL; sm.D[t & 1][1][o + f] = h.hfR; sm.D[t & 1][2][o + f] = h.air;
                    }
                }
            }
            __syncthreads();
        }
        if (live) {
            if (role == 0) {
                stw(0, s, m.amount); stw(1, s, m.sens); stw(2, s, m.detL.s1); stw(3, s, m.detL.s2); stw(4, s, m.detR.s1); stw(5, s, m.detR.s2);
                stw(16, s, m.hfLvl); stw(17, s, m.sm1L); stw(18, s, m.sm2L); stw(19, s, m.sm1R); stw(20, s, m.sm2R);
                stw(21, s, m.s1L); stw(22, s, m.s2L); stw(23, s, m.s1R); stw(24, s, m.s2R);
            } else if (role == 1) {
                stw(25, s, m.env);
            } else {
                stw(6, s, m.hf1L.s1); stw(7, s, m.hf1L.s2); stw(8, s, m.hf2L.s1); stw(9, s, m.hf2L.s2); stw(10, s, m.hf1R.s1); stw(11, s, m.hf1R.s2);
                stw(12, s, m.hf2R.s1); stw(13, s, m.hf2R.s2); stw(14, s, m.airZ.s1); stw(15, s, m.airZ.s2); a.state[26 * ns + s] = (unsigned)m.noise;
            }
        }
    }
}
static_assert(sizeof(faust::Var::State) == 27 * 4 && offsetof(faust::Var::State, hfLvl) == 16 * 4 && offsetof(faust::Var::State, env) == 25 * 4,
              "var_tiled_kernel addresses the state words of faust::Var::State by number");

template <class L> struct TiledLaunch { static int go(const FaustArgs&, const typename L::Konst&, cudaStream_t) { return 0; } };        // 0 = no tiled kernel for this leaf
template <> struct TiledLaunch<faust::Var> {
    static int go(const FaustArgs& a, const faust::Var::Konst& k, cudaStream_t st)
    {
        // 16 streams per CTA, two CTAs per SM: the time grows with every 4736 streams, while the generic kernel (eight one-warp
        // CTAs per SM) stays flat up to 37 888.  Measured per 0.5 s, tiled / generic: 9472 streams 6.0 / 15.5 ms, 18 944: 12.0 /
        // 15.5, 37 888: 24.0 / 17.3 -- the generic kernel takes over at ~160 streams per SM.  (RED and ClickBeGoneSG tiled stay
        // 1.8-5x ahead up to 37 888 streams.)
        if (!a.force_tiled) {
            int dev = 0, n_sm = 148;
            if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
            if (a.n_streams > 160 * n_sm) return 0;
        }
        static DeviceOnce attr_once;
        if (attr_once.pending()) {
            if (!cuda_ok(cudaFuncSetAttribute(var_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(var::Smem)), "VAR smem attr")) return -1;
            attr_once.mark();
        }
        var_tiled_kernel<<<(unsigned)((a.n_streams + var::kStreams - 1) / var::kStreams), var::kThreads, sizeof(var::Smem), st>>>(a, k);
        return cuda_ok(cudaGetLastError(), "var_tiled_kernel") ? 1 : -1;
    }
};
template <> struct TiledLaunch<faust::Red> {
    static int go(const FaustArgs& a, const faust::Red::Konst& k, cudaStream_t st)
    {
        static DeviceOnce attr_once;
        if (attr_once.pending()) {
            if (!cuda_ok(cudaFuncSetAttribute(red_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(red::Smem)), "RED smem attr")) return -1;
            attr_once.mark();
        }
        red_tiled_kernel<<<(unsigned)((a.n_streams + 31) / 32), red::kThreads, sizeof(red::Smem), st>>>(a, k);
        return cuda_ok(cudaGetLastError(), "red_tiled_kernel") ? 1 : -1;
    }
};
template <> struct TiledLaunch<faust::Cbg> {
    static int go(const FaustArgs& a, const faust::Cbg::Konst&, cudaStream_t st)
    {
        static DeviceOnce attr_once;
        if (attr_once.pending()) {
            if (!cuda_ok(cudaFuncSetAttribute(cbg_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(cbg::Smem)), "ClickBeGoneSG smem attr")) return -1;
            attr_once.mark();
        }
        cbg_tiled_kernel<<<(unsigned)((a.n_streams + 31) / 32), cbg::kThreads, sizeof(cbg::Smem), st>>>(a);
        return cuda_ok(cudaGetLastError(), "cbg_tiled_kernel") ? 1 : -1;
    }
};

template <class L>
class FaustLeaf : public PluginImpl {
    typename L::Konst k_{};
    unsigned* d_state = nullptr;
    typename L::Slow* d_slow = nullptr;
    size_t slow_cap = 0;
    std::vector<typename L::Slow> h_slow, h_slow_dev;
    bool slow_valid = false;
    float fs = 48000.0f;
    static constexpr int kWords = (int)(sizeof(typename L::State) / 4);

public:
    ~FaustLeaf() override { cudaFree(d_state); cudaFree(d_slow); }

    bool create(Instance& I) override
    {
        if (I.precision != ZA_F32) {
            set_error(std::string(L::key()) + ": Faust leaves compute in float32 (no -double in the reference build); create it with ZA_F32");
            I.create_status = ZA_ERR_UNSUPPORTED;
            return false;
        }
        if ((int)I.decls.size() != L::kSliders) { set_error(std::string(L::key()) + ": slider layout mismatch"); return false; }
        fs = std::fmin(192000.0f, std::fmax(1.0f, (float)(int)I.srate));      // ma.SR
        L::konst(fs, k_);
        return cuda_ok(cudaMalloc(&d_state, (size_t)kWords * I.n_streams * 4), "Faust leaf state");
    }
    bool reset(Instance& I) override    // mydsp::init -> instanceClear: every recursive signal starts at 0
    {
        slow_valid = false;
        return cuda_ok(cudaMemsetAsync(d_state, 0, (size_t)kWords * I.n_streams * 4, I.stream), "Faust leaf reset");
    }
    bool sliders_changed(Instance&, int, int, const std::vector<int>&) override { return true; }
    bool takes_block_params() const override { return true; }
    bool run(Instance& I, int s0, int ns, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        return run_blocks(I, s0, ns, &I.sliders[(size_t)s0 * I.decls.size()], 1, in, out, tap, n, stride);
    }
    bool run_blocks(Instance& I, int s0, int ns, const double* q, int nbq, const float* in, float* out, double* tap, int64_t n,
                    int64_t stride) override
    {
        if (s0 != 0 || ns != I.n_streams) { set_error(std::string(L::key()) + ": stream sub-ranges are not supported"); return false; }
        if (tap) { set_error(std::string(L::key()) + ": no double tap (float32 leaf)"); return false; }
        constexpr int nd = L::kSliders;
        h_slow.resize((size_t)nbq * ns);
        parallel_for(ns, 64, [&](int64_t sb, int64_t se) {
            for (int b = 0; b < nbq; ++b)
                for (int64_t s = sb; s < se; ++s) {
                    const double* sl = q + ((size_t)b * ns + s) * nd;
                    typename L::Slow& c = h_slow[(size_t)b * ns + s];
                    if (b > 0 && std::memcmp(sl, sl - (size_t)ns * nd, sizeof(double) * nd) == 0) { c = h_slow[(size_t)(b - 1) * ns + s]; continue; }
                    float z[nd];
                    for (int j = 0; j < nd; ++j) z[j] = (float)sl[j];              // *zone = (FAUSTFLOAT) value
                    L::slow(z, fs, c);
                }
        });
        const bool same = slow_valid && h_slow.size() == h_slow_dev.size() &&
                          std::memcmp(h_slow.data(), h_slow_dev.data(), h_slow.size() * sizeof(typename L::Slow)) == 0;
        if (!same) {
            slow_valid = false;
            if (h_slow.size() > slow_cap) {
                cudaFree(d_slow); d_slow = nullptr; slow_cap = 0;
                if (!cuda_ok(cudaMalloc(&d_slow, h_slow.size() * sizeof(typename L::Slow)), "Faust leaf rows")) return false;
                slow_cap = h_slow.size();
            }
            if (!cuda_ok(cudaMemcpyAsync(d_slow, h_slow.data(), h_slow.size() * sizeof(typename L::Slow), cudaMemcpyHostToDevice, I.stream), "Faust leaf rows upload"))
                return false;
            slow_valid = h_slow.size() <= (1u << 20);
            if (slow_valid) h_slow_dev = h_slow;
        }
        if (n <= 0) return true;
        FaustArgs a{in, out, stride, n, I.n_streams, I.block_len, nbq, d_slow, d_state, I.opt_lanes, I.opt_scalar_f32 == 2};
        const bool timed = I.time_begin();
        const int tiled = I.opt_scalar_f32 == 1 ? 0 : TiledLaunch<L>::go(a, k_, I.stream);      // option scalar_f32 = 1: the generic kernel for every leaf, 2: the tiled one
        if (tiled < 0) return false;
        if (!tiled) {
            if (!launch_generic<L>(a, k_, I.stream)) return false;
        }
        if (timed) I.time_end();
        ++I.launches;
        return true;
    }
    void state_io(Instance& I, StateIO& io) override { io.dev(d_state, (size_t)kWords * I.n_streams * 4); }
    bool after_state_load(Instance&) override { slow_valid = false; return true; }
    bool read_state(Instance& I, const char* name, double* out) override     // "state[k]": word k of the leaf's State as float
    {
        int k = -1;
        if (!std::strncmp(name, "state[", 6)) k = std::atoi(name + 6);
        if (k < 0 || k >= kWords) { set_error(std::string(L::key()) + ": unknown state variable " + name); return false; }
        std::vector<float> t((size_t)I.n_streams);
        if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync") ||
            !cuda_ok(cudaMemcpy(t.data(), d_state + (size_t)k * I.n_streams, t.size() * 4, cudaMemcpyDeviceToHost), "state read")) return false;
        for (int s = 0; s < I.n_streams; ++s) out[s] = (double)t[s];
        return true;
    }
};

}  // namespace

PluginImpl* make_red() { return new FaustLeaf<faust::Red>(); }
PluginImpl* make_clickbegonesg() { return new FaustLeaf<faust::Cbg>(); }
PluginImpl* make_var() { return new FaustLeaf<faust::Var>(); }
const char* red_slider_text() { return faust::Red::slider_text(); }
const char* clickbegonesg_slider_text() { return faust::Cbg::slider_text(); }
const char* var_slider_text() { return faust::Var::slider_text(); }

}  // namespace za
