// k_stft.cu -- STFT class: plugin "stft_mask2048" (BASELINE config 4) on sm_100a.
//
// Replaces the generated @sample loop of plugins/Workloads/stft_mask2048/src/stft_mask2048.jsfx (structure of
// "DECRUST", plugins/Spectral/Sample/src/Sample.jsfx:3503-3751) and the FFT runtime hooks it calls
// (jsfx_fft / jsfx_fft_permute / jsfx_fft_ipermute / jsfx_ifft, src/JSFXJuceProcessor.cpp:1267-1447 over
// src/WDL/fft.c) for n_streams instances at once.
//
// Mapping.  128 threads per stream, NS streams per CTA (4 in fp32, 2 in fp64: one CTA per SM, the twiddle and
// window tables are shared by the CTA's streams), persistent over the call's time axis.  Everything between two
// hops lives on chip: the last four hops of input (analysis window + the 2048-frame dry delay), four hops of
// overlap-add accumulator and the FFT exchange buffer in shared memory; the per-bin recursive state (baseline
// EMA, previous magnitude) in registers of the thread that owns the bin.  Per hop (stft_core.h):
//   (1) 512 stereo frames in / out as float4 per thread, next hop's input prefetched during the transforms;
//   (2) window both channels into ONE complex signal z = w (l + i r), forward FFT 16 x 16 x 8 (registers,
//       two shared-memory exchanges, padded so every access pattern is bank-conflict free);
//   (3) stereo split by conjugate symmetry inside one thread's registers, magnitudes, per-bin gain rule with
//       its recursive state, 3-tap smoothing (magnitudes and gains cross threads through shared memory),
//       gains applied to Z[k] and Z[N-k] (a real, symmetric gain commutes with the stereo packing);
//   (4) inverse FFT 8 x 16 x 16, x 1/N x synthesis window, overlap-add.
// One complex FFT pair per stereo hop instead of the reference's two: same mathematics, rounding differs at
// the 1e-15 level (criterion 1e-12).  Window / twiddle tables are computed on the host with glibc.
#include "stft_core.h"
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

using namespace stft;

static_assert(sizeof(Coef) == sizeof(StftCoef), "Coef layout");

size_t stft_state_bytes() { return sizeof(StftState); }

namespace {

__device__ __forceinline__ void bar_stream(int sub) { asm volatile("bar.sync %0, 128;" ::"r"(sub + 1) : "memory"); }

template <typename R> __device__ __forceinline__ void ld4r(const R* p, R (&x)[4]);
template <> __device__ __forceinline__ void ld4r<float>(const float* p, float (&x)[4])
{
    const float4 v = *reinterpret_cast<const float4*>(p);
    x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
}
template <> __device__ __forceinline__ void ld4r<double>(const double* p, double (&x)[4])
{
    const double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
    x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y;
}
template <typename R> __device__ __forceinline__ void zero4(R* p);
template <> __device__ __forceinline__ void zero4<float>(float* p) { *reinterpret_cast<float4*>(p) = make_float4(0.f, 0.f, 0.f, 0.f); }
template <> __device__ __forceinline__ void zero4<double>(double* p)
{
    *reinterpret_cast<double2*>(p) = make_double2(0.0, 0.0);
    *reinterpret_cast<double2*>(p + 2) = make_double2(0.0, 0.0);
}

}  // namespace

template <typename R, int NS, bool TAP>
__global__ void __launch_bounds__(NS * kT, 1)
stft_mask_kernel(const StftArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    Tables<R>& tb = *reinterpret_cast<Tables<R>*>(smem_raw);
    const int sub = threadIdx.x >> 7, t = threadIdx.x & (kT - 1);
    StreamSmem<R>& S = reinterpret_cast<StreamSmem<R>*>(smem_raw + sizeof(Tables<R>))[sub];

    // ---- tables -> shared memory (whole CTA) ------------------------------------------------------------
    for (int i = threadIdx.x; i < 15 * kT; i += NS * kT) { tb.tw1[i / kT][i % kT].re = (R)a.tw1[2 * i]; tb.tw1[i / kT][i % kT].im = (R)a.tw1[2 * i + 1]; }
    for (int i = threadIdx.x; i < 15 * 8; i += NS * kT) { tb.tw2[i / 8][i % 8].re = (R)a.tw2[2 * i]; tb.tw2[i / 8][i % 8].im = (R)a.tw2[2 * i + 1]; }
    for (int i = threadIdx.x; i < kN; i += NS * kT) { tb.win[i] = (R)a.win[i]; tb.syn[i] = (R)a.syn[i]; }
    __syncthreads();
    const int stream = blockIdx.x * NS + sub;
    if (stream >= a.n_streams) return;
    StftState& gs = a.state[stream];

    // ---- per-stream state -> shared memory / registers ---------------------------------------------------
    Regs<R> g;
    for (int i = t; i < 2 * kN; i += kT) { S.acc[i / kN][i % kN] = (R)gs.acc[i / kN][i % kN]; S.xin[i / kN][i % kN] = gs.xin[i / kN][i % kN]; }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int k = slot_bin_rt(t, s);
        g.base[s] = k >= 0 ? (R)gs.base[k] : (R)0;
        g.prev[s] = k >= 0 ? (R)gs.prev[k] : (R)0;
        g.magr[s] = (R)0;
    }
    bar_stream(sub);

    const float* inL = a.in + (size_t)stream * 2 * a.row_stride;
    const float* inR = inL + a.row_stride;
    float* outL = a.out + (size_t)stream * 2 * a.row_stride;
    float* outR = outL + a.row_stride;
    const bool aligned = ((a.row_stride & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.in) & 15) == 0) &&
                         ((reinterpret_cast<uintptr_t>(a.out) & 15) == 0) &&
                         (!TAP || (reinterpret_cast<uintptr_t>(a.tap) & 15) == 0);

    int64_t pos = 0;
    int64_t hop = a.t0 / kHop;                 // absolute hop index of the frame position
    int off = (int)(a.t0 % kHop);              // offset inside the hop (= the script's hop_ctr)
    float4 nl = make_float4(0.f, 0.f, 0.f, 0.f), nr = nl;
    bool have_next = false;                    // nl / nr hold frames pos + 4t .. of a full, aligned hop
    if (aligned && off == 0 && kHop <= a.n_frames) {
        nl = *reinterpret_cast<const float4*>(inL + 4 * t); nr = *reinterpret_cast<const float4*>(inR + 4 * t);
        have_next = true;
    }
    while (pos < a.n_frames) {
        // segment = up to the next hop boundary, the next host-block boundary, or the end of the call
        const int64_t blk = pos / a.block_len;
        int64_t end = (blk + 1) * a.block_len;
        if (end > a.n_frames) end = a.n_frames;
        if (end - pos > kHop - off) end = pos + (kHop - off);
        const int len = (int)(end - pos);
        const StftCoef c = a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * a.n_streams + stream];
        const R mixv = (R)c.mixv;
        const int slot = (int)(hop & 3) * kHop;     // input slot of this hop = dry slot (hop - 4) = accumulator slot

        // ---- (1) stream the segment: dry read, input ring write, accumulator read + clear, dry/wet mix ------
        if (have_next && len == kHop) {
            const int o = slot + 4 * t;
            const float4 dl = *reinterpret_cast<const float4*>(&S.xin[0][o]), dr = *reinterpret_cast<const float4*>(&S.xin[1][o]);
            *reinterpret_cast<float4*>(&S.xin[0][o]) = nl; *reinterpret_cast<float4*>(&S.xin[1][o]) = nr;
            R rl[4], rr[4];
            ld4r<R>(&S.acc[0][o], rl); ld4r<R>(&S.acc[1][o], rr);
            zero4<R>(&S.acc[0][o]); zero4<R>(&S.acc[1][o]);
            const R d0[4] = {(R)dl.x, (R)dl.y, (R)dl.z, (R)dl.w}, d1[4] = {(R)dr.x, (R)dr.y, (R)dr.z, (R)dr.w};
            R o0[4], o1[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { o0[i] = add_(d0[i], mul_(mixv, sub_(rl[i], d0[i]))); o1[i] = add_(d1[i], mul_(mixv, sub_(rr[i], d1[i]))); }
            *reinterpret_cast<float4*>(outL + pos + 4 * t) = make_float4((float)o0[0], (float)o0[1], (float)o0[2], (float)o0[3]);
            *reinterpret_cast<float4*>(outR + pos + 4 * t) = make_float4((float)o1[0], (float)o1[1], (float)o1[2], (float)o1[3]);
            if (TAP) {
                double* tp = a.tap + (size_t)stream * 2 * a.row_stride + pos + 4 * t;
                *reinterpret_cast<double2*>(tp) = make_double2((double)o0[0], (double)o0[1]);
                *reinterpret_cast<double2*>(tp + 2) = make_double2((double)o0[2], (double)o0[3]);
                *reinterpret_cast<double2*>(tp + a.row_stride) = make_double2((double)o1[0], (double)o1[1]);
                *reinterpret_cast<double2*>(tp + a.row_stride + 2) = make_double2((double)o1[2], (double)o1[3]);
            }
        } else {
            for (int i = t; i < len; i += kT) {
                const int o = slot + off + i;
                const float l = inL[pos + i], r = inR[pos + i];
                const R dryL = (R)S.xin[0][o], dryR = (R)S.xin[1][o];
                S.xin[0][o] = l; S.xin[1][o] = r;
                const R rawL = S.acc[0][o], rawR = S.acc[1][o];
                S.acc[0][o] = (R)0; S.acc[1][o] = (R)0;
                const R o0 = add_(dryL, mul_(mixv, sub_(rawL, dryL))), o1 = add_(dryR, mul_(mixv, sub_(rawR, dryR)));
                outL[pos + i] = (float)o0; outR[pos + i] = (float)o1;
                if (TAP) {
                    double* tp = a.tap + (size_t)stream * 2 * a.row_stride + pos + i;
                    tp[0] = (double)o0; tp[a.row_stride] = (double)o1;
                }
            }
        }
        pos = end; off += len;
        // prefetch the next hop's input when the next segment is a full aligned hop inside one host block
        have_next = false;
        if (aligned && off == kHop && (pos & 3) == 0 && pos + kHop <= a.n_frames &&
            (pos / a.block_len) == ((pos + kHop - 1) / a.block_len)) {
            nl = *reinterpret_cast<const float4*>(inL + pos + 4 * t); nr = *reinterpret_cast<const float4*>(inR + pos + 4 * t);
            have_next = true;
        }
        bar_stream(sub);
        if (off < kHop) continue;
        off = 0;

        // ---- (2)-(4) frame `hop` -----------------------------------------------------------------------------
        Coef cc; cc.amt = c.amt; cc.thresh = c.thresh; cc.ab = c.ab; cc.bb = c.bb; cc.floor_g = c.floor_g; cc.mixv = c.mixv;
        fwd1<R>(t, g, S, tb, hop);
        bar_stream(sub);
        fwd2<R>(t, g, S, tb);
        bar_stream(sub);
        fwd3<R>(t, g, S);
        t0_permute<R>(t == 0, g);              // thread 0's self-paired blocks into the regular pair layout
        bar_stream(sub);                       // every Z read done: the buffer now carries magnitudes and gains
        mags_u<R>(t, g, S);
        bar_stream(sub);
        gains_u<R>(t, g, S, cc);
        bar_stream(sub);
        apply_gains_u<R>(t, g, S);
        t0_unpermute<R>(t == 0, g);
        bar_stream(sub);                       // every gain read done: the buffer carries Z again
        inv3<R>(t, g, S);
        bar_stream(sub);
        inv2<R>(t, g, S, tb);
        bar_stream(sub);
        inv1_ola<R>(t, g, S, tb, hop);
        ++hop;
        bar_stream(sub);
    }

    // ---- state -> HBM (carried into the next call) ------------------------------------------------------------
    for (int i = t; i < 2 * kN; i += kT) { gs.acc[i / kN][i % kN] = (double)S.acc[i / kN][i % kN]; gs.xin[i / kN][i % kN] = S.xin[i / kN][i % kN]; }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int k = slot_bin_rt(t, s);
        if (k >= 0) { gs.base[k] = (double)g.base[s]; gs.prev[k] = (double)g.prev[s]; }
    }
}

template <typename R, int NS, bool TAP>
static bool launch_stft(const StftArgs& a, cudaStream_t stream)
{
    auto kern = stft_mask_kernel<R, NS, TAP>;
    const size_t smem = sizeof(Tables<R>) + NS * sizeof(StreamSmem<R>);
    static DeviceOnce attr_once;
    if (attr_once.pending()) {
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "stft smem attr")) return false;
        attr_once.mark();
    }
    kern<<<(a.n_streams + NS - 1) / NS, NS * kT, smem, stream>>>(a);
    return cuda_ok(cudaGetLastError(), "stft_mask_kernel launch");
}

// streams per CTA: as many as fit one SM's shared memory, but no more than spreads n_streams over the SMs
bool stft_launch(const StftArgs& a, int precision, int n_sm, cudaStream_t stream)
{
    if (a.n_frames <= 0 || a.n_streams <= 0) return true;
    const int per_sm = (a.n_streams + n_sm - 1) / n_sm;
    if (precision) {
        if (a.tap) return per_sm >= 2 ? launch_stft<double, 2, true>(a, stream) : launch_stft<double, 1, true>(a, stream);
        return per_sm >= 2 ? launch_stft<double, 2, false>(a, stream) : launch_stft<double, 1, false>(a, stream);
    }
    if (per_sm >= 4) return launch_stft<float, 4, false>(a, stream);
    if (per_sm == 3) return launch_stft<float, 3, false>(a, stream);
    if (per_sm == 2) return launch_stft<float, 2, false>(a, stream);
    return launch_stft<float, 1, false>(a, stream);
}

}  // namespace za
