// k_stft.cu -- STFT class: plugin "stft_mask2048" (BASELINE config 4) on sm_100a.
//
// Replaces the generated @sample loop of plugins/Workloads/stft_mask2048/src/stft_mask2048.jsfx (structure of
// "DECRUST", plugins/Spectral/Sample/src/Sample.jsfx:3503-3751) and the FFT runtime hooks it calls
// (jsfx_fft / jsfx_fft_permute / jsfx_fft_ipermute / jsfx_ifft, src/JSFXJuceProcessor.cpp:1267-1447 over
// src/WDL/fft.c) for n_streams instances at once.
//
// Mapping.  One CTA per stream, persistent over the call's time axis; everything between two hops lives in
// shared memory: the last five hops of input (analysis window + dry delay), four hops of overlap-add
// accumulator, the per-bin recursive state (baseline EMA, previous magnitude), the FFT buffer and twiddles.
// Per hop: (1) stream 512 stereo frames in / out (out = finished accumulator slot, mixed with the delayed
// dry signal), (2) window both channels into ONE complex buffer z = w (l + i r), (3) in-place radix-2 DIF
// FFT in shared memory (natural in, bit-reversed out: the permutation the reference does with fft_permute is
// folded into the bin addressing), (4) split the stereo spectra by conjugate symmetry, magnitudes,
// per-bin gain rule with its recursive state, 3-tap smoothing, (5) scale Z[k] and Z[N-k] (a real,
// symmetric gain commutes with the stereo packing), (6) in-place radix-2 DIT inverse FFT (bit-reversed in,
// natural out), (7) x 1/N x synthesis window, overlap-add.  One complex FFT pair per stereo hop instead of
// the reference's two: same mathematics, rounding differs at the 1e-15 level (criterion 1e-12).
// Window / twiddle tables are computed on the host with glibc in the script's expression order.
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

namespace {

constexpr int kN = kStftN, kHop = kStftHop, kBins = kStftBins;

template <typename R> struct Cx { R re, im; };

template <typename R> __device__ __forceinline__ R sqrt_r(R v);
template <> __device__ __forceinline__ double sqrt_r<double>(double v) { return __dsqrt_rn(v); }
template <> __device__ __forceinline__ float sqrt_r<float>(float v) { return sqrtf(v); }
template <typename R> __device__ __forceinline__ R div_r(R a, R b);
template <> __device__ __forceinline__ double div_r<double>(double a, double b) { return __ddiv_rn(a, b); }
template <> __device__ __forceinline__ float div_r<float>(float a, float b) { return a / b; }
template <typename R> __device__ __forceinline__ R max_r(R a, R b) { return a > b ? a : b; }   // ogt select (dsp_jsfx_aot.py:5223-5232)
template <typename R> __device__ __forceinline__ R min_r(R a, R b) { return a < b ? a : b; }
template <typename R> __device__ __forceinline__ R clamp01(R x) { return x < (R)0 ? (R)0 : (x > (R)1 ? (R)1 : x); }

__device__ __forceinline__ int brev11(int k) { return (int)(__brev((unsigned)k) >> 21); }   // 2048 = 2^11

template <typename R, int NT> struct StftSmem {
    Cx<R> z[kN];               // FFT buffer
    Cx<R> tw[kN / 2];          // exp(-2 pi i k / N)
    R acc[2][kN];              // overlap-add accumulator: 4 slots of one hop per channel
    R mag[kBins + 3];
    R gain[kBins + 3];
    R base[kBins + 3];
    R prev[kBins + 3];
    float xin[2][5 * kHop];    // last five hops of input per channel (slot = hop index mod 5)
};

}  // namespace

size_t stft_state_bytes() { return sizeof(StftState); }

template <typename R, int NT, bool TAP>
__global__ void __launch_bounds__(NT)
stft_mask_kernel(const StftArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    StftSmem<R, NT>& sm = *reinterpret_cast<StftSmem<R, NT>*>(smem_raw);
    const int tid = threadIdx.x, stream = blockIdx.x;
    StftState& gs = a.state[stream];

    // ---- state and tables -> shared memory -----------------------------------------------------------
    for (int i = tid; i < kN / 2; i += NT) { sm.tw[i].re = (R)a.tw[2 * i]; sm.tw[i].im = (R)a.tw[2 * i + 1]; }
    for (int i = tid; i < 2 * kN; i += NT) sm.acc[i / kN][i % kN] = (R)gs.acc[i / kN][i % kN];
    for (int i = tid; i < 2 * 5 * kHop; i += NT) sm.xin[i / (5 * kHop)][i % (5 * kHop)] = gs.xin[i / (5 * kHop)][i % (5 * kHop)];
    for (int i = tid; i < kBins; i += NT) { sm.base[i] = (R)gs.base[i]; sm.prev[i] = (R)gs.prev[i]; }
    __syncthreads();

    const float* inL = a.in + (size_t)stream * 2 * a.row_stride;
    const float* inR = inL + a.row_stride;
    float* outL = a.out + (size_t)stream * 2 * a.row_stride;
    float* outR = outL + a.row_stride;

    int64_t pos = 0;
    int64_t hop = a.t0 / kHop;                 // absolute hop index of the frame position
    int off = (int)(a.t0 % kHop);              // offset inside the hop (= the script's hop_ctr)
    while (pos < a.n_frames) {
        // segment = up to the next hop boundary, the next host-block boundary, or the end of the call
        const int64_t blk = pos / a.block_len;
        int64_t end = (blk + 1) * a.block_len;
        if (end > a.n_frames) end = a.n_frames;
        if (end - pos > kHop - off) end = pos + (kHop - off);
        const int len = (int)(end - pos);
        const StftCoef c = a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * a.n_streams + stream];
        const R mixv = (R)c.mixv;
        const int s_in = (int)(hop % 5), s_dry = (int)((hop + 1) % 5), s_acc = (int)(hop % 4);   // (hop - 4) mod 5 = (hop + 1) mod 5

        // ---- (1) stream the segment: input ring write, accumulator read + clear, dry/wet mix ------------
        for (int i = tid; i < len; i += NT) {
            const int o = off + i;
            const float l = inL[pos + i], r = inR[pos + i];
            const R dryL = (R)sm.xin[0][s_dry * kHop + o], dryR = (R)sm.xin[1][s_dry * kHop + o];
            sm.xin[0][s_in * kHop + o] = l; sm.xin[1][s_in * kHop + o] = r;
            const R rawL = sm.acc[0][s_acc * kHop + o], rawR = sm.acc[1][s_acc * kHop + o];
            sm.acc[0][s_acc * kHop + o] = (R)0; sm.acc[1][s_acc * kHop + o] = (R)0;
            const R o0 = add(dryL, mul(mixv, sub(rawL, dryL))), o1 = add(dryR, mul(mixv, sub(rawR, dryR)));
            outL[pos + i] = (float)o0; outR[pos + i] = (float)o1;
            if (TAP) {
                double* tp = a.tap + (size_t)stream * 2 * a.row_stride + pos + i;
                tp[0] = (double)o0; tp[a.row_stride] = (double)o1;
            }
        }
        pos = end; off += len;
        __syncthreads();
        if (off < kHop) continue;
        off = 0;

        // ---- (2) frame `hop`: window hops hop-3 .. hop into z = w (l + i r) ---------------------------------
        for (int i = tid; i < kN; i += NT) {
            const int hs = (int)((hop + 2 + (i >> 9)) % 5);            // (hop - 3 + i/512) mod 5
            const R w = (R)a.win[i];
            sm.z[i].re = mul((R)sm.xin[0][hs * kHop + (i & 511)], w);
            sm.z[i].im = mul((R)sm.xin[1][hs * kHop + (i & 511)], w);
        }
        __syncthreads();
        // ---- (3) forward FFT, radix-2 decimation in frequency: natural in, bit-reversed out ----------------
        for (int half = kN / 2, tstep = 1; half >= 1; half >>= 1, tstep <<= 1) {
            for (int j = tid; j < kN / 2; j += NT) {
                const int p = j & (half - 1), i0 = ((j - p) << 1) + p, i1 = i0 + half;
                const Cx<R> u = sm.z[i0], v = sm.z[i1], w = sm.tw[p * tstep];
                const R dr = u.re - v.re, di = u.im - v.im;
                sm.z[i0].re = u.re + v.re; sm.z[i0].im = u.im + v.im;
                sm.z[i1].re = dr * w.re - di * w.im; sm.z[i1].im = dr * w.im + di * w.re;
            }
            __syncthreads();
        }
        // ---- (4) stereo split, magnitudes (stft_mask2048.jsfx process_frame; Sample.jsfx:3575-3585) ---------
        for (int k = tid; k < kBins; k += NT) {
            const Cx<R> p = sm.z[brev11(k)], q = sm.z[brev11((kN - k) & (kN - 1))];
            const R reL = mul((R)0.5, add(p.re, q.re)), imL = mul((R)0.5, sub(p.im, q.im));
            const R reR = mul((R)0.5, add(p.im, q.im)), imR = mul((R)0.5, sub(q.re, p.re));
            const R m2L = max_r((R)0, add(mul(reL, reL), mul(imL, imL)));
            const R m2R = max_r((R)0, add(mul(reR, reR), mul(imR, imR)));
            sm.mag[k] = sqrt_r<R>(mul((R)0.5, add(m2L, m2R)));
        }
        __syncthreads();
        // ---- per-bin gain rule with recursive state ----------------------------------------------------------
        {
            const R tiny = (R)1e-18, ab = (R)c.ab, bb = (R)c.bb, thresh = (R)c.thresh, amt = (R)c.amt, floor_g = (R)c.floor_g;
            for (int k = tid; k < kBins; k += NT) {
                if (k == 0 || k == kBins - 1) { sm.gain[k] = (R)1; continue; }
                const R magbin = sm.mag[k];
                R base_old = sm.base[k];
                if (base_old < tiny) base_old = max_r(magbin, tiny);
                sm.base[k] = add(mul(ab, base_old), mul(bb, magbin));
                const R flux = div_r<R>(max_r((R)0, sub(magbin, sm.prev[k])), add(base_old, tiny));
                sm.prev[k] = magbin;
                R neigh = mul((R)0.25, add(add(add(sm.mag[k - 2 < 0 ? 0 : k - 2], sm.mag[k - 1]), sm.mag[k + 1]),
                                           sm.mag[k + 2 > kBins - 1 ? kBins - 1 : k + 2]));
                if (neigh < tiny) neigh = base_old;
                const R peak = div_r<R>(magbin, add(neigh, tiny));
                const R dp = clamp01(div_r<R>(sub(peak, thresh), max_r((R)0.18, thresh)));
                const R df = clamp01(div_r<R>(sub(flux, (R)1), (R)2));
                const R control = mul(amt, clamp01(add(mul((R)0.7, dp), mul((R)0.3, df))));
                const R ref = add(mul((R)0.75, neigh), mul((R)0.25, base_old));
                const R hard = min_r((R)1, div_r<R>(add(ref, tiny), add(magbin, tiny)));
                R raw = sub((R)1, mul(control, sub((R)1, hard)));
                raw = max_r(floor_g, min_r((R)1, raw));
                sm.gain[k] = raw;
            }
        }
        __syncthreads();
        // ---- (5) 3-tap smoothing, applied to bins k and N-k of the packed spectrum ----------------------------
        for (int k = tid + 1; k < kBins - 1; k += NT) {
            const R sg = add(add(mul((R)0.16, sm.gain[k - 1]), mul((R)0.68, sm.gain[k])), mul((R)0.16, sm.gain[k + 1]));
            const int i0 = brev11(k), i1 = brev11(kN - k);
            sm.z[i0].re = mul(sm.z[i0].re, sg); sm.z[i0].im = mul(sm.z[i0].im, sg);
            sm.z[i1].re = mul(sm.z[i1].re, sg); sm.z[i1].im = mul(sm.z[i1].im, sg);
        }
        __syncthreads();
        // ---- (6) inverse FFT, radix-2 decimation in time: bit-reversed in, natural out (unscaled, like WDL) ----
        for (int half = 1, tstep = kN / 2; half <= kN / 2; half <<= 1, tstep >>= 1) {
            for (int j = tid; j < kN / 2; j += NT) {
                const int p = j & (half - 1), i0 = ((j - p) << 1) + p, i1 = i0 + half;
                const Cx<R> u = sm.z[i0], v = sm.z[i1], w = sm.tw[p * tstep];
                const R tr = v.re * w.re + v.im * w.im, ti = v.im * w.re - v.re * w.im;      // v * conj(w)
                sm.z[i0].re = u.re + tr; sm.z[i0].im = u.im + ti;
                sm.z[i1].re = u.re - tr; sm.z[i1].im = u.im - ti;
            }
            __syncthreads();
        }
        // ---- (7) x 1/N x synthesis window, overlap-add into output hops hop+1 .. hop+4 -------------------------
        {
            const R scale = (R)(1.0 / kN);
            for (int i = tid; i < kN; i += NT) {
                const int slot = (int)((hop + 1 + (i >> 9)) % 4), o = slot * kHop + (i & 511);
                const R sy = (R)a.syn[i];
                sm.acc[0][o] = add(sm.acc[0][o], mul(mul(sm.z[i].re, scale), sy));
                sm.acc[1][o] = add(sm.acc[1][o], mul(mul(sm.z[i].im, scale), sy));
            }
        }
        ++hop;
        __syncthreads();
    }

    // ---- state -> HBM (carried into the next call) ------------------------------------------------------------
    for (int i = tid; i < 2 * kN; i += NT) gs.acc[i / kN][i % kN] = (double)sm.acc[i / kN][i % kN];
    for (int i = tid; i < 2 * 5 * kHop; i += NT) gs.xin[i / (5 * kHop)][i % (5 * kHop)] = sm.xin[i / (5 * kHop)][i % (5 * kHop)];
    for (int i = tid; i < kBins; i += NT) { gs.base[i] = (double)sm.base[i]; gs.prev[i] = (double)sm.prev[i]; }
}

template <typename R, int NT, bool TAP>
static bool launch_stft(const StftArgs& a, cudaStream_t stream)
{
    auto kern = stft_mask_kernel<R, NT, TAP>;
    const size_t smem = sizeof(StftSmem<R, NT>);
    static bool attr_done = false;
    if (!attr_done) {
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "stft smem attr")) return false;
        attr_done = true;
    }
    kern<<<a.n_streams, NT, smem, stream>>>(a);
    return cuda_ok(cudaGetLastError(), "stft_mask_kernel launch");
}

bool stft_launch(const StftArgs& a, int precision, cudaStream_t stream)
{
    if (a.n_frames <= 0 || a.n_streams <= 0) return true;
    if (precision) return a.tap ? launch_stft<double, 512, true>(a, stream) : launch_stft<double, 512, false>(a, stream);
    return launch_stft<float, 256, false>(a, stream);
}

}  // namespace za
