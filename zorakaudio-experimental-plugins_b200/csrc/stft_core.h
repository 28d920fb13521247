// stft_core.h -- per-hop phases of the STFT class kernel (plugin "stft_mask2048"), written once for the
// device (k_stft.cu: 128 threads per stream, phases separated by named barriers) and for a host
// emulation of the same thread decomposition (tools/emu_stft.cpp: index arithmetic checked on the CPU
// against the reference build before GPU time is spent).
//
// Replaces process_frame() of plugins/Workloads/stft_mask2048/src/stft_mask2048.jsfx (structure of "DECRUST",
// plugins/Spectral/Sample/src/Sample.jsfx:3551-3688) and the FFT hooks it calls (jsfx_fft / fft_permute /
// fft_ipermute / ifft, src/JSFXJuceProcessor.cpp:1267-1447 over src/WDL/fft.c).
//
// FFT: 2048 = 16 x 16 x 8, decimation in frequency forward / decimation in time inverse, 16 points per
// thread in registers, two shared-memory exchanges per transform.  Stereo is packed as z = w (l + i r); the
// split into the two channel spectra needs Z[k] and Z[N-k]: thread t (1..127) runs the two last-pass
// radix-8 blocks k = t (mod 256) and k = 256 - t (mod 256), thread 0 the self-paired blocks 0 and 128, so
// every pair already sits in ONE thread's registers: magnitudes, the per-bin recursive state (registers,
// fixed bins per thread), gain application and the first inverse pass need no exchange of Z.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define ZA_HD __host__ __device__ __forceinline__
#else
#define ZA_HD inline
#endif

namespace za {
namespace stft {

constexpr int kN = 2048, kHop = 512, kBins = 1025, kT = 128;   // kT = threads per stream
constexpr int kZLen = kN + kN / 128;                          // one pad entry per 128 (conflict-free strides)

template <typename R> struct Cx { R re, im; };

// ---- rounding-pinned scalar ops for the script-order parts (bins, window, overlap-add) ----------------
#if defined(__CUDA_ARCH__)
ZA_HD float mul_(float a, float b) { return a * b; }
ZA_HD float add_(float a, float b) { return a + b; }
ZA_HD float sub_(float a, float b) { return a - b; }
ZA_HD double mul_(double a, double b) { return __dmul_rn(a, b); }
ZA_HD double add_(double a, double b) { return __dadd_rn(a, b); }
ZA_HD double sub_(double a, double b) { return __dsub_rn(a, b); }
ZA_HD float div_(float a, float b) { return __fdividef(a, b); }
ZA_HD double div_(double a, double b) { return __ddiv_rn(a, b); }
ZA_HD float sqrt_(float a) { return sqrtf(a); }
ZA_HD double sqrt_(double a) { return __dsqrt_rn(a); }
#else
template <typename R> ZA_HD R mul_(R a, R b) { return a * b; }
template <typename R> ZA_HD R add_(R a, R b) { return a + b; }
template <typename R> ZA_HD R sub_(R a, R b) { return a - b; }
template <typename R> ZA_HD R div_(R a, R b) { return a / b; }
ZA_HD float sqrt_(float a) { return __builtin_sqrtf(a); }
ZA_HD double sqrt_(double a) { return __builtin_sqrt(a); }
#endif
template <typename R> ZA_HD R max_(R a, R b) { return a > b ? a : b; }      // ogt select (dsp_jsfx_aot.py:5223-5232)
template <typename R> ZA_HD R min_(R a, R b) { return a < b ? a : b; }
template <typename R> ZA_HD R clamp01(R x) { return x < (R)0 ? (R)0 : (x > (R)1 ? (R)1 : x); }

ZA_HD int pz(int i) { return i + (i >> 7); }                   // padded index into z[]

// ---- shared memory ------------------------------------------------------------------------------------------
template <typename R> struct Tables {       // per CTA, read-only
    Cx<R> tw1[15][kT];                      // tw1[q-1][t] = exp(-2 pi i t q / 2048)
    Cx<R> tw2[15][8];                       // tw2[q-1][j] = exp(-2 pi i j q / 128)
    R win[kN];                              // analysis window
    R syn[kN];                              // synthesis window x 1/N (a power of two: exact)
};
template <typename R> struct StreamSmem {   // per stream
    Cx<R> z[kZLen];                         // FFT exchange buffer; between the transforms: mag[1028] | gain[1028]
    float xin[2][4 * kHop];                 // last four hops of input per channel (slot = hop index mod 4)
    R acc[2][4 * kHop];                     // overlap-add accumulator (slot = hop index mod 4)
    ZA_HD R* mag() { return reinterpret_cast<R*>(z); }
    ZA_HD R* gain() { return reinterpret_cast<R*>(z) + 1028; }
};
struct Coef { double amt, thresh, ab, bb, floor_g, mixv; };   // @slider results (same layout as za::StftCoef)

template <typename R> struct Regs {         // per thread, lives across hops
    Cx<R> v[16];                            // FFT working set; between the transforms: A = v[0..7], B = v[8..15]
    R base[8], prev[8];                     // per-bin recursive state of this thread's bins (slot order)
    R magr[8];                              // this hop's magnitudes of the same bins
};

// ---- small DFTs in registers (natural order in and out); SIGN = -1 forward, +1 inverse -----------------------
template <typename R, int SIGN> ZA_HD Cx<R> rot(Cx<R> a)       // a * (SIGN * i)
{
    return SIGN < 0 ? Cx<R>{a.im, -a.re} : Cx<R>{-a.im, a.re};
}
template <typename R, int SIGN> ZA_HD Cx<R> cmulw(Cx<R> a, Cx<R> w)   // a * w (forward) or a * conj(w) (inverse)
{
    return SIGN < 0 ? Cx<R>{a.re * w.re - a.im * w.im, a.re * w.im + a.im * w.re}
                    : Cx<R>{a.re * w.re + a.im * w.im, a.im * w.re - a.re * w.im};
}
template <typename R, int SIGN> ZA_HD void dft4(Cx<R>& a0, Cx<R>& a1, Cx<R>& a2, Cx<R>& a3)
{
    const Cx<R> t0{a0.re + a2.re, a0.im + a2.im}, t1{a0.re - a2.re, a0.im - a2.im};
    const Cx<R> t2{a1.re + a3.re, a1.im + a3.im};
    const Cx<R> t3 = rot<R, SIGN>(Cx<R>{a1.re - a3.re, a1.im - a3.im});
    a0 = {t0.re + t2.re, t0.im + t2.im}; a2 = {t0.re - t2.re, t0.im - t2.im};
    a1 = {t1.re + t3.re, t1.im + t3.im}; a3 = {t1.re - t3.re, t1.im - t3.im};
}
template <typename R, int SIGN> ZA_HD void dft16(Cx<R> (&v)[16])
{
    const R c1 = (R)0.92387953251128675613, s1 = (R)0.38268343236508977173, r = (R)0.70710678118654752440;
#pragma unroll
    for (int p2 = 0; p2 < 4; ++p2) dft4<R, SIGN>(v[p2], v[p2 + 4], v[p2 + 8], v[p2 + 12]);
    // v[p2 + 4 q1] *= w16^(p2 q1); constants are the forward values (cos, -sin), conjugated for SIGN > 0
    v[5] = cmulw<R, SIGN>(v[5], Cx<R>{c1, -s1});      // p2 1, q1 1: w^1
    v[9] = cmulw<R, SIGN>(v[9], Cx<R>{r, -r});        // 1,2: w^2
    v[13] = cmulw<R, SIGN>(v[13], Cx<R>{s1, -c1});    // 1,3: w^3
    v[6] = cmulw<R, SIGN>(v[6], Cx<R>{r, -r});        // 2,1: w^2
    v[10] = rot<R, SIGN>(v[10]);                      // 2,2: w^4
    v[14] = cmulw<R, SIGN>(v[14], Cx<R>{-r, -r});     // 2,3: w^6
    v[7] = cmulw<R, SIGN>(v[7], Cx<R>{s1, -c1});      // 3,1: w^3
    v[11] = cmulw<R, SIGN>(v[11], Cx<R>{-r, -r});     // 3,2: w^6
    v[15] = cmulw<R, SIGN>(v[15], Cx<R>{-c1, s1});    // 3,3: w^9
#pragma unroll
    for (int q1 = 0; q1 < 4; ++q1) dft4<R, SIGN>(v[4 * q1], v[4 * q1 + 1], v[4 * q1 + 2], v[4 * q1 + 3]);
    // v[4 q1 + q2] holds X[q1 + 4 q2]: transpose 4x4 (register renaming)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = i + 1; j < 4; ++j) { const Cx<R> tmp = v[4 * i + j]; v[4 * i + j] = v[4 * j + i]; v[4 * j + i] = tmp; }
}
template <typename R, int SIGN> ZA_HD void dft8(Cx<R>* v)
{
    const R r = (R)0.70710678118654752440;
    dft4<R, SIGN>(v[0], v[2], v[4], v[6]);           // E[q1] at v[2 q1]
    dft4<R, SIGN>(v[1], v[3], v[5], v[7]);           // O[q1] at v[2 q1 + 1]
    v[3] = cmulw<R, SIGN>(v[3], Cx<R>{r, -r});       // w8^1
    v[5] = rot<R, SIGN>(v[5]);                       // w8^2
    v[7] = cmulw<R, SIGN>(v[7], Cx<R>{-r, -r});      // w8^3
    Cx<R> o[8];
#pragma unroll
    for (int q1 = 0; q1 < 4; ++q1) {
        const Cx<R> e = v[2 * q1], d = v[2 * q1 + 1];
        o[q1] = {e.re + d.re, e.im + d.im};
        o[q1 + 4] = {e.re - d.re, e.im - d.im};
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = o[q];
}

// ---- which bins a thread owns --------------------------------------------------------------------------------
// Thread t >= 1: block A = bins t + 256 q (v[q]), block B = bins 256 - t + 256 q (v[8 + q]); Z[N - k] of A[q] is B[7 - q].
// Thread 0: A = bins 256 q (pairs q <-> 8 - q, 0 and 1024 self-paired), B = bins 128 + 256 q (pairs q <-> 7 - q).
// Slot s = 0..7 -> (bin, register of Z[bin], register of Z[N - bin]); thread 0 has no slot 3.
template <bool T0> ZA_HD constexpr int slot_p(int s) { return T0 ? (s < 4 ? s + 1 : 4 + s) : (s < 4 ? s : 4 + s); }
template <bool T0> ZA_HD constexpr int slot_q(int s) { return T0 ? (s < 4 ? 7 - s : 19 - s) : (s < 4 ? 15 - s : 11 - s); }
template <bool T0> ZA_HD int slot_bin(int t, int s)
{
    return T0 ? (s < 3 ? 256 * (s + 1) : (s == 3 ? -1 : 128 + 256 * (s - 4)))
              : (s < 4 ? t + 256 * s : 256 - t + 256 * (s - 4));
}
ZA_HD int slot_bin_rt(int t, int s) { return t == 0 ? slot_bin<true>(0, s) : slot_bin<false>(t, s); }

// position of the radix-8 block with residue m = k mod 256 (m = q1 + 16 q2 sits at 128 q1 + 8 q2)
ZA_HD int block_pos(int m) { return 128 * (m & 15) + 8 * (m >> 4); }
ZA_HD int block_a(int t) { return t; }
ZA_HD int block_b(int t) { return t == 0 ? 128 : 256 - t; }

// ---- forward transform -------------------------------------------------------------------------------------------
template <typename R>
ZA_HD void fwd1(int t, Regs<R>& g, StreamSmem<R>& S, const Tables<R>& tb, int64_t hop)
{
#pragma unroll
    for (int p = 0; p < 16; ++p) {          // frame `hop` = input hops hop-3 .. hop (stft_mask2048.jsfx process_frame)
        const int i = t + 128 * p;
        const int o = (int)((hop + 1 + (p >> 2)) & 3) * kHop + t + 128 * (p & 3);
        const R w = tb.win[i];
        g.v[p].re = mul_((R)S.xin[0][o], w);
        g.v[p].im = mul_((R)S.xin[1][o], w);
    }
    dft16<R, -1>(g.v);
    S.z[pz(t)] = g.v[0];
#pragma unroll
    for (int q = 1; q < 16; ++q) S.z[pz(128 * q + t)] = cmulw<R, -1>(g.v[q], tb.tw1[q - 1][t]);
}
template <typename R>
ZA_HD void fwd2(int t, Regs<R>& g, StreamSmem<R>& S, const Tables<R>& tb)
{
    const int b = t & 15, j = t >> 4;
#pragma unroll
    for (int p = 0; p < 16; ++p) g.v[p] = S.z[pz(128 * b + j + 8 * p)];
    dft16<R, -1>(g.v);
    S.z[pz(128 * b + j)] = g.v[0];
#pragma unroll
    for (int q = 1; q < 16; ++q) S.z[pz(128 * b + 8 * q + j)] = cmulw<R, -1>(g.v[q], tb.tw2[q - 1][j]);
}
template <typename R>
ZA_HD void fwd3(int t, Regs<R>& g, StreamSmem<R>& S)
{
    const int pa = pz(block_pos(block_a(t))), pb = pz(block_pos(block_b(t)));
#pragma unroll
    for (int p = 0; p < 8; ++p) { g.v[p] = S.z[pa + p]; g.v[8 + p] = S.z[pb + p]; }
    dft8<R, -1>(g.v);
    dft8<R, -1>(g.v + 8);
}

// ---- bins: magnitudes, gain rule with recursive state, smoothing, application -------------------------------------
template <typename R> ZA_HD R mag_of(Cx<R> p, Cx<R> q)         // stereo split by conjugate symmetry, then the script's magnitude
{
    const R reL = mul_((R)0.5, add_(p.re, q.re)), imL = mul_((R)0.5, sub_(p.im, q.im));
    const R reR = mul_((R)0.5, add_(p.im, q.im)), imR = mul_((R)0.5, sub_(q.re, p.re));
    const R m2L = max_((R)0, add_(mul_(reL, reL), mul_(imL, imL)));
    const R m2R = max_((R)0, add_(mul_(reR, reR), mul_(imR, imR)));
    return sqrt_(mul_((R)0.5, add_(m2L, m2R)));
}
template <typename R, bool T0>
ZA_HD void mags(int t, Regs<R>& g, StreamSmem<R>& S)
{
    R* mag = S.mag();
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        if (T0 && s == 3) continue;
        const R m = mag_of<R>(g.v[slot_p<T0>(s)], g.v[slot_q<T0>(s)]);
        g.magr[s] = m;
        mag[slot_bin<T0>(t, s)] = m;
    }
    if (T0) { mag[0] = mag_of<R>(g.v[0], g.v[0]); mag[1024] = mag_of<R>(g.v[4], g.v[4]); }
}
template <typename R, bool T0>
ZA_HD void gains(int t, Regs<R>& g, StreamSmem<R>& S, const Coef& c)
{
    const R* mag = S.mag();
    R* gain = S.gain();
    const R tiny = (R)1e-18, ab = (R)c.ab, bb = (R)c.bb, thresh = (R)c.thresh, amt = (R)c.amt, floor_g = (R)c.floor_g;
    const R thr_den = max_((R)0.18, thresh);
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        if (T0 && s == 3) continue;
        const int k = slot_bin<T0>(t, s);
        const R magbin = g.magr[s];
        R base_old = g.base[s];
        if (base_old < tiny) base_old = max_(magbin, tiny);
        g.base[s] = add_(mul_(ab, base_old), mul_(bb, magbin));
        const R flux = div_(max_((R)0, sub_(magbin, g.prev[s])), add_(base_old, tiny));
        g.prev[s] = magbin;
        R neigh = mul_((R)0.25, add_(add_(add_(mag[k - 2 < 0 ? 0 : k - 2], mag[k - 1]), mag[k + 1]),
                                     mag[k + 2 > kBins - 1 ? kBins - 1 : k + 2]));
        if (neigh < tiny) neigh = base_old;
        const R peak = div_(magbin, add_(neigh, tiny));
        const R dp = clamp01(div_(sub_(peak, thresh), thr_den));
        const R df = clamp01(div_(sub_(flux, (R)1), (R)2));
        const R control = mul_(amt, clamp01(add_(mul_((R)0.7, dp), mul_((R)0.3, df))));
        const R ref = add_(mul_((R)0.75, neigh), mul_((R)0.25, base_old));
        const R hard = min_((R)1, div_(add_(ref, tiny), add_(magbin, tiny)));
        R raw = sub_((R)1, mul_(control, sub_((R)1, hard)));
        raw = max_(floor_g, min_((R)1, raw));
        gain[k] = raw;
    }
    if (T0) { gain[0] = (R)1; gain[1024] = (R)1; }
}
template <typename R, bool T0>
ZA_HD void apply_gains(int t, Regs<R>& g, StreamSmem<R>& S)
{
    const R* gain = S.gain();
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        if (T0 && s == 3) continue;
        const int k = slot_bin<T0>(t, s);
        const R sg = add_(add_(mul_((R)0.16, gain[k - 1]), mul_((R)0.68, gain[k])), mul_((R)0.16, gain[k + 1]));
        Cx<R>& p = g.v[slot_p<T0>(s)];
        Cx<R>& q = g.v[slot_q<T0>(s)];
        p.re = mul_(p.re, sg); p.im = mul_(p.im, sg);
        q.re = mul_(q.re, sg); q.im = mul_(q.im, sg);
    }
}

// ---- uniform bins path: thread 0 first renames its registers into the regular pair layout (slot s < 4: P = v[s],
// Q = v[15 - s]; s >= 4: P = v[4 + s], Q = v[11 - s]) so all 128 threads run ONE instruction stream; its slot 3 holds
// the self-paired bins 0 (v[3]) and 1024 (v[12]), which only contribute magnitudes --------------------------------------
template <typename R>
ZA_HD void t0_permute(bool t0, Regs<R>& g)       // v' = [v1 v2 v3 v0 | v12..v15 | v8..v11 | v4..v7]
{
    Cx<R> n[16];
    n[0] = g.v[1]; n[1] = g.v[2]; n[2] = g.v[3]; n[3] = g.v[0];
#pragma unroll
    for (int i = 0; i < 4; ++i) { n[4 + i] = g.v[12 + i]; n[8 + i] = g.v[8 + i]; n[12 + i] = g.v[4 + i]; }
#pragma unroll
    for (int i = 0; i < 16; ++i) { g.v[i].re = t0 ? n[i].re : g.v[i].re; g.v[i].im = t0 ? n[i].im : g.v[i].im; }
}
template <typename R>
ZA_HD void t0_unpermute(bool t0, Regs<R>& g)
{
    Cx<R> n[16];
    n[1] = g.v[0]; n[2] = g.v[1]; n[3] = g.v[2]; n[0] = g.v[3];
#pragma unroll
    for (int i = 0; i < 4; ++i) { n[12 + i] = g.v[4 + i]; n[8 + i] = g.v[8 + i]; n[4 + i] = g.v[12 + i]; }
#pragma unroll
    for (int i = 0; i < 16; ++i) { g.v[i].re = t0 ? n[i].re : g.v[i].re; g.v[i].im = t0 ? n[i].im : g.v[i].im; }
}
template <typename R>
ZA_HD void mags_u(int t, Regs<R>& g, StreamSmem<R>& S)
{
    R* mag = S.mag();
    const bool t0 = t == 0;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        // thread 0, slot 3: bin 0 is self-paired (P = Q = v[3]); bin 1024 (v[12]) follows below
        const Cx<R> p = g.v[slot_p<false>(s)];
        const Cx<R> q = (t0 && s == 3) ? p : g.v[slot_q<false>(s)];
        const R m = mag_of<R>(p, q);
        g.magr[s] = m;
        const int k = slot_bin_rt(t, s);
        mag[k < 0 ? 0 : k] = m;
    }
    if (t0) mag[1024] = mag_of<R>(g.v[12], g.v[12]);
}
template <typename R>
ZA_HD void gains_u(int t, Regs<R>& g, StreamSmem<R>& S, const Coef& c)
{
    const R* mag = S.mag();
    R* gain = S.gain();
    const R tiny = (R)1e-18, ab = (R)c.ab, bb = (R)c.bb, thresh = (R)c.thresh, amt = (R)c.amt, floor_g = (R)c.floor_g;
    const R thr_den = max_((R)0.18, thresh);
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int k = slot_bin_rt(t, s);
        if (k < 0) continue;                               // thread 0, slot 3: bins 0 / 1024 keep gain 1 and no state
        const R magbin = g.magr[s];
        R base_old = g.base[s];
        if (base_old < tiny) base_old = max_(magbin, tiny);
        g.base[s] = add_(mul_(ab, base_old), mul_(bb, magbin));
        const R flux = div_(max_((R)0, sub_(magbin, g.prev[s])), add_(base_old, tiny));
        g.prev[s] = magbin;
        R neigh = mul_((R)0.25, add_(add_(add_(mag[k - 2 < 0 ? 0 : k - 2], mag[k - 1]), mag[k + 1]),
                                     mag[k + 2 > kBins - 1 ? kBins - 1 : k + 2]));
        if (neigh < tiny) neigh = base_old;
        const R peak = div_(magbin, add_(neigh, tiny));
        const R dp = clamp01(div_(sub_(peak, thresh), thr_den));
        const R df = clamp01(div_(sub_(flux, (R)1), (R)2));
        const R control = mul_(amt, clamp01(add_(mul_((R)0.7, dp), mul_((R)0.3, df))));
        const R ref = add_(mul_((R)0.75, neigh), mul_((R)0.25, base_old));
        const R hard = min_((R)1, div_(add_(ref, tiny), add_(magbin, tiny)));
        R raw = sub_((R)1, mul_(control, sub_((R)1, hard)));
        raw = max_(floor_g, min_((R)1, raw));
        gain[k] = raw;
    }
    if (t == 0) { gain[0] = (R)1; gain[1024] = (R)1; }
}
template <typename R>
ZA_HD void apply_gains_u(int t, Regs<R>& g, StreamSmem<R>& S)
{
    const R* gain = S.gain();
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int k = slot_bin_rt(t, s);
        if (k < 0) continue;
        const R sg = add_(add_(mul_((R)0.16, gain[k - 1]), mul_((R)0.68, gain[k])), mul_((R)0.16, gain[k + 1]));
        Cx<R>& p = g.v[slot_p<false>(s)];
        Cx<R>& q = g.v[slot_q<false>(s)];
        p.re = mul_(p.re, sg); p.im = mul_(p.im, sg);
        q.re = mul_(q.re, sg); q.im = mul_(q.im, sg);
    }
}

// ---- inverse transform (unscaled, like WDL's) + overlap-add ---------------------------------------------------------
template <typename R>
ZA_HD void inv3(int t, Regs<R>& g, StreamSmem<R>& S)
{
    dft8<R, +1>(g.v);
    dft8<R, +1>(g.v + 8);
    const int pa = pz(block_pos(block_a(t))), pb = pz(block_pos(block_b(t)));
#pragma unroll
    for (int p = 0; p < 8; ++p) { S.z[pa + p] = g.v[p]; S.z[pb + p] = g.v[8 + p]; }
}
template <typename R>
ZA_HD void inv2(int t, Regs<R>& g, StreamSmem<R>& S, const Tables<R>& tb)
{
    const int b = t & 15, j = t >> 4;
    g.v[0] = S.z[pz(128 * b + j)];
#pragma unroll
    for (int q = 1; q < 16; ++q) g.v[q] = cmulw<R, +1>(S.z[pz(128 * b + 8 * q + j)], tb.tw2[q - 1][j]);
    dft16<R, +1>(g.v);
#pragma unroll
    for (int p = 0; p < 16; ++p) S.z[pz(128 * b + j + 8 * p)] = g.v[p];
}
template <typename R>
ZA_HD void inv1_ola(int t, Regs<R>& g, StreamSmem<R>& S, const Tables<R>& tb, int64_t hop)
{
    g.v[0] = S.z[pz(t)];
#pragma unroll
    for (int q = 1; q < 16; ++q) g.v[q] = cmulw<R, +1>(S.z[pz(128 * q + t)], tb.tw1[q - 1][t]);
    dft16<R, +1>(g.v);
#pragma unroll
    for (int p = 0; p < 16; ++p) {          // x 1/N x synthesis window into output hops hop+1 .. hop+4
        const int i = t + 128 * p;
        const int o = (int)((hop + 1 + (p >> 2)) & 3) * kHop + t + 128 * (p & 3);
        const R sy = tb.syn[i];
        S.acc[0][o] = add_(S.acc[0][o], mul_(g.v[p].re, sy));
        S.acc[1][o] = add_(S.acc[1][o], mul_(g.v[p].im, sy));
    }
}

}  // namespace stft
}  // namespace za
