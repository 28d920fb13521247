// jsfx_rt.cuh -- device runtime under the translated DSP-JSFX scripts (csrc/gen/g_*.cu, written by
// za_gpu/jsfxc.py): the per-thread context (one stream per thread), mem[] as a stream-minor
// structure-of-arrays heap in HBM, the arithmetic the reference compiler pins (dsp_jsfx_aot.py:4263-5590),
// its per-instance MT19937 (:3880-4060), and the two kernels every script instantiates:
//   jsfx_prepare_kernel  = prepareToPlay: sliders pushed, @init, alias re-apply, @slider
//                          (src/JSFXJuceProcessor.cpp:3297-3318)
//   jsfx_process_kernel  = processBlock -> jsfx_process_block for every host block of the call: push the
//                          block's quantised sliders, @slider if one changed (:3536-3556), samplesblock,
//                          @block, then per frame float -> spl[], @sample, spl[] -> float
//                          (dsp_jsfx_aot.py:5766-5892)
// Audio moves through shared memory in [32 streams] x [32 frames] tiles so HBM sees 128-byte rows although a
// thread walks along its own stream.
#pragma once
#include <math.h>

#ifdef ZA_JSFX_HOST_EMU
// Host emulation (tools/emu_jsfx.py): the generated translation units compiled as plain C++ so the translator's
// semantics can be checked against the reference build on the CPU, one stream at a time.  Development / test aid only.
#include <stdint.h>
#include <string.h>
#include <algorithm>
#include <vector>
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__
inline double __dadd_rn(double a, double b) { return a + b; }
inline double __dsub_rn(double a, double b) { return a - b; }
inline double __dmul_rn(double a, double b) { return a * b; }
inline double __ddiv_rn(double a, double b) { return a / b; }
inline double __dsqrt_rn(double a) { return sqrt(a); }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fdiv_rn(float a, float b) { return a / b; }
inline float __fdividef(float a, float b) { return a / b; }
inline float __fsqrt_rn(float a) { return sqrtf(a); }
inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
using std::min;
using std::max;
namespace za {
struct GenLaunch {
    const float* in; float* out; double* tap;
    int64_t row_stride, n_frames, mem_slots;
    int n_streams, block_len, n_rows, n_decl;
    const double* q; const int* decl_index0;
    void* state; void* mem; unsigned* mt; int* err; double srate;
    int64_t fft_scratch; const double* tw;
    int lanes, unstaged;
};
}
#else
#include "za_device.cuh"
#include "za_internal.h"
#endif

// User functions of a script: inlined by default, so the variable file v[] (constant indices only) is promoted to
// registers; a leaf whose straight-line code then outgrows the instruction cache can be built with
// -DZA_JSFX_FN=__noinline__ (v[] goes to local memory, calls are real calls).
#ifndef ZA_JSFX_FN
#define ZA_JSFX_FN __forceinline__
#endif

namespace za {
namespace jsfx {

// ---- arithmetic (fp64: separate, never contracted operations in the reference's order; fp32: plain float) ----
__device__ __forceinline__ double add_(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double mul_(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double div_(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float add_(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float mul_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float div_(float a, float b) { return __fdividef(a, b); }      // fp32 mode: MUFU.RCP + FMUL (2 ulp), no slow-path call
// Conditions and `!=` are ORDERED compares in the reference compiler (fcmp one: dsp_jsfx_aot.py:3725-3726 `_truthy`,
// :4349-4352 `!=` -> "one"), so a NaN is false / compares "not different"; C++'s != is unordered (true on NaN).
__device__ __forceinline__ bool truthy_(double x) { return x < 0.0 || x > 0.0; }
__device__ __forceinline__ bool truthy_(float x) { return x < 0.0f || x > 0.0f; }
__device__ __forceinline__ bool ne_(double a, double b) { return a < b || a > b; }
__device__ __forceinline__ bool ne_(float a, float b) { return a < b || a > b; }
__device__ __forceinline__ double pow_(double a, double b) { return pow(a, b); }
__device__ __forceinline__ float pow_(float a, float b) { return powf(a, b); }

#define ZA_M1(name, fd, ff)                                                  \
    __device__ __forceinline__ double m_##name(double a) { return fd(a); }   \
    __device__ __forceinline__ float m_##name(float a) { return ff(a); }
ZA_M1(sin, sin, sinf) ZA_M1(cos, cos, cosf) ZA_M1(tan, tan, tanf) ZA_M1(asin, asin, asinf) ZA_M1(acos, acos, acosf)
ZA_M1(atan, atan, atanf) ZA_M1(exp, exp, expf) ZA_M1(log, log, logf) ZA_M1(log10, log10, log10f)
ZA_M1(floor, floor, floorf) ZA_M1(ceil, ceil, ceilf) ZA_M1(abs, fabs, fabsf)
#undef ZA_M1
__device__ __forceinline__ double m_sqrt(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ float m_sqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ double m_pow(double a, double b) { return pow(a, b); }
__device__ __forceinline__ float m_pow(float a, float b) { return powf(a, b); }
__device__ __forceinline__ double m_atan2(double a, double b) { return atan2(a, b); }
__device__ __forceinline__ float m_atan2(float a, float b) { return atan2f(a, b); }
template <typename R> __device__ __forceinline__ R m_invsqrt(R x)      // dsp_jsfx_aot.py:5256-5277
{
    const int bits = __float_as_int((float)x);
    const R y0 = (R)__int_as_float(0x5f3759df - (bits >> 1));
    return mul_(y0, sub_((R)1.5, mul_(mul_((R)0.5, x), mul_(y0, y0))));
}

template <typename R> __device__ __forceinline__ int to_i32(R x) { return (int)(long long)x; }      // :4107-4110
__device__ __forceinline__ int shl_i32(int a, int b) { return (int)((unsigned)a << (b & 31)); }
__device__ __forceinline__ int rem_i32(int a, int b) { return (b == 0 || (a == (int)0x80000000 && b == -1)) ? 0 : a % b; }
template <typename R> __device__ __forceinline__ long long to_count(R x) { const long long n = (long long)x; return n < 0 ? 0 : n; }

// ---- per-thread context ---------------------------------------------------------------------------------------
template <typename R> struct Rt {
    R* mem;                 // this stream's column of the heap: slot i at mem[i * stride]
    long long stride, cap;  // n_streams, slots per stream
    unsigned* mt;           // this stream's column of the MT19937 state: word k at mt[k * stride], index at word 624
    int* err;               // device error flag: 1 = mem[] index beyond the heap the plugin was sized for
    long long amax;         // largest mem[] index this thread has touched in this launch (fp64 mode)
    int amax32;             // the same, fp32 mode
    int cap32;              // cap as int
    unsigned stride_bytes;  // n_streams * sizeof(R): distance between two slots of one stream
    R srate, samplesblock;
    R sliders[64];
    R spl[64];
    long long scr, scr_n;         // FFT scratch: slots [scr, scr + scr_n) of this stream's heap column, beyond the script's own slots
    const double* tw;             // e^{-2 pi i k / 32768}, k = 0 .. 16383, as (cos, -sin) pairs (host libm)
    unsigned long long pending;   // pendingSliderChange | Automate | AutomateEnd masks of this host block (DSPJSFX_State, :3321-3387)


    // ---- FFT runtime hooks: jsfx_fft / ifft / fft_real / ifft_real / fft_permute / fft_ipermute / convolve_c
    // (src/JSFXJuceProcessor.cpp:1267-1447 over src/WDL/fft.c), for ONE stream per thread on its heap column.
    // Contract kept: sizes 16 .. 32768 powers of two, the region inside one 65536-slot page (:1131-1158), index rounding
    // v +- 1e-5 (:1121-1124), invalid arguments do nothing and return 0; forward kernel e^{-2 pi i k n / N}, unscaled both
    // ways (round trip gain N); between fft() and fft_permute() bin k lives at position WDL_fft_permute_tab(N)[k] =
    // wdl_pos(k, N) below (fft.c:990-1025); real forward = 2 x DFT with the Nyquist bin in bin0.im (fft.h:62-71).
    // The transform itself is a Stockham radix-2 autosort ping-ponging between the region and the per-stream scratch the
    // heap carries, followed by a scatter into WDL's order: same values as WDL's split-radix code to rounding.
    __device__ __forceinline__ static long long round_index(R v) { const double d = (double)v; return (long long)(d + (d >= 0.0 ? 1.0e-5 : -1.0e-5)); }
    __device__ __forceinline__ static bool size_ok(long long n) { return n >= 16 && n <= 32768 && (n & (n - 1)) == 0; }
    __device__ __forceinline__ static bool in_page(long long base, long long span)
    {
        if (base < 0 || span <= 0) return false;
        return (base / 65536) == ((base + span - 1) / 65536);
    }
    __device__ __forceinline__ R& at(long long a) { return *slot((unsigned)a); }
    // fftfreq_c (fft.c:990-1005) inverted: the buffer position of frequency j in djbfft's output order
    __device__ static int pos_of_freq(unsigned j, unsigned n)
    {
        int p = 0;
        while (n > 2) {
            if (!(j & 1)) { j >>= 1; n >>= 1; }
            else if ((j & 3) == 1) { p += n >> 1; j = (j - 1) >> 2; n >>= 2; }
            else { p += (n >> 1) + (n >> 2); j = ((j + 1) & (n - 1)) >> 2; n >>= 2; }
        }
        return p + (int)j;
    }
    __device__ __forceinline__ static int wdl_pos(int k, int n) { return k == 0 ? 0 : pos_of_freq((unsigned)(n - k), (unsigned)n); }   // _idxperm[n - j] = i
    __device__ __forceinline__ bool note(long long last)        // heap bounds: like every mem[] access
    {
        if (sizeof(R) == 8) amax = last > amax ? last : amax; else amax32 = max(amax32, (int)min(last, (long long)0x7fffffff));
        return last < cap;
    }
    __device__ bool complex_region(R baseR, long long n, long long& base)
    {
        long long start = round_index(baseR);
        if (start < 0) start = 0;
        if (n <= 0 || n > 32768 || !in_page(start, 2 * n)) return false;
        base = start;
        return note(start + 2 * n - 1);
    }
    __device__ bool need_scratch(long long slots)
    {
        if (slots <= scr_n) return true;
        *err = 1;                                  // the heap of this plugin was built without (enough) FFT scratch: fail the call loudly
        return false;
    }
    // natural-order DFT of the N complex values at `src` (slot index), result left at `dst`; src / dst are the region and the
    // scratch in either order; inverse = conjugate twiddles.
    __device__ void stockham(long long a, long long b, int N, bool inverse, bool want_in_a)
    {
        long long src = a, dst = b;
        const int half = N >> 1;
        for (int Ns = 1; Ns < N; Ns <<= 1) {
            const int tstep = 16384 / Ns;
            for (int j = 0; j < half; ++j) {
                const int k = j & (Ns - 1);
                const double wr = tw[2 * (k * tstep)], wi = inverse ? -tw[2 * (k * tstep) + 1] : tw[2 * (k * tstep) + 1];
                const R v0r = at(src + 2 * j), v0i = at(src + 2 * j + 1), xr = at(src + 2 * (j + half)), xi = at(src + 2 * (j + half) + 1);
                const R v1r = sub_(mul_((R)wr, xr), mul_((R)wi, xi)), v1i = add_(mul_((R)wr, xi), mul_((R)wi, xr));
                const long long j0 = (long long)(((j - k) << 1) + k);
                at(dst + 2 * j0) = add_(v0r, v1r); at(dst + 2 * j0 + 1) = add_(v0i, v1i);
                at(dst + 2 * (j0 + Ns)) = sub_(v0r, v1r); at(dst + 2 * (j0 + Ns) + 1) = sub_(v0i, v1i);
            }
            const long long t = src; src = dst; dst = t;
        }
        if ((src == a) != want_in_a)
            for (int i = 0; i < 2 * N; ++i) at(dst + i) = at(src + i);
    }
    __device__ void fft_c(long long base, int N, bool inverse)
    {
        const long long s0 = scr;
        if (!inverse) {         // region -> natural order in the scratch -> scattered to WDL's order in the region
            stockham(base, s0, N, false, false);
            for (int k = 0; k < N; ++k) { const int p = wdl_pos(k, N); at(base + 2 * p) = at(s0 + 2 * k); at(base + 2 * p + 1) = at(s0 + 2 * k + 1); }
        } else {                // WDL's order in the region -> natural order in the scratch -> time samples in the region
            for (int k = 0; k < N; ++k) { const int p = wdl_pos(k, N); at(s0 + 2 * k) = at(base + 2 * p); at(s0 + 2 * k + 1) = at(base + 2 * p + 1); }
            stockham(s0, base, N, true, false);
        }
    }
    __device__ R fft(R b, R sz, bool inverse)
    {
        const long long N = round_index(sz);
        long long base;
        if (!size_ok(N) || !complex_region(b, N, base) || !need_scratch(2 * N)) return R(0);
        fft_c(base, (int)N, inverse);
        return R(0);
    }
    __device__ R fft_permute(R b, R sz, bool inverse)
    {
        const long long N = round_index(sz);
        long long base;
        if (!size_ok(N) || !complex_region(b, N, base) || !need_scratch(2 * N)) return R(0);
        for (int i = 0; i < (int)N; ++i) {
            const int p = wdl_pos(i, (int)N);
            if (!inverse) { at(scr + 2 * i) = at(base + 2 * p); at(scr + 2 * i + 1) = at(base + 2 * p + 1); }        // sc[i] = buf[perm[i]]
            else { at(scr + 2 * p) = at(base + 2 * i); at(scr + 2 * p + 1) = at(base + 2 * i + 1); }               // sc[perm[i]] = buf[i]
        }
        for (int i = 0; i < 2 * (int)N; ++i) at(base + i) = at(scr + i);
        return R(0);
    }
    // WDL_real_fft (fft.c:1110-1178): N reals as N/2 complex values, two-for-one split with tw = e^{2 pi i k / N}
    __device__ R fft_real(R b, R sz, bool inverse)
    {
        const long long len = round_index(sz);
        long long start = round_index(b);
        if (start < 0) start = 0;
        if (!size_ok(len) || !in_page(start, len) || !note(start + len - 1) || !need_scratch(len)) return R(0);
        const int half = (int)len >> 1, quart = half >> 1;
        if (!inverse) {
            fft_c(start, half, false);
            const R t1 = add_(at(start), at(start + 1)), t2 = sub_(at(start), at(start + 1));
            at(start) = mul_(t1, R(2)); at(start + 1) = mul_(t2, R(2));                                          // r2
        } else {
            const R t1 = add_(at(start), at(start + 1)), t2 = sub_(at(start), at(start + 1));
            at(start) = t1; at(start + 1) = t2;                                                                  // v2
        }
        const int tstep = 32768 / (int)len;
        for (int i = 1; i < quart; ++i) {
            const long long p = start + 2 * wdl_pos(i, half), q = start + 2 * wdl_pos(half - i, half);
            double twr = tw[2 * (i * tstep)];                       // cos(2 pi i / len)
            const double twi = -tw[2 * (i * tstep) + 1];            // sin(2 pi i / len)
            if (!inverse) twr = -twr;
            const R sr = add_(at(p), at(q)), si = add_(at(p + 1), at(q + 1)), dr = sub_(at(p), at(q)), di = sub_(at(p + 1), at(q + 1));
            const R tw1 = add_(mul_((R)twr, si), mul_((R)twi, dr)), tw2 = sub_(mul_((R)twi, si), mul_((R)twr, dr));
            at(p) = sub_(sr, tw1); at(p + 1) = sub_(di, tw2);
            at(q) = add_(sr, tw1); at(q + 1) = sub_(R(0), add_(di, tw2));
        }
        {
            const long long p = start + 2 * wdl_pos(quart, half);
            at(p) = mul_(at(p), R(2)); at(p + 1) = mul_(at(p + 1), R(-2));
        }
        if (inverse) fft_c(start, half, true);
        return R(0);
    }
    __device__ R convolve_c(R d, R sR, R nR)                      // dest[i] *= src[i] (complex), :1341-1383
    {
        const long long n = round_index(nR);
        long long db, sb;
        if (!complex_region(d, n, db) || !complex_region(sR, n, sb)) return R(0);
        const bool overlaps = (db < sb + 2 * n) && (sb < db + 2 * n);
        long long src = sb;
        if (overlaps && db != sb) {
            if (!need_scratch(2 * n)) return R(0);
            for (long long i = 0; i < 2 * n; ++i) at(scr + i) = at(sb + i);
            src = scr;
        }
        for (long long i = 0; i < n; ++i) {
            const R ar = at(db + 2 * i), ai = at(db + 2 * i + 1), br = at(src + 2 * i), bi = at(src + 2 * i + 1);
            at(db + 2 * i) = sub_(mul_(ar, br), mul_(ai, bi));
            at(db + 2 * i + 1) = add_(mul_(ar, bi), mul_(ai, br));
        }
        return R(0);
    }

    // sliderchange(mask) / slider_automate(mask[, end]) inside a DSP section: the host hook ORs round(mask) into the pending
    // masks and returns 1 (src/JSFXJuceProcessor.cpp:2448-2476); @slider then re-runs once after @block
    // (dsp_jsfx_aot.py:5786-5804) and the host clears the masks after the block (consumeDspSliderChanges, :5665-5738).
    __device__ __forceinline__ R sliderchange(R mask)
    {
        const long long m = mask > R(0) ? llround((double)mask) : 0;
        if (m <= 0) return R(0);
        pending |= (unsigned long long)m;
        return R(1);
    }

    // a[b] = mem[trunc(a + b + 1e-5)], negative -> 0 (:4062-4103).  fp64 mode follows the reference's double
    // arithmetic; fp32 mode keeps the whole index computation in float / int32 (heaps are far below 2^24 slots, so
    // integer-valued offsets are exact either way) instead of paying two F2F, two DADD and an F2I.S64 per access.
    __device__ __forceinline__ long long addr(double base, double idx) const
    {
        const long long a = (long long)__dadd_rn(__dadd_rn(base, idx), 1.0e-5);
        return a < 0 ? 0 : a;
    }
    __device__ __forceinline__ int addr(float base, float idx) const
    {
        const int a = (int)__fadd_rn(__fadd_rn(base, idx), 1.0e-5f);     // saturating conversion, NaN -> 0
        return a < 0 ? 0 : a;
    }
    // mem[x] with no base (the common form): 0 + x is x for every x the truncation can tell apart
    __device__ __forceinline__ long long addr1(double idx) const
    {
        const long long a = (long long)__dadd_rn(idx, 1.0e-5);
        return a < 0 ? 0 : a;
    }
    __device__ __forceinline__ int addr1(float idx) const
    {
        const int a = (int)__fadd_rn(idx, 1.0e-5f);
        return a < 0 ? 0 : a;
    }
    // Accesses beyond the heap the plugin was sized for (the reference would grow it, :1452-1491): the index is
    // clamped to the spare slot `cap` every heap carries and the largest index seen is kept in a register; the kernel
    // raises the device error flag once, when it stores the context back -- no branch, flag store or constant load
    // on the per-access path.
    // slot a (already clamped to [0, cap]) of this stream's column: base + a * (n_streams * sizeof(R)) as a
    // 32 x 32 -> 64 bit multiply (IMAD.WIDE.U32) + add; a generic 64-bit index * 64-bit stride costs seven instructions
    __device__ __forceinline__ R* slot(unsigned a) const
    {
        return reinterpret_cast<R*>(reinterpret_cast<char*>(mem) + (unsigned long long)a * stride_bytes);
    }
    int pf;                 // 1 in the stage pipelines (jsfx_staged_kernel): rand() reads its state through the shared-memory window
    __device__ __forceinline__ R ldi(long long a)
    {
        amax = a > amax ? a : amax;
        return *slot((unsigned)(a < cap ? a : cap));
    }
    __device__ __forceinline__ void st(long long a, R v)
    {
        amax = a > amax ? a : amax;
        *slot((unsigned)(a < cap ? a : cap)) = v;
    }
    __device__ __forceinline__ R ldi(int a)     // fp32 mode: 32-bit index
    {
        amax32 = max(amax32, a);
        return *slot((unsigned)min(a, cap32));
    }
    __device__ __forceinline__ void st(int a, R v)
    {
        amax32 = max(amax32, a);
        *slot((unsigned)min(a, cap32)) = v;
    }
    __device__ __forceinline__ R ld(R base, R idx) { return ldi(addr(base, idx)); }
    __device__ __forceinline__ void memset(R dest, R value, R len)                         // :5439-5495
    {
        long long d = (long long)__dadd_rn((double)dest, 1.0e-5);
        if (d < 0) d = 0;
        long long n = (long long)len;
        if (n < 0) n = 0;
        if (d + n > cap) { *err = 1; n = cap > d ? cap - d : 0; }
        for (long long i = 0; i < n; ++i) mem[(d + i) * stride] = value;
    }
    __device__ __forceinline__ void memcpy(R dest, R src, R len)                           // overlap permitted (memmove), returns 0
    {
        long long d = (long long)__dadd_rn((double)dest, 1.0e-5), s = (long long)__dadd_rn((double)src, 1.0e-5);
        if (d < 0) d = 0;
        if (s < 0) s = 0;
        long long n = (long long)len;
        if (n < 0) n = 0;
        if (d + n > cap || s + n > cap) { *err = 1; return; }
        if (d <= s) for (long long i = 0; i < n; ++i) mem[(d + i) * stride] = mem[(s + i) * stride];
        else        for (long long i = n - 1; i >= 0; --i) mem[(d + i) * stride] = mem[(s + i) * stride];
    }
    __device__ __forceinline__ R spl_dyn(R i) const { const int k = to_i32(i); return (k >= 0 && k < 64) ? spl[k] : (R)0; }
    __device__ __forceinline__ R slider_dyn(R i) const { const int k = to_i32(i) - 1; return (k >= 0 && k < 64) ? sliders[k] : (R)0; }

    // MT19937, seed 0x4141F00D, lazily initialised while the index is 0 (:3880-4060); rand([max]) = genrand_int32 *
    // (1 / 0xFFFFFFFF) * max(1, floor(arg)) (:5294-5323)
    // (static: the context object itself never has its address taken, so its arrays can live in registers)
    // Slow path, out of line: lazy seeding (stored index 0) and the block regeneration (stored index 624).  The regeneration
    // runs in groups of 24 words with all 49 loads of a group issued before its stores (word k needs the OLD words k, k + 1 and
    // word k + 397, which is either old or was stored at least 227 words earlier): one memory latency per group instead of
    // three per word -- with the one-word-at-a-time loop the state column in L2 made rand() 40 % of SOMA's frame time.
    static __device__ __noinline__ void mt_refill(unsigned* mt, long long stride, unsigned idx)
    {
        const int N = 624, M = 397;
        if (idx == 0) {
            unsigned prev = 0x4141F00Du;
            mt[0] = prev;
            for (int k = 1; k < N; ++k) { prev = 1812433253u * (prev ^ (prev >> 30)) + (unsigned)k; mt[k * stride] = prev; }
        }
        constexpr int G = 24;                              // 624 = 26 x 24
        for (int k0 = 0; k0 < N; k0 += G) {
            unsigned a[G + 1], m[G];
#pragma unroll
            for (int j = 0; j < G + 1; ++j) a[j] = mt[((k0 + j) % N) * stride];
#pragma unroll
            for (int j = 0; j < G; ++j) m[j] = mt[((k0 + j + M) % N) * stride];
#pragma unroll
            for (int j = 0; j < G; ++j) {
                const unsigned y = (a[j] & 0x80000000u) | (a[j + 1] & 0x7fffffffu);
                mt[(k0 + j) * stride] = m[j] ^ (y >> 1) ^ ((y & 1u) ? 0x9908B0DFu : 0u);
            }
        }
    }
    // Fast path: the index lives in a register for the length of a launch (the stored copy at word 624 is kept up to date).
    // Stage pipelines (mtw != nullptr, flag pf) also keep a 16-word window of the column in shared memory, filled by 4-byte
    // cp.async eight draws ahead: one commit group per draw (empty past word 623), so `wait_group 8` is exactly "the word of this
    // draw has landed" and a draw never waits for L2.  (Measured alternatives: words held in registers make the draw wait on
    // the loads in flight as soon as they are shifted or selected; prefetch.global.L1 eight draws ahead changed nothing.)
    unsigned mt_idx;
    unsigned* mtw;                                         // this lane's column of the window: word k at mtw[(k & 15) * 32]
    __device__ __forceinline__ void mt_fetch(unsigned k)
    {
#ifndef ZA_JSFX_HOST_EMU
        if (k < 624u) {
            const unsigned d = (unsigned)__cvta_generic_to_shared(mtw + (k & 15u) * 32u);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(mt + (long long)k * stride) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
#else
        (void)k;
#endif
    }
    __device__ __forceinline__ unsigned genrand()
    {
        const unsigned N = 624, kAhead = 8;
        bool fill = false;
        if (mt_idx == 0xFFFFFFFFu) { mt_idx = mt[624 * stride]; fill = true; }        // first draw of this launch
        if (mt_idx == 0 || mt_idx >= N) { mt_refill(mt, stride, mt_idx); mt_idx = 0; fill = true; }
        unsigned y;
#ifndef ZA_JSFX_HOST_EMU
        if (pf) {
            if (fill) {
                __threadfence_block();
                for (unsigned d = 0; d < kAhead; ++d) mt_fetch(mt_idx + d);
            }
            mt_fetch(mt_idx + kAhead);
            asm volatile("cp.async.wait_group 8;" ::: "memory");
            y = mtw[(mt_idx & 15u) * 32u];
        } else
#endif
        {
            (void)fill; (void)kAhead;
            y = mt[mt_idx * stride];
        }
        mt_idx += 1;
        mt[624 * stride] = mt_idx;
        y ^= y >> 11; y ^= (y << 7) & 0x9D2C5680u; y ^= (y << 15) & 0xEFC60000u; y ^= y >> 18;
        return y;
    }
    __device__ __forceinline__ R rand(R arg, int has_arg)
    {
        R mx = has_arg ? m_floor(arg) : (R)1;
        if (mx < (R)1) mx = (R)1;
        const double r = __dmul_rn((double)genrand(), 1.0 / 4294967295.0);
        return (R)__dmul_rn(r, (double)mx);
    }
};

// stage pipelines (jsfx_staged_kernel below): the plan a generated unit carries, and the batch counter its stages index rings with
struct NoStages { static constexpr int kStages = 1, kMailSlots = 1; };
// batch number of a stage modulo every ring depth a plan can ask for (2 .. 9), kept incrementally
struct BatchMod {
    int r2 = 0, r3 = 0, r4 = 0, r5 = 0, r6 = 0, r7 = 0, r8 = 0, r9 = 0;
    __device__ __forceinline__ void next()
    {
        r2 = r2 + 1 == 2 ? 0 : r2 + 1; r3 = r3 + 1 == 3 ? 0 : r3 + 1; r4 = r4 + 1 == 4 ? 0 : r4 + 1; r5 = r5 + 1 == 5 ? 0 : r5 + 1;
        r6 = r6 + 1 == 6 ? 0 : r6 + 1; r7 = r7 + 1 == 7 ? 0 : r7 + 1; r8 = r8 + 1 == 8 ? 0 : r8 + 1; r9 = r9 + 1 == 9 ? 0 : r9 + 1;
    }
};

// state rows per stream (all R, [row][n_streams]): the script's variables, then st.sliders[64], then lastSlidersValid
template <class S> constexpr int state_rows() { return S::kVars + 64 + 1; }

template <class S, typename R>
__device__ __forceinline__ void ctx_open(Rt<R>& c, R* v, const GenLaunch& a, int sc)
{
    const long long ns = a.n_streams;
    c.mem = reinterpret_cast<R*>(a.mem) + sc; c.stride = ns; c.cap = a.mem_slots;
    c.mt = a.mt + sc; c.mt_idx = 0xFFFFFFFFu; c.mtw = nullptr; c.pf = 0; c.err = a.err; c.amax = -1; c.amax32 = -1;
    c.cap32 = (int)a.mem_slots; c.stride_bytes = (unsigned)a.n_streams * (unsigned)sizeof(R);
    c.srate = (R)a.srate; c.samplesblock = (R)0; c.pending = 0;
    c.scr = a.mem_slots + 1; c.scr_n = a.fft_scratch; c.tw = a.tw;       // scratch sits behind the spare slot out-of-range accesses land in
    const R* st = reinterpret_cast<const R*>(a.state) + sc;
#pragma unroll
    for (int k = 0; k < S::kVars; ++k) v[k] = st[(long long)k * ns];
#pragma unroll
    for (int k = 0; k < 64; ++k) { c.sliders[k] = st[(long long)(S::kVars + k) * ns]; c.spl[k] = (R)0; }
}
template <class S, typename R>
__device__ __forceinline__ void ctx_close(const Rt<R>& c, const R* v, const GenLaunch& a, int s, R valid)
{
    const long long ns = a.n_streams;
    R* st = reinterpret_cast<R*>(a.state) + s;
#pragma unroll
    for (int k = 0; k < S::kVars; ++k) st[(long long)k * ns] = v[k];
#pragma unroll
    for (int k = 0; k < 64; ++k) st[(long long)(S::kVars + k) * ns] = c.sliders[k];
    st[(long long)(S::kVars + 64) * ns] = valid;
    if (c.amax >= c.cap || c.amax32 >= (int)c.cap) *c.err = 1;
}

// push one row of quantised slider values (pushParamsToStateSliders, :9286-9357): returns "changed"
template <class S, typename R>
__device__ __forceinline__ bool push_sliders(Rt<R>& c, R* v, const GenLaunch& a, int row, int sc, bool valid)
{
    bool changed = !valid;
    const double* q = a.q + ((size_t)row * a.n_streams + sc) * S::kDecl;
#pragma unroll
    for (int j = 0; j < S::kDecl; ++j) {             // compile-time slider slots: sliders[] / vars stay in registers
        const int i0 = S::decl(j);
        const R nv = (R)q[j];
        if (nv != c.sliders[i0]) changed = true;
        c.sliders[i0] = nv;
        const int al = S::alias(i0);
        if (al >= 0) v[al] = nv;
    }
    return changed;
}

#ifndef ZA_JSFX_HOST_EMU
template <class S, typename R>
__global__ void __launch_bounds__(32) jsfx_prepare_kernel(const GenLaunch a)
{
    const int s = blockIdx.x * 32 + threadIdx.x;
    if (s >= a.n_streams) return;
    Rt<R> c;
    R v[S::kVars];
    ctx_open<S, R>(c, v, a, s);                    // state and heap were zeroed by the host (resetStateStructOnly)
    push_sliders<S, R>(c, v, a, 0, s, false);
    S::template init<R>(c, v);
#pragma unroll
    for (int j = 0; j < S::kDecl; ++j) {           // alias re-apply after @init (:3308-3316)
        const int al = S::alias(S::decl(j));
        if (al >= 0) v[al] = c.sliders[S::decl(j)];
    }
    S::template slider<R>(c, v);
    ctx_close<S, R>(c, v, a, s, (R)1);
}

// CTA shape of jsfx_process_kernel: independent warps of 32 streams each.
//  * Leaves without a mem[] heap and at most four channels: FOUR warps per CTA.  The warps of one CTA are spread over the
//    SM's four schedulers, whereas single-warp CTAs were all placed on the same one (measured with RTT: solo time flat
//    up to 2 warps per SM, then growing linearly -- 24.7 ms at 8192 streams, 72 ms at 37 888; with 4-warp CTAs flat up
//    to 4 warps per SM).
//  * Leaves WITH a heap: one warp per CTA.  Their hot heap lines (SpectralStabilizer: ~34 KB per warp) live in L1,
//    which shares the SM's 228 KB with the audio tiles; four such warps on one SM thrash it (SpectralStabilizer
//    8192 streams x 0.5 s: 113 ms in 1-warp CTAs, 245 ms in 4-warp CTAs).
//  * Leaves with more than four channels (RTT, 8 inputs): one warp per CTA, because four warps' double-buffered
//    32-frame tiles do not fit an SM and 16-frame tiles measured 35-45 % slower.
template <class S> struct GenWarps { static constexpr int value = (S::kMemSlots == 0 && S::kCh <= 4) ? 4 : 1; };
// Frames per audio tile: 32 (128-byte row segments).  16 (half the shared memory, two rows per copy instruction) was
// measured too: RTT / ATTACK lose 35-45 % alone, SpectralStabilizer 5 %, ERBTilt 5 %, Roomalizer 16 %, and the mixed
// cfg5 batch does not gain from the smaller tiles.
template <class S> struct TileFrames { static constexpr int value = 32; };

// 4-byte asynchronous global -> shared copy (LDGSTS); `ok` false = write zero without reading
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gmem_src, bool ok)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    const int n = ok ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(gmem_src), "r"(n) : "memory");
}

template <class S, typename R, bool TAP>
__global__ void __launch_bounds__(128, 1) jsfx_process_kernel(const GenLaunch a)
{
    // Audio tiles, [2 buffers][kCh][32 streams][kF frames + 1]: the tile after the one being computed is already on
    // its way from HBM (cp.async), so the only warp of the SM does not sit out a DRAM latency per tile (ncu, profiles/
    // r01_generic_rtt_ncu.md: 26 % of the samples were the tile-in stores waiting for their loads).  kF: see TileFrames.
    extern __shared__ float tiles_cta[];
    // Streams per warp L = a.lanes (8, 16 or 32): a stream's @sample is one long dependent chain, so a warp with 8 busy
    // lanes runs as fast as one with 32; at a few thousand streams the host spreads them over 4x as many warps (one per SM
    // scheduler instead of one per SM).  All 32 lanes move the audio tiles either way.
    constexpr int kF = TileFrames<S>::value, kP = kF + 1, kRows = 32 / kF;   // rows per copy instruction
    const int L = a.lanes, kTile = S::kCh * L * kP;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, s0 = (blockIdx.x * (blockDim.x >> 5) + warp) * L, s = s0 + lane;
    if (s0 >= a.n_streams) return;                 // the warps of a CTA never synchronise with each other
    float* const tiles = tiles_cta + warp * 2 * kTile;
    const bool live = lane < L && s < a.n_streams; // dead lanes only help move the audio tiles
    const int ns_cta = min(L, a.n_streams - s0);
    const int half = lane / kF, fl = lane % kF;    // tile I/O: kF lanes per row, lane = frame
    Rt<R> c;
    R v[S::kVars];
    bool valid = false;
    if (live) {
        ctx_open<S, R>(c, v, a, s);
        valid = reinterpret_cast<const R*>(a.state)[(long long)(S::kVars + 64) * a.n_streams + s] != (R)0;
    }

    auto fetch = [&](float* tile, int64_t frame0, int nf) {
        const int nq = ns_cta * S::kIn;
        const float* src = a.in + (size_t)s0 * S::kIn * a.row_stride + frame0 + (fl < nf ? fl : 0);
        for (int q0 = 0; q0 < nq; q0 += kRows) {
            const int q = q0 + half;
            if (q < nq) cp_async4(tile + ((q % S::kIn) * L + q / S::kIn) * kP + fl, src + (size_t)q * a.row_stride, fl < nf);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    int64_t pos = 0;
    int nb = (int)min((int64_t)a.block_len, a.n_frames - pos);
    int buf = 0;
    if (a.n_frames > 0) fetch(tiles, 0, min(kF, nb));
    for (int blk = 0; pos < a.n_frames; ++blk) {
        if (live) {
            if (a.n_rows > 0) {
                const bool changed = push_sliders<S, R>(c, v, a, blk < a.n_rows ? blk : a.n_rows - 1, s, valid);
                valid = true;
                if (changed) S::template slider<R>(c, v);
            }
            c.samplesblock = (R)nb;
            c.pending = 0;                                   // cleared by the host after every block
            S::template block<R>(c, v);
            if (c.pending) S::template slider<R>(c, v);      // @block changed sliders itself: @slider once before @sample
        }
        const int nb_next = (int)min((int64_t)a.block_len, a.n_frames - (pos + nb));      // <= 0 after the last block
        for (int f0 = 0; f0 < nb; f0 += kF) {
            const int nf = min(kF, nb - f0);
            float* tile = tiles + buf * kTile;
            // the next tile of this block, else the first of the next block
            const bool more_here = f0 + kF < nb;
            const int64_t nx0 = more_here ? pos + f0 + kF : pos + nb;
            const int nxf = more_here ? min(kF, nb - f0 - kF) : min(kF, nb_next);
            if (nxf > 0) {
                fetch(tiles + (buf ^ 1) * kTile, nx0, nxf);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            __syncwarp();
            if (live) {
                for (int f = 0; f < nf; ++f) {
#pragma unroll
                    for (int ch = 0; ch < S::kCh; ++ch) c.spl[ch] = ch < S::kIn ? (R)tile[(ch * L + lane) * kP + f] : (R)0;
                    S::template sample<R>(c, v);
#pragma unroll
                    for (int ch = 0; ch < S::kOut; ++ch) {
                        if (TAP) a.tap[((size_t)s * S::kOut + ch) * a.row_stride + pos + f0 + f] = (double)c.spl[ch];
                        tile[(ch * L + lane) * kP + f] = (float)c.spl[ch];
                    }
                }
            }
            __syncwarp();
            {
                const int nq = ns_cta * S::kOut;
                float* dst = a.out + (size_t)s0 * S::kOut * a.row_stride + pos + f0 + fl;
#pragma unroll 4
                for (int q0 = 0; q0 < nq; q0 += kRows) {
                    const int q = q0 + half;
                    if (q < nq && fl < nf) dst[(size_t)q * a.row_stride] = tile[((q % S::kOut) * L + q / S::kOut) * kP + fl];
                }
            }
            __syncwarp();
            buf ^= 1;
        }
        pos += nb;
        nb = nb_next;
    }
    if (live) ctx_close<S, R>(c, v, a, s, valid ? (R)1 : (R)0);
}

// ---- stage pipeline (leaves the translator could cut, jsfxc.py::plan_stages) --------------------------------------------
// One CTA = 32 streams x K warps: warp k runs statements [cut k, cut k+1) of @sample (S::Staged::run<R, F, k>) for frame batch
// i - k while warp k + 1 runs the next group on batch i - k - 1 -- the EasyExpander role pipeline, generated.  A stream's
// time axis stays sequential, but the dependent chain of one frame is spread over K warps that work at the same time.
// Every warp keeps its OWN copy of the variables and replays the host-block protocol (slider push, @slider, @block) when
// its own frame counter enters a block; the planner has checked that those sections do not read what @sample writes.  What
// crosses a cut inside a frame travels through per-value rings in shared memory (mb[slot][F][32], ring depth = distance to
// the furthest consumer + 1), one __syncthreads per batch of F frames.  Same statements in the same order on the same
// values: in float64 the outputs and states are bit-identical to jsfx_process_kernel.
// shared memory of a CTA: [rings][audio tiles: 2 x kIn in, kOut out][MT19937 window]
// frames per audio tile.  (16 for the 8-input RTT, which leaves its rings room for 5-frame instead of 2-frame batches, was
// measured: 45.7 ms against 27.2 ms per 5041 x 1 s -- the first stage's warp pays for twice the copy instructions.)
template <class S> __host__ __device__ constexpr int staged_tile_frames() { return 32; }
// input tiles: two (the next one on its way while this one is computed), or ONE where many input channels (RTT: 8) would leave
// the rings room for 2-frame batches only -- the first stage then waits out one copy latency per 32 frames, which it has time for
template <class S> __host__ __device__ constexpr int staged_in_tiles() { return S::kIn > 4 ? 1 : 2; }
template <class S> __host__ __device__ constexpr size_t staged_fixed_smem()
{
    return (size_t)(staged_in_tiles<S>() * S::kIn + S::kOut) * 32 * (staged_tile_frames<S>() + 1) * sizeof(float) + 16 * 32 * sizeof(unsigned);
}
// CTAs per SM the stage pipelines are compiled for (register cap): two for five and more stages.  (Three -- 85-113 registers
// per thread, with a 72 KB shared-memory budget -- is a measurement build: -DZA_STAGED_MIN_CTAS\(K\)=3.)
#ifndef ZA_STAGED_MIN_CTAS
#define ZA_STAGED_MIN_CTAS(K) ((K) > 4 ? 2 : 1)
#endif
// Shared-memory budget of a stage-pipeline CTA (two CTAs per SM at 110 KB).  72 KB was measured too (more room for other
// kernels' CTAs in the cfg5 mix): the mixed step stays at 149-150 ms, SOMA alone goes 27.8 -> 30.4 ms with its shorter batches.
#ifndef ZA_STAGED_SMEM_KB
#define ZA_STAGED_SMEM_KB 110
#endif
constexpr size_t kStagedSmemBudget = (size_t)ZA_STAGED_SMEM_KB * 1024;
template <class S, typename R> __host__ __device__ constexpr int staged_frames()     // frames per batch: two CTAs per SM (110 KB each)
{
    int f = 32;
    while (f > 2 && (size_t)S::Staged::kMailSlots * f * 32 * sizeof(R) + staged_fixed_smem<S>() > kStagedSmemBudget) --f;
    return f;
}
template <class S, typename R> __host__ __device__ constexpr size_t staged_smem()
{
    return (size_t)S::Staged::kMailSlots * staged_frames<S, R>() * 32 * sizeof(R) + staged_fixed_smem<S>();
}

template <class S, typename R, bool TAP, int K>
__device__ __forceinline__ void staged_role(const GenLaunch& a, R* mb, float* tiles, int lane, int s0)
{
    using P = typename S::Staged;
    constexpr int KS = P::kStages, F = staged_frames<S, R>(), kF = staged_tile_frames<S>(), kP = kF + 1, kTile = S::kIn * 32 * kP, kRows = 32 / kF;
    const int half = lane / kF, fl = lane % kF;                  // tile I/O: kF lanes per row, lane = frame
    constexpr bool first = K == 0, last = K == KS - 1;
    const int s = s0 + lane;
    const bool live = s < a.n_streams;
    const int ns_cta = min(32, a.n_streams - s0);
    constexpr int kInTiles = staged_in_tiles<S>();
    float* const out_tile = tiles + kInTiles * kTile;
    Rt<R> c;
    R v[S::kVars];
    bool valid = false;
    if (live) {
        ctx_open<S, R>(c, v, a, s);
        c.pf = 1;
        c.mtw = reinterpret_cast<unsigned*>(tiles + (kInTiles * S::kIn + S::kOut) * 32 * kP) + lane;
        valid = reinterpret_cast<const R*>(a.state)[(long long)(S::kVars + 64) * a.n_streams + s] != (R)0;
    }
    auto fetch = [&](float* tile, int64_t frame0, int nf) {
        const int nq = ns_cta * S::kIn;
        const float* src = a.in + (size_t)s0 * S::kIn * a.row_stride + frame0 + (fl < nf ? fl : 0);
        for (int q0 = 0; q0 < nq; q0 += kRows) {
            const int q = q0 + half;
            if (q < nq) cp_async4(tile + ((q % S::kIn) * 32 + q / S::kIn) * kP + fl, src + (size_t)q * a.row_stride, fl < nf);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int j = 0; j < K; ++j) __syncthreads();                 // pipeline fill: batch 0 reaches this stage at iteration K
    int64_t pos = 0;
    int nb = (int)min((int64_t)a.block_len, a.n_frames - pos);
    int buf = 0;
    BatchMod bt;
    if (first && kInTiles == 2 && a.n_frames > 0) fetch(tiles, 0, min(kF, nb));
    for (int blk = 0; pos < a.n_frames; ++blk) {
        if (live) {
            if (a.n_rows > 0) {
                const bool changed = push_sliders<S, R>(c, v, a, blk < a.n_rows ? blk : a.n_rows - 1, s, valid);
                valid = true;
                if (changed) S::template slider<R>(c, v);
            }
            c.samplesblock = (R)nb;
            c.pending = 0;
            S::template block<R>(c, v);
            if (c.pending) S::template slider<R>(c, v);
        }
        const int nb_next = (int)min((int64_t)a.block_len, a.n_frames - (pos + nb));
        for (int f0 = 0; f0 < nb; f0 += kF) {
            const int nf = min(kF, nb - f0);
            const float* tile = tiles + buf * kTile;
            if (first && kInTiles == 1) {
                fetch(tiles, pos + f0, nf);
                asm volatile("cp.async.wait_group 0;" ::: "memory");
                __syncwarp();
            } else if (first) {
                const bool more_here = f0 + kF < nb;
                const int64_t nx0 = more_here ? pos + f0 + kF : pos + nb;
                const int nxf = more_here ? min(kF, nb - f0 - kF) : min(kF, nb_next);
                if (nxf > 0) {
                    fetch(tiles + (buf ^ 1) * kTile, nx0, nxf);
                    asm volatile("cp.async.wait_group 1;" ::: "memory");
                } else {
                    asm volatile("cp.async.wait_group 0;" ::: "memory");
                }
                __syncwarp();
            }
            for (int b0 = 0; b0 < nf; b0 += F) {
                const int nbf = min(F, nf - b0);
                if (live) {
                    for (int f = 0; f < nbf; ++f) {
                        if (first) {
#pragma unroll
                            for (int ch = 0; ch < S::kCh; ++ch) c.spl[ch] = ch < S::kIn ? (R)tile[(ch * 32 + lane) * kP + b0 + f] : (R)0;
                        }
                        P::template run<R, F, K>(c, v, mb, bt, f * 32 + lane);
                        if (last) {
#pragma unroll
                            for (int ch = 0; ch < S::kOut; ++ch) {
                                if (TAP) a.tap[((size_t)s * S::kOut + ch) * a.row_stride + pos + f0 + b0 + f] = (double)c.spl[ch];
                                out_tile[(ch * 32 + lane) * kP + b0 + f] = (float)c.spl[ch];
                            }
                        }
                    }
                }
                bt.next();
                __syncthreads();
            }
            if (last) {
                __syncwarp();
                const int nq = ns_cta * S::kOut;
                float* dst = a.out + (size_t)s0 * S::kOut * a.row_stride + pos + f0 + fl;
                if (fl < nf) {
#pragma unroll 4
                    for (int q0 = 0; q0 < nq; q0 += kRows) {
                        const int q = q0 + half;
                        if (q < nq) dst[(size_t)q * a.row_stride] = out_tile[((q % S::kOut) * 32 + q / S::kOut) * kP + fl];
                    }
                }
                __syncwarp();
            }
            if (first) { __syncwarp(); if (kInTiles == 2) buf ^= 1; }
        }
        pos += nb;
        nb = nb_next;
    }
    for (int j = K; j < KS - 1; ++j) __syncthreads();            // drain
    if (live) {                                                  // every variable from the stage that owns it; the rest from stage 0
        R* st = reinterpret_cast<R*>(a.state) + s;
        const long long ns = a.n_streams;
#pragma unroll
        for (int k = 0; k < S::kVars; ++k)
            if (P::owner(k) == K) st[(long long)k * ns] = v[k];
        if (first) {
#pragma unroll
            for (int k = 0; k < 64; ++k) st[(long long)(S::kVars + k) * ns] = c.sliders[k];
            st[(long long)(S::kVars + 64) * ns] = valid ? (R)1 : (R)0;
        }
        if (c.amax >= c.cap || c.amax32 >= (int)c.cap) *c.err = 1;
    }
}

template <class S, typename R, bool TAP, int K>
__device__ __forceinline__ void staged_pick(const GenLaunch& a, R* mb, float* tiles, int warp, int lane, int s0)
{
    if constexpr (K < S::Staged::kStages) {
        if (warp == K) staged_role<S, R, TAP, K>(a, mb, tiles, lane, s0);
        else staged_pick<S, R, TAP, K + 1>(a, mb, tiles, warp, lane, s0);
    }
}

template <class S, typename R, bool TAP>
__global__ void __launch_bounds__(32 * S::Staged::kStages, ZA_STAGED_MIN_CTAS(S::Staged::kStages)) jsfx_staged_kernel(const GenLaunch a)
{
    extern __shared__ __align__(16) unsigned char staged_smem_raw[];
    R* mb = reinterpret_cast<R*>(staged_smem_raw);
    float* tiles = reinterpret_cast<float*>(staged_smem_raw + (size_t)S::Staged::kMailSlots * staged_frames<S, R>() * 32 * sizeof(R));
    staged_pick<S, R, TAP, 0>(a, mb, tiles, threadIdx.x >> 5, threadIdx.x & 31, blockIdx.x * 32);
}

template <class S> bool launch_staged(const GenLaunch& a, int precision, cudaStream_t st)
{
    const int blocks = (a.n_streams + 31) / 32;
    auto go = [&](auto kern, size_t smem) {
        if (smem > 48 * 1024 && !cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "jsfx_staged_kernel smem attr")) return false;
        kern<<<blocks, 32 * S::Staged::kStages, smem, st>>>(a);
        return cuda_ok(cudaGetLastError(), "jsfx_staged_kernel launch");
    };
    if (precision) return a.tap ? go(jsfx_staged_kernel<S, double, true>, staged_smem<S, double>()) : go(jsfx_staged_kernel<S, double, false>, staged_smem<S, double>());
    return go(jsfx_staged_kernel<S, float, false>, staged_smem<S, float>());
}

template <class S> bool launch_prepare(const GenLaunch& a, int precision, cudaStream_t st)
{
    const int blocks = (a.n_streams + 31) / 32;
    if (precision) jsfx_prepare_kernel<S, double><<<blocks, 32, 0, st>>>(a);
    else jsfx_prepare_kernel<S, float><<<blocks, 32, 0, st>>>(a);
    return cuda_ok(cudaGetLastError(), "jsfx_prepare_kernel launch");
}
template <class S> bool launch_process(const GenLaunch& a, int precision, cudaStream_t st)
{
    // Streams per warp: 32 unless the caller asks for 16 or 8 (option "lanes_per_warp", a measurement switch).  A stream's
    // time axis is sequential, so at a few thousand streams the launch takes frames x (latency of one frame's dependent
    // chain) however the streams are spread over warps and SMs: measured on B200, RTT 4096 streams x 10 s takes 629 ms as
    // 128 one-warp CTAs and as 64 two-warp CTAs, 520 ms with 8 streams per warp (less divergence per warp), and the
    // concurrent cfg5 mix loses 15 % with narrow warps (a warp with 8 live lanes costs a full issue slot).  Only a shorter
    // chain per frame or time-parallel kernels (the scan kernels) make these leaves faster.
    if constexpr (S::Staged::kStages > 1) {
        // The pipeline holds 32 streams per CTA and two CTAs per SM.  Leaves whose one-thread kernel runs four warps per CTA
        // (no heap, at most four channels) catch up once the batch is more than that: measured on B200 (0.5 s, pipeline /
        // one-thread): SaliencePush 9472 streams 10.4 / 14.5 ms, 18 944: 20.8 / 14.5, 37 888: 41.6 / 29.0; ATTACK 8.8 / 17.7,
        // 17.6 / 17.7, 35.1 / 35.3.  RTT (8 inputs) and SOMA (heap) stay 1.5-2.8x ahead up to 37 888 streams (13.8 / 27.8 ...
        // 55.2 / 81.5; 13.4 / 37.4 ... 55.0 / 86.1).  Option "unstaged": 0 = this rule, 1 = one-thread kernel, 2 = pipeline.
        bool staged = a.unstaged != 1;
        if (a.unstaged == 0 && GenWarps<S>::value == 4) {
            int dev = 0, n_sm = 148;
            if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
            staged = a.n_streams <= 2 * 32 * n_sm;
        }
        if (staged) return launch_staged<S>(a, precision, st);
    }
    GenLaunch b = a;
    int L = a.lanes, W = GenWarps<S>::value;
    auto smem_of = [&](int w, int l) { return (size_t)w * 2 * S::kCh * l * (TileFrames<S>::value + 1) * sizeof(float); };
    if (L != 8 && L != 16) L = 32;
    // 32 streams per warp: the round-1 shapes (GenWarps).  Narrower warps: four per CTA when the tiles fit.
    if (L != 32 && S::kMemSlots == 0) { W = 4; while (W > 1 && smem_of(W, L) > 160 * 1024) W >>= 1; }
    b.lanes = L;
    const int kGenWarps = W;
    const int blocks = (a.n_streams + L * kGenWarps - 1) / (L * kGenWarps);
    const size_t smem = smem_of(W, L);
    auto go = [&](auto kern) {
        if (smem > 48 * 1024 && !cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "jsfx_process_kernel smem attr")) return false;
        // Leaves with a heap live off the L1 (their hot mem[] lines; SpectralStabilizer ~34 KB per warp).  Left alone, the
        // driver sizes the shared-memory carve-out for the 8 CTAs per SM the register file allows (164 KB), which leaves
        // 64 KB of L1: fine for 2 warps per SM, thrashing beyond (ncu at 16 384 streams: L1 hit rate 98.8 % -> 89.8 %,
        // long_scoreboard 2 % -> 31 %, 2.3x the time per frame).  Ask for a 100 KB carve-out instead: 5 CTAs per SM, 128 KB L1 (64 KB / 3 CTAs measured too: see DESIGN.md 4.5).
        if (S::kMemSlots > 0 && !cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 44), "jsfx_process_kernel carve-out")) return false;
        kern<<<blocks, 32 * kGenWarps, smem, st>>>(b);
        return cuda_ok(cudaGetLastError(), "jsfx_process_kernel launch");
    };
    if (precision) return a.tap ? go(jsfx_process_kernel<S, double, true>) : go(jsfx_process_kernel<S, double, false>);
    return go(jsfx_process_kernel<S, float, false>);
}

#else   // ZA_JSFX_HOST_EMU: the same two entry points for ONE stream on the host
template <class S, typename R> void emu_prepare(const GenLaunch& a)
{
    Rt<R> c;
    std::vector<R> v(S::kVars);
    ctx_open<S, R>(c, v.data(), a, 0);
    push_sliders<S, R>(c, v.data(), a, 0, 0, false);
    S::template init<R>(c, v.data());
    for (int j = 0; j < a.n_decl; ++j) {
        const int al = S::alias(a.decl_index0[j]);
        if (al >= 0) v[al] = c.sliders[a.decl_index0[j]];
    }
    S::template slider<R>(c, v.data());
    ctx_close<S, R>(c, v.data(), a, 0, (R)1);
}
template <class S, typename R> void emu_process(const GenLaunch& a)
{
    Rt<R> c;
    std::vector<R> v(S::kVars);
    ctx_open<S, R>(c, v.data(), a, 0);
    bool valid = reinterpret_cast<const R*>(a.state)[S::kVars + 64] != (R)0;
    int64_t pos = 0;
    for (int blk = 0; pos < a.n_frames; ++blk) {
        const int nb = (int)std::min<int64_t>(a.block_len, a.n_frames - pos);
        if (a.n_rows > 0) {
            const bool changed = push_sliders<S, R>(c, v.data(), a, blk < a.n_rows ? blk : a.n_rows - 1, 0, valid);
            valid = true;
            if (changed) S::template slider<R>(c, v.data());
        }
        c.samplesblock = (R)nb;
        c.pending = 0;
        S::template block<R>(c, v.data());
        if (c.pending) S::template slider<R>(c, v.data());
        for (int f = 0; f < nb; ++f) {
            for (int ch = 0; ch < S::kCh; ++ch) c.spl[ch] = ch < S::kIn ? (R)a.in[ch * a.row_stride + pos + f] : (R)0;
            S::template sample<R>(c, v.data());
            for (int ch = 0; ch < S::kOut; ++ch) {
                if (a.tap) a.tap[ch * a.row_stride + pos + f] = (double)c.spl[ch];
                a.out[ch * a.row_stride + pos + f] = (float)c.spl[ch];
            }
        }
        pos += nb;
    }
    ctx_close<S, R>(c, v.data(), a, 0, valid ? (R)1 : (R)0);
}
#endif

}  // namespace jsfx
}  // namespace za

// Emitted at the end of every generated translation unit.  (A unit with a stage pipeline defines ZA_JSFX_STAGED_IMPL first.)
#ifndef ZA_JSFX_STAGED_IMPL
#define ZA_JSFX_STAGED_IMPL ::za::jsfx::NoStages
#endif
#ifdef ZA_JSFX_HOST_EMU
#define ZA_JSFX_DEFINE_PLUGIN(IDENT, KEY, MEM_SLOTS, HAS_BLOCK, HAS_SLIDER)                                            \
    struct Script {                                                                                                     \
        static constexpr int kVars = g_##IDENT::kVars, kIn = g_##IDENT::kIn, kOut = g_##IDENT::kOut, kCh = g_##IDENT::kCh, kDecl = g_##IDENT::kDecl; \
        static constexpr long long kMemSlots = (MEM_SLOTS);                                                             \
        static constexpr long long kFftScratch = g_##IDENT::kFftScratch;                                                \
        static constexpr int alias(int i0) { return g_##IDENT::alias_slot(i0); }                                        \
        static constexpr int decl(int j) { return g_##IDENT::decl_slot(j); }                                            \
        template <typename R> static void init(Rt<R>& c, R* v) { sec_init<R>(c, v); }                                   \
        template <typename R> static void slider(Rt<R>& c, R* v) { sec_slider<R>(c, v); }                               \
        template <typename R> static void block(Rt<R>& c, R* v) { sec_block<R>(c, v); }                                 \
        template <typename R> static void sample(Rt<R>& c, R* v) { sec_sample<R>(c, v); }                               \
    };                                                                                                                  \
    }  /* namespace g_IDENT */                                                                                          \
    }}                                                                                                                  \
    extern "C" {                                                                                                        \
    int emu_info(int what) { using S = za::jsfx::g_##IDENT::Script; return what == 0 ? S::kVars : what == 1 ? S::kIn : what == 2 ? S::kOut : what == 3 ? (int)(MEM_SLOTS) : (int)S::kFftScratch; } \
    const char* emu_var_name(int k) { return za::jsfx::g_##IDENT::kVarNames[k]; }                                       \
    void emu_prepare(const za::GenLaunch* a, int f64) { using S = za::jsfx::g_##IDENT::Script; if (f64) za::jsfx::emu_prepare<S, double>(*a); else za::jsfx::emu_prepare<S, float>(*a); } \
    void emu_process(const za::GenLaunch* a, int f64) { using S = za::jsfx::g_##IDENT::Script; if (f64) za::jsfx::emu_process<S, double>(*a); else za::jsfx::emu_process<S, float>(*a); } \
    }                                                                                                                   \
    namespace za { namespace jsfx { namespace g_##IDENT {
#else
#define ZA_JSFX_DEFINE_PLUGIN(IDENT, KEY, MEM_SLOTS, HAS_BLOCK, HAS_SLIDER)                                            \
    struct Script {                                                                                                     \
        static constexpr int kVars = g_##IDENT::kVars, kIn = g_##IDENT::kIn, kOut = g_##IDENT::kOut, kCh = g_##IDENT::kCh, kDecl = g_##IDENT::kDecl; \
        static constexpr long long kMemSlots = (MEM_SLOTS);                                                             \
        static constexpr long long kFftScratch = g_##IDENT::kFftScratch;                                                \
        using Staged = ZA_JSFX_STAGED_IMPL;                                                                             \
        static __host__ __device__ constexpr int alias(int i0) { return g_##IDENT::alias_slot(i0); }                    \
        static __host__ __device__ constexpr int decl(int j) { return g_##IDENT::decl_slot(j); }                        \
        template <typename R> static __device__ __forceinline__ void init(Rt<R>& c, R* v) { sec_init<R>(c, v); }       \
        template <typename R> static __device__ __forceinline__ void slider(Rt<R>& c, R* v) { sec_slider<R>(c, v); }   \
        template <typename R> static __device__ __forceinline__ void block(Rt<R>& c, R* v) { sec_block<R>(c, v); }     \
        template <typename R> static __device__ __forceinline__ void sample(Rt<R>& c, R* v) { sec_sample<R>(c, v); }   \
    };                                                                                                                  \
    }  /* namespace g_IDENT */                                                                                          \
    const ::za::GenDesc* desc_##IDENT()                                                                                 \
    {                                                                                                                   \
        using S = g_##IDENT::Script;                                                                                    \
        static const ::za::GenDesc d = {KEY, g_##IDENT::kSliderText, S::kIn, S::kOut, S::kVars, (int64_t)(MEM_SLOTS),   \
                                        g_##IDENT::kVarNames, g_##IDENT::kAlias, &launch_prepare<S>, &launch_process<S>,  \
                                        S::kDecl, g_##IDENT::kDeclIndex0, (int64_t)S::kFftScratch};                    \
        return &d;                                                                                                      \
    }                                                                                                                   \
    namespace g_##IDENT {
#endif
