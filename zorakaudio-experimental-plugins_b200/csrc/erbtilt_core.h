// erbtilt_core.h -- ERB Tilt (plugins/Spectral/ERBTilt/src/ERB Tilt.jsfx) as blocked parallel linear-recurrence scans over
// TIME: one warp renders one stream, lane l owns C consecutive frames of a 32 x C frame tile (lanes.h vocabulary: the same
// text is the CUDA kernel's tile body, k_erbtilt.cu, and the CPU emulation, tools/emu_scan.cpp).
//
// @sample of the script (:381-480) per frame: 15 one-pole low-passes per channel on the SAME input (telescoping bands),
// per band an energy EMA before and after the gain, and on the bands above 2 kHz two envelope followers feeding a
// memoryless soft saturator.  Every recurrence is linear with sample-rate-only coefficients, so per tile and band:
//   * the low-passes and the fast / slow followers are FULL scans (all values are needed): zero-state pass A over the lane's
//     chunk, Kogge-Stone over the 32 chunk end values with tabulated powers, start state = prefix + Q[lane] x carry,
//     pass B = the script's own recurrence in its own operation order (fp64: separate multiply / add, no contraction);
//   * the pre / post energy EMAs (:402-403, :431-432) are read only by @block (:305-337), i.e. at host-block ends, so
//     they are END-STATE-ONLY: a weighted sum of the band powers over the tile (weights 0.5 c (1-c)^k, tabulated) plus
//     W x carry -- 3 FMAs per frame instead of a scan;
//   * the carries of all bands live lane-distributed (lane i = band i) in six registers; tile-to-tile they are the exact
//     sequential states of the last valid lane.
// Float mode: the slow EMAs (c = 6.9e-5) stall in a float recurrence (resolution eps / c = 1e-3 relative: the translated
// float path misses the -100 dBFS criterion by 12 dB through global_gain); the weighted-sum form has no such stall, so
// float mode here meets the criterion with float state.
// Partial tiles are chunk-granular (m = number of valid lanes): with C = 1 that is frame-granular, which is how ragged
// block / call ends are rendered (k_erbtilt.cu runs C = 8 (4 in fp64) tiles, then C = 1 tiles, then one partial one).
#pragma once
#include "lanes.h"

namespace za {
namespace erb {
using namespace lanes;

constexpr int kNB = 16, kNBm1 = 15;

// Host-built constants for one (precision, chunk size): powers are formed in long double from the script's double
// coefficients and rounded once.
template <class T, int C>
struct Tables {
    T lp_a[kNBm1], lp_1ma[kNBm1];        // a = lpA[i] (:153-161), (1 - a) as the script forms it (fp64) / from double (fp32)
    T lp_P[kNBm1][5];                    // (a^C)^(2^k)
    T lp_Q[kNBm1][33];                   // (a^C)^j, j = 0 .. 32
    T f_c, f_P[5], f_Q[33];              // rough_fast (:168): c, ((1-c)^C)^(2^k), ((1-c)^C)^j
    T s_c, s_P[5], s_Q[33];              // rough_slow (:171)
    T e_v[C];                            // energy EMA (:165): 0.5 c (1-c)^(C-1-i)
    T e_W[33];                           // ((1-c)^C)^j
    int guarded[kNB];                    // fcBand[i] > rough_start_hz (:410)
};

template <class T>
struct Carry {                           // lane i = band i
    LV<T> lpL, lpR, envF, envS, preE, postE;
};

template <class T>
struct Block {                           // per host block, uniform unless noted
    LV<T> g;                             // gBand, lane i = band i (@slider recalc_gains, :87-109)
    T guard, gg;                         // slider4 / 100 (:300), global_gain after @block (:336)
};

// Sample-rate constants of @init build_filterbank (:113-175), formed on the host with glibc (double), used by the
// per-block prologue below.
struct BlockConsts {
    double erbBand[kNB], wBand[kNB];
    double fmax, erb_range;
};
struct BlockVars {                       // the script's block-rate variables (kept for za_gpu_read_state)
    double tilt_db, pivot_hz, erb_pivot, slope, comp_mix, guard, Lpre, Lpost, target, blend, alpha, global_gain;
};

// @slider (:298-301, recalc_gains :87-109) + @block loudness compensation (:316-337) for one host block, in double with
// the script's operation order; q = the block's quantised sliders.  Pure functions of the sliders and the carried
// energies, so evaluating @slider every block equals evaluating it on change.
template <class T>
ZA_LN void block_prologue(const BlockConsts& bc, const double* q, double alpha, const Carry<T>& cy, BlockVars& bv, Block<T>& bk)
{
    bv.comp_mix = div_(q[2], 100.0);
    bv.guard = div_(q[3], 100.0);
    bv.tilt_db = q[0];
    bv.pivot_hz = q[1] < 20.0 ? 20.0 : (q[1] > bc.fmax ? bc.fmax : q[1]);
    bv.erb_pivot = mul(mul(21.4, log_(add(mul(0.00437, bv.pivot_hz), 1.0))), 0.4342944819032518);
    bv.slope = div_(mul(2.0, bv.tilt_db), add(bc.erb_range, 0.000001));
    const double slope = bv.slope, erb_pivot = bv.erb_pivot;
    const double* erbBand = bc.erbBand;
    bk.g = lane_generate<T>([=](int l) {
        double gdB = mul(slope, sub(erbBand[l & 15], erb_pivot));
        gdB = gdB < -24.0 ? -24.0 : (gdB > 24.0 ? 24.0 : gdB);
        return (T)exp_(mul(gdB, 0.11512925464970229));
    });
    double Lpre = 0.0, Lpost = 0.0;
    for (int i = 0; i < kNB; ++i) {
        Lpre = add(Lpre, mul(bc.wBand[i], (double)lane_value(cy.preE, i)));
        Lpost = add(Lpost, mul(bc.wBand[i], (double)lane_value(cy.postE, i)));
    }
    bv.Lpre = Lpre; bv.Lpost = Lpost;
    double target = sqrt_(div_(add(Lpre, 0.0000000001), add(Lpost, 0.0000000001)));
    target = target < 0.25 ? 0.25 : (target > 4.0 ? 4.0 : target);
    bv.target = target;
    bv.blend = add(1.0, mul(bv.comp_mix, sub(target, 1.0)));
    bv.alpha = alpha;
    bv.global_gain = add(mul(bv.global_gain, alpha), mul(bv.blend, sub(1.0, alpha)));
    bk.guard = (T)bv.guard;
    bk.gg = (T)bv.global_gain;
}

// Full scan of y_n = a y_{n-1} + v_n (in: v, out: y) over the tile; `m` valid lanes; returns the state after the last valid
// frame.  Pass B order: (the script's) t3 = a * y; y = v + t3  -- callers pass v = (1 - a) * x.
template <class T, int C>
ZA_LN T lp_tile(LV<T> (&v)[C], T a, const T* P, const LV<T>& Q, T carry, int m)
{
    LV<T> e = v[0];
#pragma unroll
    for (int i = 1; i < C; ++i) e = fma_rn(e, a, v[i]);
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const LV<T> t = shfl_up(e, 1 << k);
        e = select(lane_ge(1 << k), fma_rn(t, P[k], e), e);
    }
    LV<T> y = fma_rn(Q, carry, select(lane_ge(1), shfl_up(e, 1), T(0)));
#pragma unroll
    for (int i = 0; i < C; ++i) { y = add(v[i], mul(a, y)); v[i] = y; }
    return lane_value(y, m - 1);
}

// Full scan of y_n = y_{n-1} + (x_n - y_{n-1}) c (the script's follower form, :415-416); in: x, out: y.
template <class T, int C>
ZA_LN T ema_tile(const LV<T> (&x)[C], LV<T> (&out)[C], T c, const T* P, const LV<T>& Q, T carry, int m)
{
    LV<T> e = mul(x[0], c);
#pragma unroll
    for (int i = 1; i < C; ++i) e = fma_rn(sub(x[i], e), c, e);
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const LV<T> t = shfl_up(e, 1 << k);
        e = select(lane_ge(1 << k), fma_rn(t, P[k], e), e);
    }
    LV<T> y = fma_rn(Q, carry, select(lane_ge(1), shfl_up(e, 1), T(0)));
#pragma unroll
    for (int i = 0; i < C; ++i) { y = add(y, mul(sub(x[i], y), c)); out[i] = y; }
    return lane_value(y, m - 1);
}

template <class T>
ZA_LN T warp_sum(LV<T> v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = add(v, shfl_down(v, d));
    return lane_value(v, 0);
}

// One tile.  IO: load(xL, xR) / store(yL, yR) of this lane's C frames.  m = valid lanes (1 .. 32; lanes >= m hold zeros).
template <class T, int C, bool FASTDIV, class IO>
ZA_LN void tile(const Tables<T, C>& tb, Carry<T>& cy, const Block<T>& bk, int m, IO& io)
{
    using V = LV<T>;
    V xL[C], xR[C], oL[C], oR[C], pL[C], pR[C];
    io.load(xL, xR);
#pragma unroll
    for (int i = 0; i < C; ++i) { oL[i] = lane_zero<T>(); oR[i] = oL[i]; pL[i] = oL[i]; pR[i] = oL[i]; }
    // weights of the end-state-only energy EMAs for this tile: lane l contributes with ((1-c)^C)^(m-1-l)
    const V laneW = lane_load_rev<T>(tb.e_W, m);
    V yL[C], yR[C];
#pragma unroll 1
    for (int band = 0; band < kNB; ++band) {
        V tL[C], tR[C];
        if (band < kNBm1) {                                                     // :386-400
            const T a = tb.lp_a[band], oma = tb.lp_1ma[band];
#pragma unroll
            for (int i = 0; i < C; ++i) { tL[i] = mul(oma, xL[i]); tR[i] = mul(oma, xR[i]); }
            const V q = lane_load<T>(tb.lp_Q[band]);
            lane_set(cy.lpL, band, lp_tile<T, C>(tL, a, tb.lp_P[band], q, lane_value(cy.lpL, band), m));
            lane_set(cy.lpR, band, lp_tile<T, C>(tR, a, tb.lp_P[band], q, lane_value(cy.lpR, band), m));
        } else {                                                                // :436-438  last band = input - last LP
#pragma unroll
            for (int i = 0; i < C; ++i) { tL[i] = xL[i]; tR[i] = xR[i]; }
        }
        const T g = lane_value(bk.g, band);
        V accPre = lane_zero<T>(), accPost = accPre;
#pragma unroll
        for (int i = 0; i < C; ++i) {
            const V bL = sub(tL[i], pL[i]), bR = sub(tR[i], pR[i]);              // :392-393 (band 0: lp - 0)
            pL[i] = tL[i]; pR[i] = tR[i];
            accPre = fma_rn(fma_rn(bR, bR, mul(bL, bL)), tb.e_v[i], accPre);     // :402-403, end state only
            yL[i] = mul(bL, g); yR[i] = mul(bR, g);                              // :406-408
        }
        if (tb.guarded[band]) {                                                 // :410-424
            V amp[C], eF[C], eS[C];
#pragma unroll
            for (int i = 0; i < C; ++i) amp[i] = mul(T(0.5), add(abs_(yL[i]), abs_(yR[i])));
            const V qf = lane_load<T>(tb.f_Q), qs = lane_load<T>(tb.s_Q);
            lane_set(cy.envF, band, ema_tile<T, C>(amp, eF, tb.f_c, tb.f_P, qf, lane_value(cy.envF, band), m));
            lane_set(cy.envS, band, ema_tile<T, C>(amp, eS, tb.s_c, tb.s_P, qs, lane_value(cy.envS, band), m));
#pragma unroll
            for (int i = 0; i < C; ++i) {
                V mod = FASTDIV ? div_fast_(sub(eF[i], eS[i]), add(eS[i], T(0.000001))) : div_(sub(eF[i], eS[i]), add(eS[i], T(0.000001)));
                mod = select(lt(mod, T(0)), T(0), mod);
                const V sat = mul(mul(bk.guard, mod), T(6));
                const V dl = add(T(1), mul(sat, abs_(yL[i]))), dr = add(T(1), mul(sat, abs_(yR[i])));
                yL[i] = FASTDIV ? div_fast_(yL[i], dl) : div_(yL[i], dl);
                yR[i] = FASTDIV ? div_fast_(yR[i], dr) : div_(yR[i], dr);
            }
        }
#pragma unroll
        for (int i = 0; i < C; ++i) {
            accPost = fma_rn(fma_rn(yR[i], yR[i], mul(yL[i], yL[i])), tb.e_v[i], accPost);   // :431-432, end state only
            if (band < kNBm1) { oL[i] = add(oL[i], yL[i]); oR[i] = add(oR[i], yR[i]); }        // :434-435
        }
        const T sPre = warp_sum<T>(mul(accPre, laneW)), sPost = warp_sum<T>(mul(accPost, laneW));
        lane_set(cy.preE, band, fma_rn(tb.e_W[m], lane_value(cy.preE, band), sPre));
        lane_set(cy.postE, band, fma_rn(tb.e_W[m], lane_value(cy.postE, band), sPost));
    }
#pragma unroll
    for (int i = 0; i < C; ++i) { yL[i] = mul(add(oL[i], yL[i]), bk.gg); yR[i] = mul(add(oR[i], yR[i]), bk.gg); }   // :465-466
    io.store(yL, yR);
}

}  // namespace erb
}  // namespace za
