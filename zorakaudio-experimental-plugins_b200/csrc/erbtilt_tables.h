// erbtilt_tables.h -- HOST side of ERB Tilt: @init build_filterbank (ERB Tilt.jsfx:113-175) in double with the host libm, in
// the script's expression order (the reference's compiled @init calls the same libm, so the recursion coefficients lpA[]
// are bit-identical), and the scan tables of erbtilt_core.h derived from them in long double.  Shared by p_erbtilt.cu and
// the CPU emulation (tools/emu_scan.cpp).  Compile with -ffp-contract=off.
#pragma once
#include <cmath>
#include "erbtilt_core.h"

namespace za {
namespace erb {

struct Init {                            // what @init leaves behind for one sample rate
    double lpA[kNBm1], fcBand[kNB];
    BlockConsts bc;
    double energy_smooth, rough_fast, rough_slow, gain_tau, work_srate;
};

inline double erb_from_hz(double f) { return 21.4 * std::log(0.00437 * f + 1) * 0.4342944819032518; }           // :64-68
inline double hz_from_erb(double e) { return (std::exp((e / 21.4) * 2.302585092994046) - 1) / 0.00437; }         // :70-74
inline double a_weight_amp(double f)                                                                            // :76-90
{
    const double f2 = f * f;
    const double num = (12200.0 * 12200.0) * (f2 * f2);
    const double den1 = f2 + (20.6 * 20.6);
    const double den2 = std::sqrt((f2 + (107.7 * 107.7)) * (f2 + (737.9 * 737.9)));
    const double den3 = f2 + (12200.0 * 12200.0);
    const double ra = num / (den1 * den2 * den3);
    return ra * 1.2589254117941673;
}

inline Init build_filterbank(double srate)
{
    Init in{};
    const double pi = 3.141592653589793;
    double work_srate = srate;
    if (work_srate < 1000) work_srate = 48000;                 // last_good_srate is 48000 after @init (:208)
    in.work_srate = work_srate;
    const double nyq = work_srate * 0.5;
    double fmax = 20000;
    if (fmax > nyq * 0.98) fmax = nyq * 0.98;
    if (fmax < 40) fmax = 40;
    const double erb_min = erb_from_hz(20), erb_max = erb_from_hz(fmax);
    const double erb_range = erb_max - erb_min, erb_step = erb_range / kNB;
    for (int i = 0; i < kNB; ++i) {
        const double erb_c = erb_min + (i + 0.5) * erb_step;
        in.bc.erbBand[i] = erb_c;
        double fc = hz_from_erb(erb_c);
        if (fc > fmax) fc = fmax;
        in.fcBand[i] = fc;
        const double aw = a_weight_amp(fc);
        in.bc.wBand[i] = aw * aw;
    }
    for (int i = 0; i < kNBm1; ++i) {
        const double erb_b = erb_min + (i + 1) * erb_step;
        double f = hz_from_erb(erb_b);
        if (f > nyq * 0.98) f = nyq * 0.98;
        if (f < 10) f = 10;
        in.lpA[i] = std::exp(-2 * pi * f / work_srate);
    }
    in.energy_smooth = 1 - std::exp(-1 / (0.300 * work_srate));
    in.rough_fast = 1 - std::exp(-1 / (0.008 * work_srate));
    in.rough_slow = 1 - std::exp(-1 / (0.080 * work_srate));
    in.gain_tau = 0.120;
    in.bc.fmax = fmax; in.bc.erb_range = erb_range;
    return in;
}

template <class T, int C>
inline void build_tables(const Init& in, Tables<T, C>& tb)
{
    for (int i = 0; i < kNBm1; ++i) {
        const T a = (T)in.lpA[i];
        tb.lp_a[i] = a;
        tb.lp_1ma[i] = (T)(1 - in.lpA[i]);                      // the script's `(1 - a)` (fp64: that double; fp32: rounded from it)
        for (int j = 0; j <= 32; ++j) tb.lp_Q[i][j] = (T)std::pow((long double)a, (long double)(C * j));
        for (int k = 0; k < 5; ++k) tb.lp_P[i][k] = tb.lp_Q[i][1 << k];
    }
    auto follower = [](double c, T& tc, T* P, T* Q) {
        tc = (T)c;
        const long double mlt = 1.0L - (long double)tc;
        for (int j = 0; j <= 32; ++j) Q[j] = (T)std::pow(mlt, (long double)(C * j));
        for (int k = 0; k < 5; ++k) P[k] = Q[1 << k];
    };
    follower(in.rough_fast, tb.f_c, tb.f_P, tb.f_Q);
    follower(in.rough_slow, tb.s_c, tb.s_P, tb.s_Q);
    const long double ce = (long double)in.energy_smooth, me = 1.0L - ce;
    for (int i = 0; i < C; ++i) tb.e_v[i] = (T)(0.5L * ce * std::pow(me, (long double)(C - 1 - i)));
    for (int j = 0; j <= 32; ++j) tb.e_W[j] = (T)std::pow(me, (long double)(C * j));
    for (int i = 0; i < kNB; ++i) tb.guarded[i] = in.fcBand[i] > 2000 ? 1 : 0;
}

}  // namespace erb
}  // namespace za
