// specstab_tables.h -- HOST side of the Spectral Stabilizer: @init (Spectral Stabilizer.jsfx:21-187) in double with the host
// libm in the script's expression order (the reference's compiled @init calls the same libm, so the biquad coefficients
// are bit-identical), and the scan tables of specstab_core.h derived from them in long double.  Shared by p_specstab.cu
// and the CPU emulation (tools/emu_scan.cpp).  Compile with -ffp-contract=off.
#pragma once
#include <cmath>
#include <vector>
#include "specstab_core.h"

namespace za {
namespace sstab {

struct Init {
    double bq[kBands][5];
    Consts k;
};

inline double jmin(double a, double b) { return a < b ? a : b; }        // olt / ogt selects (dsp_jsfx_aot.py:5223-5232)
inline double jmax(double a, double b) { return a > b ? a : b; }

inline Init build_init(double srate)
{
    Init in{};
    const double ln10 = std::log(10.0), dbfac = ln10 / 20, two_pi = 6.283185307179586;
    const double minHz = 100, maxHz = 12000;
    const double maxHz_eff = jmin(maxHz, srate * 0.45);
    for (int bi = 0; bi < kBands; ++bi) {
        const double ear_min = 21.4 * std::log10(4.37 * minHz * 0.001 + 1.0), ear_max = 21.4 * std::log10(4.37 * maxHz_eff * 0.001 + 1.0);   // :33-39
        const double t = bi / jmax(kBands - 1, 1);
        const double cf = (std::pow(10.0, (ear_min + t * (ear_max - ear_min)) / 21.4) - 1.0) / 4.37 * 1000.0;
        const double bw = 24.7 * (4.37 * cf * 0.001 + 1.0);                  // :41
        double q = cf / jmax(bw, 1.0);
        double fc = cf < 10 ? 10 : (cf > srate * 0.45 ? srate * 0.45 : cf);  // biquad_bp_coefs :73-96
        q = jmax(q, 0.10);
        const double w0 = two_pi * fc / srate, alpha = std::sin(w0) / (2.0 * q);
        double b0 = alpha, b1 = 0.0, b2 = -alpha;
        const double a0 = 1.0 + alpha;
        double a1 = -2.0 * std::cos(w0), a2 = 1.0 - alpha;
        b0 /= a0; b1 /= a0; b2 /= a0; a1 /= a0; a2 /= a0;
        in.bq[bi][0] = b0; in.bq[bi][1] = b1; in.bq[bi][2] = b2; in.bq[bi][3] = a1; in.bq[bi][4] = a2;
    }
    in.k.atk_a = std::exp(-1000.0 / (jmax(6, 0.01) * srate));              // :194-197
    in.k.rel_a = std::exp(-1000.0 / (jmax(160, 0.01) * srate));
    in.k.env_a = std::exp(-1000.0 / (jmax(25, 0.01) * srate));
    in.k.mot_a = std::exp(-1000.0 / (jmax(40, 0.01) * srate));
    in.k.dbfac = dbfac;
    in.k.min_gain_lin = std::exp(-18.0 * dbfac);                            // :132
    in.k.action_thr = std::exp(-4.0 * dbfac);                               // :244
    return in;
}

struct M2 { long double a, b, c, d; };
inline M2 mmul(const M2& x, const M2& y) { return {x.a * y.a + x.b * y.c, x.a * y.b + x.b * y.d, x.c * y.a + x.d * y.c, x.c * y.b + x.d * y.d}; }

template <class T, int C>
inline void build_tables(const Init& in, Tables<T, C>& tb)
{
    for (int bi = 0; bi < kBands; ++bi) {
        for (int j = 0; j < 5; ++j) tb.bq[bi][j] = (T)in.bq[bi][j];
        const M2 A{-(long double)tb.bq[bi][3], 1.0L, -(long double)tb.bq[bi][4], 0.0L};       // state map of the TDF-II section
        M2 AC{1, 0, 0, 1};
        for (int i = 0; i < C; ++i) AC = mmul(AC, A);
        M2 P{1, 0, 0, 1};
        for (int j = 0; j <= 32; ++j) {
            tb.bq_Q[bi][j][0] = (T)P.a; tb.bq_Q[bi][j][1] = (T)P.b; tb.bq_Q[bi][j][2] = (T)P.c; tb.bq_Q[bi][j][3] = (T)P.d;
            P = mmul(P, AC);
        }
        for (int k = 0; k < 5; ++k) for (int e = 0; e < 4; ++e) tb.bq_M[bi][k][e] = tb.bq_Q[bi][1 << k][e];
    }
    tb.env_a = (T)in.k.env_a; tb.env_1ma = (T)(1.0 - in.k.env_a);
    for (int j = 0; j <= 32; ++j) tb.env_Q[j] = (T)std::pow((long double)tb.env_a, (long double)(C * j));
    for (int k = 0; k < 5; ++k) tb.env_P[k] = tb.env_Q[1 << k];
    const long double ma = (long double)in.k.mot_a;
    for (int i = 0; i < C; ++i) tb.mot_v[i] = (T)((1.0L - ma) * std::pow(ma, (long double)(C - 1 - i)));
    for (int j = 0; j <= 32; ++j) tb.mot_W[j] = (T)std::pow(ma, (long double)(C * j));
}

// c^n for n = 0 .. block_len (+ slack for the unused lanes of a partial tile)
template <class T>
inline std::vector<T> power_table(double c, int block_len)
{
    std::vector<T> t((size_t)block_len + 32 * 8 + 2);
    for (size_t n = 0; n < t.size(); ++n) t[n] = (T)std::pow((long double)c, (long double)n);
    return t;
}

}  // namespace sstab
}  // namespace za
