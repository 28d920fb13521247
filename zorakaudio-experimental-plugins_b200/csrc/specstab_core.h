// specstab_core.h -- DoSG Spectral Stabilizer (plugins/Spectral/SpectralStabilizer/src/Spectral Stabilizer.jsfx) as blocked
// parallel linear-recurrence scans over TIME: one warp renders one stream, lane l owns C consecutive frames of a 32 x C
// frame tile (lanes.h vocabulary: the same text is the CUDA kernel's tile body, k_specstab.cu, and the CPU emulation,
// tools/emu_scan.cpp).
//
// @sample (:274-338) per frame and band (12 bands): an RBJ band-pass biquad per channel (TDF-II, biquad_proc :98-111), an
// energy EMA, log, first and second differences of the log energy, a "motion" EMA, and a gain smoother whose target g_t
// is fixed by @block for the whole host block.  Per tile and band:
//   * the two biquads are 2-STATE scans: zero-state pass A over the lane's chunk, Kogge-Stone over the 32 chunk end states
//     with the 2x2 matrices (A^C)^(2^k), start state = prefix + (A^C)^lane x carry, pass B = biquad_proc's own operations
//     in its own order (fp64: never contracted); A = [[-a1, 1], [-a2, 0]], powers tabulated on the host in long double;
//   * the energy EMA is a 1-state full scan (the log of every value is needed);
//   * the differences d1, d2 are shifts across lanes (one shuffle each per tile) plus the carried previous values;
//   * the motion EMA (:318-320) is read only by @block (:241): END-STATE-ONLY weighted sum;
//   * the gain smoother (:322-326) switches between attack and release by `g_t < g_prev`; with g_t constant inside a host
//     block the smoother approaches it monotonically, so the branch taken at the block's first frame holds for the
//     whole block and g_n = g_t + (g_0 - g_t) c^n in closed form (c^n tabulated for n = 0 .. block_len); in floating point
//     the sequential recurrence may dither by an ulp or two around g_t once it has converged, far below the criterion.
// @slider (Gaussian weights, :189-191) and @block (difference-of-Gaussians across bands, :193-272) run once per host
// block on the device with lane = band, in double, in the script's operation order.
#pragma once
#include "lanes.h"

namespace za {
namespace sstab {
using namespace lanes;

constexpr int kBands = 12, kRadius = 5, kTaps = 11;

template <class T, int C>
struct Tables {
    T bq[kBands][5];                    // b0 b1 b2 a1 a2 (biquad_bp_coefs :73-96, host libm)
    T bq_M[kBands][5][4];               // (A^C)^(2^k), row major
    T bq_Q[kBands][33][4];              // (A^C)^j
    T env_a, env_1ma, env_P[5], env_Q[33];      // energy EMA (:199, :283-284): a, (1 - a), (a^C)^(2^k), (a^C)^j
    T mot_v[C], mot_W[33];              // motion EMA end-state weights: (1 - a) a^(C-1-i), (a^C)^j
};

template <class T>
struct Carry {                          // lane i = band i
    LV<T> z1L, z2L, z1R, z2R, env, elog, d1, d2, mot, gs;
};

template <class T>
struct Block {
    LV<T> gt, g0, atk;                  // target gain, gain at the block's first frame, 1 = attack branch (g_t < g_0)
    T action_amt;
};

struct Consts {                         // @init / @block sample-rate constants (host libm, double)
    double atk_a, rel_a, env_a, mot_a, dbfac, min_gain_lin, action_thr;
};
struct Prolog {                         // block-rate script variables kept for za_gpu_read_state (lane-distributed where per band)
    LV<double> gn, gw, res, gt;
    double depth, mb, low_min_gain, action_amt;
};

// @slider + @block for one host block: q = the block's quantised sliders (sigma, depth %, motion bias %).
template <class T>
ZA_LN void block_prologue(const Consts& k, const double* q, const Carry<T>& cy, Prolog& pv, Block<T>& bk)
{
    using D = LV<double>;
    auto fill = [&](double sigma) {                                        // fill_gauss_weights (:48-65), lanes 0..10 <-> j = -5..5
        sigma = sigma > 0.000001 ? sigma : 0.000001;
        const double den = mul(mul(2.0, sigma), sigma);
        D w = lane_generate<double>([=](int l) { const double j = (double)(l - kRadius); return l < kTaps ? exp_(div_(-mul(j, j), den)) : 0.0; });
        double sum = 0.0;
        for (int l = 0; l < kTaps; ++l) sum = add(sum, lane_value(w, l));
        return div_(w, sum > 0.00000000000000000001 ? sum : 0.00000000000000000001);
    };
    pv.gn = fill(q[0]);
    pv.gw = fill(mul(q[0], 2.50));
    pv.depth = mul(q[1], 0.01); pv.mb = mul(q[2], 0.01);                      // :199-200
    const D elog = lane_cast<double>(cy.elog), mot = lane_cast<double>(cy.mot);
    D acc_n = lane_zero<double>(), acc_w = acc_n;
    for (int kk = -kRadius; kk <= kRadius; ++kk) {                           // :208-221
        const D e = lane_gather(elog, [=](int l) { const int bj = l + kk; return bj < 0 ? 0 : (bj >= kBands ? kBands - 1 : bj); });
        acc_n = add(acc_n, mul(lane_value(pv.gn, kk + kRadius), e));
        acc_w = add(acc_w, mul(lane_value(pv.gw, kk + kRadius), e));
    }
    D r = sub(acc_n, acc_w);                                                 // :223
    r = div_(mul(r, 1.50), add(add(1.50, abs_(r)), 0.00000000000000000001));   // softclip_lim(r, res_clip)
    pv.res = r;
    const D mc = select(lt(mot, 0.0), 0.0, select(gt(mot, 1.0), 1.0, mot));
    const D mgate = sub(1.0, mul(pv.mb, mc));                                // :227
    const D rk = mul(r, div_(abs_(r), add(abs_(r), 0.50)));                  // :229
    D att = select(gt(rk, 0.0), mul(mul(mul(-pv.depth, mgate), rk), 6.0), 0.0);   // :231
    att = select(gt(att, -18.0), att, -18.0);                                // :232  max(att_db, -max_att_db)
    D gt_ = exp_(mul(att, k.dbfac));                                         // :234  db2lin
    gt_ = select(lt(gt_, k.min_gain_lin), k.min_gain_lin, select(gt(gt_, 1.0), 1.0, gt_));   // :235
    pv.gt = gt_;
    double lmg = 1.0;
    for (int bi = 0; bi < 3; ++bi) { const double g = lane_value(gt_, bi); lmg = lmg < g ? lmg : g; }   // :238
    pv.low_min_gain = lmg;
    pv.action_amt = lmg < k.action_thr ? 0.60 : 1.0;                          // :243-244
    bk.gt = lane_cast<T>(gt_);
    bk.g0 = cy.gs;
    bk.atk = select(lt(bk.gt, bk.g0), T(1), lane_zero<T>());
    bk.action_amt = (T)pv.action_amt;
}

// gain at the end of a block of nb frames: g_t + (g_0 - g_t) c^nb with the block's branch
template <class T>
ZA_LN void block_epilogue(Carry<T>& cy, const Block<T>& bk, T pwA_nb, T pwR_nb)
{
    const LV<T> a = add(lane_zero<T>(), pwA_nb), r = add(lane_zero<T>(), pwR_nb);
    cy.gs = add(bk.gt, mul(sub(bk.g0, bk.gt), select(gt(bk.atk, T(0.5)), a, r)));
}

// 2-state scan of one TDF-II biquad over the tile.  in: x, out: y; (z1, z2) carried.
template <class T, int C>
ZA_LN void biquad_tile(const LV<T> (&x)[C], LV<T> (&y)[C], const T* c5, const T (*M)[4], const T (*Q)[4], T& z1c, T& z2c, int m)
{
    using V = LV<T>;
    const T b0 = c5[0], b1 = c5[1], b2 = c5[2], a1 = c5[3], a2 = c5[4];
    V e1 = lane_zero<T>(), e2 = e1;
#pragma unroll
    for (int i = 0; i < C; ++i) {                                            // pass A: zero-state end state
        const V yy = fma_rn(b0, x[i], e1);
        e1 = fma_rn(b1, x[i], fma_rn(-a1, yy, e2));
        e2 = fma_rn(b2, x[i], mul(-a2, yy));
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const V t1 = shfl_up(e1, 1 << k), t2 = shfl_up(e2, 1 << k);
        const V n1 = fma_rn(M[k][0], t1, fma_rn(M[k][1], t2, e1)), n2 = fma_rn(M[k][2], t1, fma_rn(M[k][3], t2, e2));
        e1 = select(lane_ge(1 << k), n1, e1); e2 = select(lane_ge(1 << k), n2, e2);
    }
    const V q0 = lane_generate<T>([=](int l) { return Q[l][0]; }), q1 = lane_generate<T>([=](int l) { return Q[l][1]; });
    const V q2 = lane_generate<T>([=](int l) { return Q[l][2]; }), q3 = lane_generate<T>([=](int l) { return Q[l][3]; });
    V z1 = fma_rn(q0, z1c, fma_rn(q1, z2c, select(lane_ge(1), shfl_up(e1, 1), T(0))));
    V z2 = fma_rn(q2, z1c, fma_rn(q3, z2c, select(lane_ge(1), shfl_up(e2, 1), T(0))));
#pragma unroll
    for (int i = 0; i < C; ++i) {                                            // pass B: biquad_proc (:104-106)
        const V yy = add(mul(b0, x[i]), z1);
        z1 = add(sub(mul(b1, x[i]), mul(a1, yy)), z2);
        z2 = sub(mul(b2, x[i]), mul(a2, yy));
        y[i] = yy;
    }
    z1c = lane_value(z1, m - 1); z2c = lane_value(z2, m - 1);
}

// One tile.  fpos = frames of the host block rendered before this tile (index into the c^n tables); m = valid lanes.
template <class T, int C, bool FAST, class IO>
ZA_LN void tile(const Tables<T, C>& tb, const T* pwA, const T* pwR, Carry<T>& cy, const Block<T>& bk, int fpos, int m, IO& io)
{
    using V = LV<T>;
    V xL[C], xR[C], oL[C], oR[C], pA[C], pR[C];
    io.load(xL, xR);
#pragma unroll
    for (int i = 0; i < C; ++i) {
        oL[i] = xL[i]; oR[i] = xR[i];                                        // :277-278
        pA[i] = lane_generate<T>([=](int l) { return pwA[fpos + l * C + i + 1]; });
        pR[i] = lane_generate<T>([=](int l) { return pwR[fpos + l * C + i + 1]; });
    }
    const V motW = lane_load_rev<T>(tb.mot_W, m);
    const V envQ = lane_load<T>(tb.env_Q);
#pragma unroll 1
    for (int band = 0; band < kBands; ++band) {
        V yL[C], yR[C], en[C];
        {
            T z1 = lane_value(cy.z1L, band), z2 = lane_value(cy.z2L, band);
            biquad_tile<T, C>(xL, yL, tb.bq[band], tb.bq_M[band], tb.bq_Q[band], z1, z2, m);      // :283
            lane_set(cy.z1L, band, z1); lane_set(cy.z2L, band, z2);
            z1 = lane_value(cy.z1R, band); z2 = lane_value(cy.z2R, band);
            biquad_tile<T, C>(xR, yR, tb.bq[band], tb.bq_M[band], tb.bq_Q[band], z1, z2, m);      // :284
            lane_set(cy.z1R, band, z1); lane_set(cy.z2R, band, z2);
        }
        // energy EMA (:286-289): env = env_a * env + (1 - env_a) * e, full scan
#pragma unroll
        for (int i = 0; i < C; ++i) en[i] = mul(tb.env_1ma, mul(T(0.5), add(mul(yL[i], yL[i]), mul(yR[i], yR[i]))));
        {
            V e = en[0];
#pragma unroll
            for (int i = 1; i < C; ++i) e = fma_rn(e, tb.env_a, en[i]);
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const V t = shfl_up(e, 1 << k);
                e = select(lane_ge(1 << k), fma_rn(t, tb.env_P[k], e), e);
            }
            V y = fma_rn(envQ, lane_value(cy.env, band), select(lane_ge(1), shfl_up(e, 1), T(0)));
#pragma unroll
            for (int i = 0; i < C; ++i) { y = add(mul(tb.env_a, y), en[i]); en[i] = y; }
            lane_set(cy.env, band, lane_value(y, m - 1));
        }
        // log energy and its differences (:291-301); en[] becomes elog
#pragma unroll
        for (int i = 0; i < C; ++i) {
            const V a = add(en[i], T(0.000000000001));
            en[i] = FAST ? mul(lg2_approx(a), T(0.6931471805599453)) : log_(a);
        }
        const T pe = lane_value(cy.elog, band), pd = lane_value(cy.d1, band);
        V prev_e = select(lane_ge(1), shfl_up(en[C - 1], 1), pe);
        V d1[C];
#pragma unroll
        for (int i = 0; i < C; ++i) { d1[i] = sub(en[i], prev_e); prev_e = en[i]; }
        V prev_d = select(lane_ge(1), shfl_up(d1[C - 1], 1), pd);
        V accM = lane_zero<T>();
        V d2last = lane_zero<T>();
        const T gt_ = lane_value(bk.gt, band), g0 = lane_value(bk.g0, band);
        const bool atk = lane_value(bk.atk, band) > T(0.5);
        const T dg = sub(g0, gt_);
#pragma unroll
        for (int i = 0; i < C; ++i) {
            const V d2 = sub(d1[i], prev_d); prev_d = d1[i]; d2last = d2;
            const V onset = select(gt(d1[i], T(0.12)), T(1), lane_zero<T>());                                     // :305
            V mi = add(add(mul(abs_(d1[i]), T(6.0)), mul(abs_(d2), T(3.0))), mul(onset, T(0.75)));                // :307
            mi = select(lt(mi, T(0)), T(0), select(gt(mi, T(1)), T(1), mi));                                      // :308
            accM = fma_rn(mi, tb.mot_v[i], accM);                                                                // :310-312, end state only
            const V gn = add(gt_, mul(dg, atk ? pA[i] : pR[i]));                                                  // :314-326 in closed form
            const V gm1 = sub(add(T(1), mul(bk.action_amt, sub(gn, T(1)))), T(1));                                // :330, :332 (g_apply - 1)
            oL[i] = add(oL[i], mul(gm1, yL[i]));                                                                  // :332-333
            oR[i] = add(oR[i], mul(gm1, yR[i]));
        }
        lane_set(cy.elog, band, lane_value(en[C - 1], m - 1));
        lane_set(cy.d1, band, lane_value(d1[C - 1], m - 1));
        lane_set(cy.d2, band, lane_value(d2last, m - 1));
        {
            V s = mul(accM, motW);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) s = add(s, shfl_down(s, d));
            lane_set(cy.mot, band, fma_rn(tb.mot_W[m], lane_value(cy.mot, band), lane_value(s, 0)));
        }
    }
    io.store(oL, oR);
}

}  // namespace sstab
}  // namespace za
