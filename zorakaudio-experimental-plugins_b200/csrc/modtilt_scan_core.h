// modtilt_scan_core.h -- ModTilt (plugins/Dynamics/ModTilt/src/ModTilt.dsp) as a blocked parallel linear-recurrence scan
// over TIME: one warp renders one stream, lane l owns kC consecutive frames of a 32 x kC frame tile.
//
// Every recurrence of the leaf is onepole(a) = *(a) : (+ ~ *(1 - a)) (:29) with a coefficient that is constant inside a
// host block, and the maps between them (sqrt, the clamped ratio, 20 log10, 10^(x/20)) are memoryless, so the leaf is five
// scans in a row.  Per stage and tile:
//   pass A   each lane runs the recurrence over its chunk from state 0 (one FMA per frame; only the end value is kept);
//   scan     Kogge-Stone over the 32 chunk end values with the constants P[k] = (na^kC)^(2^k): lane l then holds the
//            zero-carry state at the END of chunk l; the state at the START of chunk l is that of lane l - 1 plus
//            Q[l] x carry, Q[l] = (na^kC)^l, carry = the exact state at the end of the previous tile;
//   pass B   each lane re-runs the .dsp's own recurrence (separate multiply and add, its rounding) over its chunk from
//            that start state and keeps all values.
// Only the 31 chunk start states per tile are re-associated; lane 0 continues the sequential recurrence bit for bit and
// a start-state error decays with the smoother itself.  The powers P / Q are tabulated on the host in double from the
// float coefficient the .dsp's expressions give (per pivot-slider grid value for the one stage that depends on a widget),
// so no power is ever formed by repeated float multiplication.  Written against lanes.h: the same text is the kernel's
// tile body (k_modtilt_scan.cu) and the CPU emulation (tools/emu_scan.cpp, tests/test_faust_port.py).
#pragma once
#include "lanes.h"
#include "za_internal.h"

namespace za {
namespace mtscan {
using namespace lanes;

constexpr int kC = 8;                 // frames per lane; tile = 256 frames
constexpr int kTile = 32 * kC;

struct StageK { float a, na, P[5]; };                       // uniform constants of a stage with a fixed coefficient
struct Consts { StageK env, base, ratio, trim; };           // kernel parameter (constant bank)
struct Carry { float env2, base, lp, rs, mean; };           // state at the end of the previous tile (ModTilt's five onepoles)
}  // namespace mtscan
struct MTScanArgs {
    MTArgs a;
    mtscan::Consts K;
    const float* qfix;      // [4][32]: Q rows of the env, base, ratio, trim stages
    const float* qpiv;      // [pivot grid values][32]: Q rows of the pivot stage per slider grid index
};
namespace mtscan {

// branch-free round-to-nearest division / square root of modtilt_core.h, in the lane vocabulary (same instruction
// sequences: reciprocal refined once, quotient corrected twice with exact FMA residuals; operands in safe ranges)
ZA_LN LF div_rn_safe(const LF& a, const LF& b)
{
    LF r = rcp_approx(b);
    r = fma_rn(r, fma_rn(neg(b), r, 1.0f), r);
    LF q = mul(a, r);
    q = fma_rn(r, fma_rn(neg(b), q, a), q);
    q = fma_rn(r, fma_rn(neg(b), q, a), q);
    return q;
}
ZA_LN LF div20_rn(const LF& a)
{
    const LF q = mul(a, 0.05f);
    return fma_rn(0.05f, fma_rn(-20.0f, q, a), q);
}
ZA_LN LF sqrt_rn_safe(const LF& x)                           // x >= 0
{
    const LF y = rsqrt_approx(x);
    LF s = mul(x, y);
    const LF h = mul(0.5f, y);
    s = fma_rn(fma_rn(neg(s), s, x), h, s);
    s = fma_rn(fma_rn(neg(s), s, x), h, s);
    return select(gt(x, 0.0f), s, 0.0f);
}

// Throughput forms of the three memoryless maps (FAST = true, the kernel's default): MUFU approximations with one
// multiply, relative error ~2^-22 each, i.e. a gain error near -130 dB relative to the signal -- far inside the fp32
// criterion (-100 dBFS residual), at a third of the instructions of the correctly rounded sequences above (ncu on the
// correctly rounded build: issue slots 80 % busy, 137 instructions per frame).  The correctly rounded forms stay
// available (FAST = false; the stream-per-thread kernel always uses them) as the cross-check.
ZA_LN LF div_fast(const LF& a, const LF& b) { return mul(a, rcp_approx(b)); }
ZA_LN LF sqrt_fast(const LF& x) { return select(gt(x, 0.0f), mul(x, rsqrt_approx(x)), 0.0f); }
ZA_LN LF db_fast(const LF& g) { return mul(lg2_approx(g), 6.0205999132796239f); }              // 20 log10(g) = 20 log10(2) lg2(g)
ZA_LN LF undb_fast(const LF& db) { return ex2_approx(mul(db, -0.16609640474436813f)); }        // 10^(-db / 20) = 2^(-db log2(10) / 20)

// One onepole stage over the tile.  In: v[i] = x_i * a (the .dsp's `*(a)`), out: v[i] = y_i.  `carry` in = state before
// the tile, out = state after it.  P(k) = (na^kC)^(2^k), Q = (na^kC)^lane.
template <class PF>
ZA_LN void onepole_tile(LF (&v)[kC], float na, PF P, const LF& Q, float& carry)
{
    LF e = v[0];
#pragma unroll
    for (int i = 1; i < kC; ++i) e = fma_rn(e, na, v[i]);                     // pass A: zero-state end value
#pragma unroll
    for (int k = 0; k < 5; ++k) {                                            // scan
        const LF t = shfl_up(e, 1 << k);
        e = select(lane_ge(1 << k), fma_rn(t, P(k), e), e);
    }
    LF y = fma_rn(Q, carry, select(lane_ge(1), shfl_up(e, 1), 0.0f));      // start state of this lane's chunk
#pragma unroll
    for (int i = 0; i < kC; ++i) { y = add(v[i], mul(y, na)); v[i] = y; }      // pass B: the .dsp's recurrence (:29)
    carry = lane_value(y, 31);
}

// One tile of `stereo` (ModTilt.dsp:51-102).  IO: load(xl, xr) fills the lane's kC frames of both channels (called twice:
// the inputs are not kept in registers across the five stages), store(xl, xr) takes the outputs.
template <bool FAST, class IO>
ZA_LN void tile(Carry& cy, const Consts& K, const LF& q_env, const LF& q_base, const LF& q_ratio, const LF& q_trim, const LF& q_piv,
                const MTCoef& c, IO& io)
{
    LF a[kC], e[kC], b[kC], m[kC];
    {
        LF xl[kC], xr[kC];
        io.load(xl, xr);
#pragma unroll
        for (int i = 0; i < kC; ++i) { const LF x = mul(0.5f, add(xl[i], xr[i])); a[i] = mul(mul(x, x), K.env.a); }      // :54,57
    }
    onepole_tile(a, K.env.na, [&](int k) { return K.env.P[k]; }, q_env, cy.env2);                                        // :58  a = env2
#pragma unroll
    for (int i = 0; i < kC; ++i) { e[i] = FAST ? sqrt_fast(max_(a[i], 0.0f)) : sqrt_rn_safe(max_(a[i], 0.0f)); a[i] = mul(e[i], K.base.a); }                   // :59
    onepole_tile(a, K.base.na, [&](int k) { return K.base.P[k]; }, q_base, cy.base);                                     // :62  a = base
#pragma unroll
    for (int i = 0; i < kC; ++i) { b[i] = a[i]; m[i] = sub(e[i], b[i]); a[i] = mul(m[i], c.a_piv); }                      // :65
    {
        float pp[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) pp[k] = lane_value(q_piv, 1 << k);                                                   // (na^kC)^(2^k) = Q[2^k]
        onepole_tile(a, c.na_piv, [&](int k) { return pp[k]; }, q_piv, cy.lp);                                            // :68  a = lp_piv
    }
#pragma unroll
    for (int i = 0; i < kC; ++i) {
        const LF lp = a[i];
        const LF m2_tilt = add(mul(lp, c.g_lo), mul(sub(m[i], lp), c.g_hi));                                              // :70,73
        const LF m2 = add(mul(m[i], 0.25f), mul(m2_tilt, 0.75f));                                                         // :74
        const LF env_tp = max_(add(b[i], m2), mul(0.05f, e[i]));                                                          // :77,80
        const LF r0 = FAST ? div_fast(add(env_tp, 1e-9f), add(e[i], 1e-9f)) : div_rn_safe(add(env_tp, 1e-9f), add(e[i], 1e-9f));                                                  // :82
        a[i] = mul(sub(min_(max_(r0, 0.67f), 1.5f), 1.0f), 0.05f);                                                        // :85,88
    }
    onepole_tile(a, K.ratio.na, [&](int k) { return K.ratio.P[k]; }, q_ratio, cy.rs);                                    // :88  a = r_s - 1
#pragma unroll
    for (int i = 0; i < kC; ++i) {
        m[i] = add(1.0f, a[i]);                                                                                           // g
        a[i] = mul(FAST ? db_fast(max_(m[i], 1e-12f)) : mul(20.0f, log10_(max_(m[i], 1e-12f))), K.trim.a);                                                     // :92
    }
    onepole_tile(a, K.trim.na, [&](int k) { return K.trim.P[k]; }, q_trim, cy.mean);                                     // :93  a = mean_rdb
    {
        LF xl[kC], xr[kC];
        io.load(xl, xr);
#pragma unroll
        for (int i = 0; i < kC; ++i) {
            const LF trim = FAST ? undb_fast(a[i]) : exp10_(div20_rn(neg(a[i])));                                                                  // :94
            xl[i] = mul(add(mul(xl[i], c.nmix), mul(mul(xl[i], m[i]), c.mix)), trim);                                     // :97,100
            xr[i] = mul(add(mul(xr[i], c.nmix), mul(mul(xr[i], m[i]), c.mix)), trim);                                     // :98,101
        }
        io.store(xl, xr);
    }
}

// Host side of the tables: (na^kC)^j for j = 0 .. 31, in double from the float coefficient, rounded once.
inline void power_row(float na, float* q32)
{
    double d = 1.0;
    for (int i = 0; i < kC; ++i) d *= (double)na;
    double p = 1.0;
    for (int j = 0; j < 32; ++j) { q32[j] = (float)p; p *= d; }
}
inline StageK make_stage(float a, float na)
{
    StageK s{};
    s.a = a; s.na = na;
    float q[32];
    power_row(na, q);
    for (int k = 0; k < 5; ++k) s.P[k] = q[1 << k];
    return s;
}

}  // namespace mtscan
}  // namespace za
