// faust_leaves.cuh -- the per-frame arithmetic of the sequential Faust leaves (RED, ClickBeGoneSG, VAR) for the generic
// stream-per-thread kernel of k_faust.cu.  Each leaf is a struct with
//   kIn / kOut / kSliders, Slow (slider-rate values of one compute() call, evaluated on the HOST with the host libm like
//   Faust's fSlow expressions), Konst (sample-rate constants, host), State (the recursive signals; all zero after init),
//   slow(zone values, fs, Slow&) / konst(fs, Konst&) host functions and the device function frame().
// Float32 like the reference's Faust build (no -double, cmake/plugin/CMakeLists.txt:144); this translation unit is
// compiled with -fmad=false so that the .dsp's separate multiplies and adds are not contracted.  These leaves are
// sequential by nature (switching followers, max-hold envelopes), hence one stream per thread; the time-parallel Faust
// leaves are k_modtilt_scan.cu and k_gts.cu.  Library definitions assumed (not in the reference tree):
// si.smooth, si.smoo = smooth(1 - 44.1/SR), ba.tau2pole, si.onePoleSwitching, an.amp_follower_ar, fi.tf22t, no.noise.
#pragma once
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

namespace za {
namespace faust {

#define ZA_FL __device__ __forceinline__
ZA_FL float clampf_(float x, float lo, float hi) { return fmaxf(lo, fminf(hi, x)); }
ZA_FL float db2lin_(float db) { return powf(10.0f, db / 20.0f); }
ZA_FL float smooth_(float& st, float pole, float x) { st = (x * (1.0f - pole)) + (st * pole); return st; }
struct Bq { float b0, b1, b2, a1, a2; };
struct BqZ { float s1, s2; };
ZA_FL float tf22t_(const Bq& c, BqZ& z, float x)
{
    const float y = (c.b0 * x) + z.s1;
    z.s1 = ((c.b1 * x) - (c.a1 * y)) + z.s2;
    z.s2 = (c.b2 * x) - (c.a2 * y);
    return y;
}
inline float host_db2lin(float db) { return std::pow(10.0f, db / 20.0f); }
inline float host_clamp(float x, float lo, float hi) { return std::fmax(lo, std::fmin(hi, x)); }

// ---- RED: plugins/Dynamics/RED/src/Reverb Expanding Downwards (RED).dsp (6 in / 6 out) -------------------------------------
struct Red {
    static constexpr int kIn = 6, kOut = 6, kSliders = 3;
    static const char* key() { return "RED"; }
    static const char* slider_text()
    {
        return "slider1:amount_db=12<0,24,0.1>Amount (max duck dB)\n"          // RED.dsp:8-10
               "slider2:sens_pct=50<0,100,1>Sensitivity (%)\n"
               "slider3:rel_ms=350<50,1200,1>Release (ms)\n";
    }
    struct Slow { float thr_db, ratio, knee_db, maxduck, pole_grace, c_rel, pad0, pad1; };
    struct Konst { float pole_rms, pole_tgt, pole_hold, pole_dry, c_atk, c_rel_in, floor_lin, dry_on_lin, ref_off_lin; };
    struct State { float wet_env2, ref_env2, offA_s, tgt_db, dryA_s, tgt_hold_s, gr_norm, gr_fast; };
    static float ms2pole(float ms, float fs) { return std::exp(-1.0f / (fs * (ms / 1000.0f))); }          // :24
    static void konst(float fs, Konst& k)
    {
        k.pole_rms = ms2pole(35.0f, fs); k.pole_tgt = ms2pole(25.0f, fs); k.pole_hold = ms2pole(80.0f, fs); k.pole_dry = ms2pole(10.0f, fs);
        k.c_atk = std::exp(-1.0f / ((12.0f / 1000.0f) * fs)); k.c_rel_in = std::exp(-1.0f / ((70.0f / 1000.0f) * fs));
        k.floor_lin = host_db2lin(-80.0f); k.dry_on_lin = host_db2lin(-50.0f); k.ref_off_lin = host_db2lin(-60.0f);
    }
    static void slow(const float* z, float fs, Slow& c)
    {
        const float sens = z[1] / 100.0f, rel_ms = z[2];
        c.maxduck = z[0];
        c.thr_db = 18.0f - (sens * 21.0f); c.ratio = 1.2f + (sens * 3.0f); c.knee_db = 10.0f - (sens * 6.0f);   // :49-51
        c.pole_grace = ms2pole(host_clamp(rel_ms * 0.25f, 60.0f, 200.0f), fs);                                  // :54,60
        c.c_rel = std::exp(-1.0f / ((rel_ms / 1000.0f) * fs));
        c.pad0 = c.pad1 = 0.0f;
    }
    ZA_FL static float smoothstep01(float x) { const float x1 = clampf_(x, 0.0f, 1.0f); return (x1 * x1) * (3.0f - (2.0f * x1)); }
    // frame() in five pieces, so that the tiled kernel (p_faust.cu, red_tiled_kernel) can run the memoryless ones lane = frame
    // and only the recurrences lane = stream: the same operations on the same values in the same order either way.
    ZA_FL static void powers(float wetL, float wetR, float refL, float refR, float& wet_p, float& ref_p)       // :68-69
    {
        wet_p = 0.5f * ((wetL * wetL) + (wetR * wetR)); ref_p = 0.5f * ((refL * refL) + (refR * refR));
    }
    ZA_FL static void envelopes(State& m, const Konst& k, float wet_p, float ref_p)                            // :72-73 (carried)
    {
        smooth_(m.wet_env2, k.pole_rms, wet_p); smooth_(m.ref_env2, k.pole_rms, ref_p);
    }
    struct Det { float dryA, offA, tgt0; };
    ZA_FL static Det detect(const Slow& c, const Konst& k, float wet_env2, float ref_env2)                     // memoryless
    {
        const float eps = 1e-12f;
        const float Ey = fmaxf(sqrtf(fmaxf(wet_env2, 0.0f)), k.floor_lin), Ex = fmaxf(sqrtf(fmaxf(ref_env2, 0.0f)), k.floor_lin);
        Det d;
        d.dryA = Ex > k.dry_on_lin ? 1.0f : 0.0f; d.offA = Ex <= k.ref_off_lin ? 1.0f : 0.0f;                     // :78-79
        const float rdB = 20.0f * log10f(fmaxf((Ey + eps) / (Ex + eps), 1e-30f));                                  // :87
        const float over = rdB - c.thr_db;
        const float over_eff = over <= 0.0f ? 0.0f : over * smoothstep01(clampf_(over / fmaxf(c.knee_db, 0.001f), 0.0f, 1.0f));   // :91-94
        d.tgt0 = over_eff > 0.0f ? fminf(c.maxduck, over_eff * c.ratio) : 0.0f;                                    // :97
        return d;
    }
    ZA_FL static float reduction_db(State& m, const Slow& c, const Konst& k, const Det& d)                     // carried: six followers
    {
        smooth_(m.offA_s, c.pole_grace, d.offA);                                                                   // :83
        const float tail_w = (1.0f - d.offA) + (d.offA * smoothstep01(m.offA_s));                                  // :84
        smooth_(m.tgt_db, k.pole_tgt, d.tgt0 * tail_w);                                                            // :98,101
        smooth_(m.dryA_s, k.pole_dry, d.dryA);                                                                     // :107
        smooth_(m.tgt_hold_s, k.pole_hold, m.tgt_db);
        const float tgt_pin = ((1.0f - d.dryA) * fmaxf(m.tgt_db, m.tgt_hold_s)) + (d.dryA * m.tgt_db);             // :112-113
        const float x = fabsf(tgt_pin);
        { const float cc = x > m.gr_norm ? k.c_atk : c.c_rel; m.gr_norm = ((1.0f - cc) * x) + (cc * m.gr_norm); }     // :120
        { const float cc = x > m.gr_fast ? k.c_atk : k.c_rel_in; m.gr_fast = ((1.0f - cc) * x) + (cc * m.gr_fast); }  // :121
        return ((1.0f - m.dryA_s) * m.gr_norm) + (m.dryA_s * m.gr_fast);                                           // :124
    }
    ZA_FL static float gain(float gr_db) { return db2lin_(-gr_db); }                                             // :127
    ZA_FL static void frame(State& m, const Slow& c, const Konst& k, const float* in, float* out)
    {
        const float wetL = in[0], wetR = in[1], refL = in[4], refR = in[5];
        float wet_p, ref_p;
        powers(wetL, wetR, refL, refR, wet_p, ref_p);
        envelopes(m, k, wet_p, ref_p);
        const Det d = detect(c, k, m.wet_env2, m.ref_env2);
        const float g = gain(reduction_db(m, c, k, d));
        out[0] = wetL * g; out[1] = wetR * g; out[2] = in[2]; out[3] = in[3]; out[4] = refL; out[5] = refR;       // :43
    }
};

// ---- ClickBeGoneSG: plugins/Restoration/ClickBeGoneSG/src/Click-Be-Gone (SG).dsp -----------------------------------------------
struct Cbg {
    static constexpr int kIn = 2, kOut = 2, kSliders = 5;
    static const char* key() { return "ClickBeGoneSG"; }
    static const char* slider_text()
    {
        return "slider1:amount=50<0,100,1>Amount [%]\n"                        // Click-Be-Gone (SG).dsp:11-15
               "slider2:sensitivity=50<0,100,1>Sensitivity [%]\n"
               "slider3:hpf_hz=1500<300,6000,10>HPF [Hz]\n"
               "slider4:mode=1<0,2,1{Fast,Medium,Slow}>Mode\n"
               "slider5:monitor=0<0,1,1{Output,Delta}>Monitor\n";
    }
    struct Slow { float a, env_rel, base_a, ratio_thr, err_thr, mix_max, relHold, range; int mode, monitor, pad0, pad1; };
    struct Konst { int unused; };
    struct State { float hL[32], hR[32]; int pos; float hpL, hpR, env, base, holdEnv; };
    static void konst(float, Konst& k) { k.unused = 0; }
    static void slow(const float* z, float fs, Slow& c)
    {
        static const float ratio_mul[3] = {1.12f, 1.00f, 0.92f}, err_mul[3] = {1.18f, 1.00f, 0.90f}, mix_mul[3] = {0.85f, 1.00f, 1.08f},
                           hold_mul[3] = {0.75f, 1.00f, 1.35f}, base_mul[3] = {0.85f, 1.00f, 1.10f}, env_mul[3] = {0.85f, 1.00f, 1.10f};
        const float amount = z[0] / 100.0f, sens = z[1] / 100.0f, eps = 1e-12f;
        const int md = (int)z[3] < 0 ? 0 : ((int)z[3] > 2 ? 2 : (int)z[3]);
        c.mode = md; c.monitor = z[4] != 0.0f; c.pad0 = c.pad1 = 0;
        const float env_rel_ms = (30.0f - (20.0f * sens)) * env_mul[md], base_ms = (300.0f - (180.0f * sens)) * base_mul[md];   // :28-35
        c.ratio_thr = (6.0f - (4.0f * sens)) * ratio_mul[md]; c.err_thr = (0.25f - (0.17f * sens)) * err_mul[md];              // :20-21,37-38
        c.mix_max = std::fmin((0.60f + (0.32f * amount)) * mix_mul[md], 0.96f);                                                // :40-41
        const float holdN = std::fmax((8.0f + (amount * 32.0f)) * hold_mul[md], 4.0f);                                         // :43-44
        c.env_rel = std::exp(-1000.0f / (fs * env_rel_ms)); c.base_a = 1.0f - std::exp(-1000.0f / (fs * base_ms));             // :46-47
        c.a = std::exp((-2.0f * 3.14159265358979323846f * z[2]) / fs);                                                        // :49
        c.relHold = std::exp(std::log(1e-3f) / (holdN + eps));                                                                // :102
        c.range = c.err_thr * 3.0f;                                                                                           // :108
    }
    // Savitzky-Golay predictors on delayed taps (:53-59), summed left to right; which = 0..3 -> SG-11 / 15 / 21 / 31.
    // `x(d)` = the input d frames ago (d = 0: the frame just pushed).  With `which` a compile-time constant after inlining
    // the tap loop unrolls and the coefficients become immediates (the tiled kernel of p_faust.cu relies on that).
    template <class X> ZA_FL static float sg(const X& x, int which)
    {
        const float c11[11] = {-36, 9, 44, 69, 84, 89, 84, 69, 44, 9, -36};
        const float c15[15] = {-78, -13, 42, 87, 122, 147, 162, 167, 162, 147, 122, 87, 42, -13, -78};
        const float c21[21] = {-171, -76, 9, 84, 149, 204, 249, 284, 309, 324, 329, 324, 309, 284, 249, 204, 149, 84, 9, -76, -171};
        const float c31[31] = {-406, -261, -126, -1, 114, 219, 314, 399, 474, 539, 594, 639, 674, 699, 714, 719, 714, 699, 674, 639, 594, 539,
                               474, 399, 314, 219, 114, -1, -126, -261, -406};
        if (which == 0) {
            float acc = c11[0] * x(20);
#pragma unroll
            for (int k = 1; k < 11; ++k) acc = acc + (c11[k] * x(20 - k));
            return acc / 429.0f;
        }
        if (which == 1) {
            float acc = c15[0] * x(22);
#pragma unroll
            for (int k = 1; k < 15; ++k) acc = acc + (c15[k] * x(22 - k));
            return acc / 1105.0f;
        }
        if (which == 2) {
            float acc = c21[0] * x(25);
#pragma unroll
            for (int k = 1; k < 21; ++k) acc = acc + (c21[k] * x(25 - k));
            return acc / 3059.0f;
        }
        float acc = c31[0] * x(30);
#pragma unroll
        for (int k = 1; k < 31; ++k) acc = acc + (c31[k] * x(30 - k));
        return acc / 9889.0f;
    }
    // the feed-forward half of a frame (:83-98): centre taps, the two predictors per channel, the better one and its error
    struct Pred { float xC_L, xC_R, pred_L, pred_R, e_norm; };
    template <class XL, class XR> ZA_FL static Pred predict(const XL& xl, const XR& xr, int mode)
    {
        Pred p;
        p.xC_L = xl(15); p.xC_R = xr(15);                                                              // :83-84
        const float small_L = sg(xl, mode), small_R = sg(xr, mode);                                   // :86-87
        const float large_L = sg(xl, mode + 1), large_R = sg(xr, mode + 1);                           // :88-89
        const float eA = fmaxf(fabsf(p.xC_L - small_L), fabsf(p.xC_R - small_R)) / (fmaxf(fabsf(small_L), fabsf(small_R)) + 1e-6f);
        const float eB = fmaxf(fabsf(p.xC_L - large_L), fabsf(p.xC_R - large_R)) / (fmaxf(fabsf(large_L), fabsf(large_R)) + 1e-6f);
        const bool useA = eA <= eB;
        p.pred_L = useA ? small_L : large_L; p.pred_R = useA ? small_R : large_R; p.e_norm = useA ? eA : eB;
        return p;
    }
    // the recurrent half (:67-79, :100-115): detector high-pass, max-hold envelope, baseline, trigger hold, mix
    struct Rec { float hpL, hpR, env, base, holdEnv; };
    ZA_FL static void tail(Rec& m, const Slow& c, float inL, float inR, float L1, float R1, const Pred& p, float* out)
    {
        const float eps = 1e-12f, T = 1e-3f;
        m.hpL = ((inL - L1) * c.a) + (m.hpL * c.a);                             // :67
        m.hpR = ((inR - R1) * c.a) + (m.hpR * c.a);
        const float ehf = fmaxf(fabsf(m.hpL), fabsf(m.hpR));                    // :76
        m.env = fmaxf(m.env * c.env_rel, ehf);                                  // :77
        m.base = (m.env * c.base_a) + (m.base * (1.0f - c.base_a));             // :78-79
        const float ratio = m.env / (m.base + eps);
        const float trig = (ratio > c.ratio_thr && p.e_norm > c.err_thr) ? 1.0f : 0.0f;                 // :100
        m.holdEnv = fmaxf(m.holdEnv * c.relHold, trig);                                                // :103-104
        const bool active = m.holdEnv > T;
        const float mix_base = active ? clampf_((p.e_norm - c.err_thr) / (c.range + eps), 0.0f, 1.0f) : 0.0f;
        const float mix = mix_base * c.mix_max;
        const float outL = (p.xC_L * (1.0f - mix)) + (p.pred_L * mix), outR = (p.xC_R * (1.0f - mix)) + (p.pred_R * mix);
        out[0] = c.monitor ? outL - p.xC_L : outL; out[1] = c.monitor ? outR - p.xC_R : outR;
    }
    struct Ring32 { const float* h; int pos; ZA_FL float operator()(int d) const { return h[(pos - d) & 31]; } };
    ZA_FL static void frame(State& m, const Slow& c, const Konst&, const float* in, float* out)
    {
        const float L1 = m.hL[m.pos & 31], R1 = m.hR[m.pos & 31];               // x@1
        m.pos = (m.pos + 1) & 31;
        m.hL[m.pos] = in[0]; m.hR[m.pos] = in[1];
        const Pred p = predict(Ring32{m.hL, m.pos}, Ring32{m.hR, m.pos}, c.mode);
        Rec r{m.hpL, m.hpR, m.env, m.base, m.holdEnv};
        tail(r, c, in[0], in[1], L1, R1, p, out);
        m.hpL = r.hpL; m.hpR = r.hpR; m.env = r.env; m.base = r.base; m.holdEnv = r.holdEnv;
    }
};

// ---- VAR: plugins/Restoration/VAR/src/Vocal Air Recovery (VAR).dsp --------------------------------------------------------------
struct Var {
    static constexpr int kIn = 2, kOut = 2, kSliders = 3;
    static const char* key() { return "VAR"; }
    static const char* slider_text()
    {
        return "slider1:amount=35<0,100,1>Air Amount [%]\n"                    // VAR.dsp:88-90
               "slider2:sens=50<0,100,1>Sensitivity [%]\n"
               "slider3:floor_db=-60<-90,-30,1>Detector Floor [dB]\n";
    }
    struct Slow { float amount_t, sens_t, floorLin, pad; };
    struct Konst { float smoo_p, detSmooth_a, hfLvl_a, c_atk, c_rel, airBase; Bq det, hf, air; };
    struct State { float amount, sens; BqZ detL, detR, hf1L, hf2L, hf1R, hf2R, airZ; float hfLvl, sm1L, sm2L, sm1R, sm2R, s1L, s2L, s1R, s2R, env; int noise; };
    static float safeFc(float fc, float fs) { return std::fmin(fc, 0.45f * fs); }
    static void rbj(float fc, float Q, float fs, int kind, Bq& c)            // :15-76; kind 0 = HP, 2 = BP constant skirt (b0 = sw/2)
    {
        const float f = safeFc(fc, fs), q = std::fmax(0.001f, Q);
        const float w0 = ((2.0f * 3.14159265358979323846f) * f) / fs, cw = std::cos(w0), sw = std::sin(w0), alpha = sw / (2.0f * q);
        float bb0, bb1, bb2;
        if (kind == 0) { bb0 = (1.0f + cw) / 2.0f; bb1 = -(1.0f + cw); bb2 = (1.0f + cw) / 2.0f; }
        else { bb0 = sw / 2.0f; bb1 = 0.0f; bb2 = -sw / 2.0f; }
        const float aa0 = 1.0f + alpha, aa1 = -2.0f * cw, aa2 = 1.0f - alpha;
        c.b0 = bb0 / aa0; c.b1 = bb1 / aa0; c.b2 = bb2 / aa0; c.a1 = aa1 / aa0; c.a2 = aa2 / aa0;
    }
    static void konst(float fs, Konst& k)
    {
        k.smoo_p = 1.0f - (44.1f / fs);
        k.detSmooth_a = std::exp(((-2.0f * 3.14159265358979323846f) * safeFc(8500.0f, fs)) / fs);       // :104
        k.hfLvl_a = std::exp(-1.0f / (fs * 0.14f));                                                    // :123
        k.c_atk = std::exp(-1.0f / (0.0025f * fs)); k.c_rel = std::exp(-1.0f / (0.080f * fs));          // :110-111
        k.airBase = host_db2lin(-34.0f);                                                               // :97
        rbj(9500.0f, 1.0f, fs, 2, k.det); rbj(11500.0f, 0.707f, fs, 0, k.hf); rbj(16000.0f, 1.2f, fs, 2, k.air);
    }
    static void slow(const float* z, float, Slow& c) { c.amount_t = z[0] / 100.0f; c.sens_t = z[1] / 100.0f; c.floorLin = host_db2lin(z[2]); c.pad = 0.0f; }
    ZA_FL static float onepole_exp(float& st, float a, float x) { st = (x * (1.0f - a)) + (st * a); return st; }   // :78
    // frame() in five pieces (as for RED), so that the tiled kernel (p_faust.cu, var_tiled_kernel) runs the memoryless ones lane =
    // frame and the three recurrences lane = stream on a warp each: the same operations on the same values either way.
    struct DetOut { float amount, sens, hfLvl, lapL, denomL, lapR, denomR; };
    ZA_FL static DetOut detector(State& m, const Slow& c, const Konst& k, float inL, float inR)                   // carried
    {
        const float eps = 1e-12f;
        smooth_(m.amount, k.smoo_p, c.amount_t); smooth_(m.sens, k.smoo_p, c.sens_t);                             // :88-89
        const float detL = tf22t_(k.det, m.detL, inL), detR = tf22t_(k.det, m.detR, inR);                        // :117-118
        onepole_exp(m.hfLvl, k.hfLvl_a, 0.5f * (fabsf(detL) + fabsf(detR)));                                     // :121-124
        onepole_exp(m.sm1L, k.detSmooth_a, detL); onepole_exp(m.sm2L, k.detSmooth_a, m.sm1L);                    // :130
        onepole_exp(m.sm1R, k.detSmooth_a, detR); onepole_exp(m.sm2R, k.detSmooth_a, m.sm1R);
        const float s0L = m.sm2L, s1L = m.s1L, s2L = m.s2L, s0R = m.sm2R, s1R = m.s1R, s2R = m.s2R;               // :134-142
        DetOut d;
        d.amount = m.amount; d.sens = m.sens; d.hfLvl = m.hfLvl;
        d.lapL = (s0L - (2.0f * s1L)) + s2L; d.lapR = (s0R - (2.0f * s1R)) + s2R;
        d.denomL = ((fabsf(s0L) + (2.0f * fabsf(s1L))) + fabsf(s2L)) + eps; d.denomR = ((fabsf(s0R) + (2.0f * fabsf(s1R))) + fabsf(s2R)) + eps;
        m.s2L = s1L; m.s1L = s0L; m.s2R = s1R; m.s1R = s0R;
        return d;
    }
    struct Ctl { float curvN, thrN, gate, mE1, airMix; };
    ZA_FL static Ctl control(const Slow& c, const DetOut& d)                                                      // memoryless
    {
        const float eps = 1e-12f;
        Ctl q;
        const float maxExp_lin = db2lin_(5.0f * d.amount);                                                        // :95-100
        q.airMix = 0.25f * d.amount; q.thrN = 0.18f - (0.13f * d.sens);
        const float gu = fminf(1.0f, fmaxf(0.0f, ((d.hfLvl / (c.floorLin + eps)) - 1.0f) / (2.0f - 1.0f)));
        q.gate = (gu * gu) * (3.0f - (2.0f * gu));                                                                // :126
        q.curvN = 0.5f * ((fabsf(d.lapL) / d.denomL) + (fabsf(d.lapR) / d.denomR));                               // :147-149
        q.mE1 = maxExp_lin - 1.0f;
        return q;
    }
    ZA_FL static float envelope(State& m, const Konst& k, float curvN)                                            // carried, :152
    {
        const float cc = curvN > m.env ? k.c_atk : k.c_rel;
        m.env = ((1.0f - cc) * curvN) + (cc * m.env);
        return m.env;
    }
    struct Hf { float hfL, hfR, air; };
    ZA_FL static Hf bands(State& m, const Konst& k, float inL, float inR)                                         // carried
    {
        Hf h;
        h.hfL = tf22t_(k.hf, m.hf2L, tf22t_(k.hf, m.hf1L, inL)); h.hfR = tf22t_(k.hf, m.hf2R, tf22t_(k.hf, m.hf1R, inR));   // :163-164
        m.noise = (int)((unsigned)m.noise * 1103515245u + 12345u);
        h.air = tf22t_(k.air, m.airZ, (float)m.noise / 2147483647.0f);                                            // :170-174 (nL == nR)
        return h;
    }
    ZA_FL static void mix(const Konst& k, const Ctl& q, float env, const Hf& h, float inL, float inR, float* out)   // memoryless
    {
        const float eps = 1e-12f;
        const float u = fmaxf(0.0f, (env / q.thrN) - 1.0f);
        const float t = (u / (1.0f + u)) * q.gate;                                                                // :155-156
        const float t2 = powf(fmaxf(eps, t), 1.8f);                                                               // :157
        const float g = 1.0f + (t * q.mE1);                                                                       // :160
        const float airGain = (t2 * k.airBase) * q.airMix;                                                        // :176
        out[0] = (inL + (h.hfL * (g - 1.0f))) + (h.air * airGain);                                                // :180 (outGain = 1)
        out[1] = (inR + (h.hfR * (g - 1.0f))) + (h.air * airGain);
    }
    ZA_FL static void frame(State& m, const Slow& c, const Konst& k, const float* in, float* out)
    {
        const DetOut d = detector(m, c, k, in[0], in[1]);
        const Ctl q = control(c, d);
        const float env = envelope(m, k, q.curvN);
        const Hf h = bands(m, k, in[0], in[1]);
        mix(k, q, env, h, in[0], in[1], out);
    }
};

}  // namespace faust
}  // namespace za
