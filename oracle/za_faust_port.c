/*
 * oracle/za_faust_port.c -- TEST INFRASTRUCTURE ONLY: plain-C CPU restatement of the catalog's Faust leaves.
 *
 * PARITY UNPINNED.  The arithmetic of a Faust leaf is defined by the Faust compiler's output (`faust -lang cpp -i
 * -cn mydsp`, cmake/plugin/CMakeLists.txt:144; version 2.81.2 pinned only for Windows CI, .github/workflows/
 * release.yml:4,95-98) plus stdfaust.lib, neither of which is under /root/reference, and no reference test renders a
 * Faust leaf.  This file restates the .dsp line by line in float32 (Faust's default -single; FAUSTFLOAT = float,
 * src/faust_support_min.h:4-6), in the .dsp's own expression order, with the library definitions written out:
 *   ba.db2linear(x) = pow(10, x/20)        ma.SR = min(192000, max(1, fSamplingFreq))        ma.PI = float(pi)
 * The order in which the real compiler would associate the float operations is not known, so agreement with the
 * reference build can only be expected to float rounding, not bit for bit.
 * Host contract mirrored: init(sr) clears the recursive state (FaustJuceProcessor.cpp:432); every processBlock writes
 * the parameter zones and calls compute(n) once (FaustJuceProcessor.cpp:463-482) under ScopedNoDenormals; slider-only
 * expressions are evaluated once per compute() call (Faust's "fSlow" values).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may load this library.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>
#include <pmmintrin.h>

#define ZO_API __attribute__((visibility("default")))

typedef struct ZoSliders {           /* same layout as in za_oracle_port.c */
    int32_t n; int32_t index0[64]; int32_t alias[64];
    float vmin[64], vmax[64], vstep[64]; int32_t is_choice[64];
} ZoSliders;

/* the host's parameter -> zone path: the APVTS range snaps to min + step * round((v - min) / step)
 * (juce::NormalisableRange<float>(min, max, step), FaustJuceProcessor.cpp:379), then `*zone = (FAUSTFLOAT) v`
 * (:467-471).  Restated with the same grid arithmetic the C ABI uses for every plugin (double, llround). */
static float zone_value(const ZoSliders* sl, int j, float host_value)
{
    double v = (double)host_value;
    const double mn = (double)sl->vmin[j], mx = (double)sl->vmax[j];
    v = v < mn ? mn : (v > mx ? mx : v);
    const double step = (double)(sl->vstep[j] > 0.0f ? sl->vstep[j] : 0.0f);
    if (step > 0.0) {
        v = mn + (double)llround((v - mn) / step) * step;
        v = v < mn ? mn : (v > mx ? mx : v);
    }
    return (float)v;
}

/* ================================================================================================
 * ModTilt  (plugins/Dynamics/ModTilt/src/ModTilt.dsp)
 * ============================================================================================== */
typedef struct {
    float fs;                                   /* ma.SR */
    float a_env, a_base, a_trim;                /* :35-36,40 (sample-rate constants, init()) */
    float a_piv, g_hi, g_lo, mix;               /* :37,44-45,17-19 (per compute() call) */
    float env2, base, lp, rs, mean;             /* the five onepole() states */
} ModTilt;

static float mt_a_from_hz(float hz, float fs)   /* :25-26  (-2*pi*max(hz,0.001)/SR) : exp : (1 - _) */
{
    const float pi = 3.14159265358979323846f;
    return 1.0f - expf(((-2.0f * pi) * fmaxf(hz, 0.001f)) / fs);
}

static void mt_init(ModTilt* m, double srate)
{
    memset(m, 0, sizeof *m);
    const float sr = (float)(int)srate;
    m->fs = fminf(192000.0f, fmaxf(1.0f, sr));
    m->a_env = mt_a_from_hz(25.0f, m->fs);
    m->a_base = mt_a_from_hz(1.0f, m->fs);
    m->a_trim = mt_a_from_hz(0.2f, m->fs);
}

static void mt_slow(ModTilt* m, float tilt_db, float pivot_hz, float mix)
{
    m->a_piv = mt_a_from_hz(pivot_hz, m->fs);
    m->g_hi = powf(10.0f, (tilt_db * 0.5f) / 20.0f);          /* :44 */
    m->g_lo = powf(10.0f, ((-tilt_db) * 0.5f) / 20.0f);       /* :45 */
    m->mix = mix;
}

/* onepole(a) = *(a) : (+ ~ *(1.0 - a)),  :29 */
#define MT_ONEPOLE(state, a, x) ((state) = ((x) * (a)) + ((state) * (1.0f - (a))))

static void mt_frame(ModTilt* m, float xL, float xR, float* oL, float* oR)
{
    const float depth = 0.75f, a_ratio = 0.05f, eps = 1e-9f;
    const float x = 0.5f * (xL + xR);                                   /* :54 */
    const float x2 = x * x;                                             /* :57 */
    const float env2 = MT_ONEPOLE(m->env2, m->a_env, x2);               /* :58 */
    const float env = sqrtf(fmaxf(env2, 0.0f));                         /* :59 */
    const float base = MT_ONEPOLE(m->base, m->a_base, env);             /* :62 */
    const float mm = env - base;                                        /* :65 */
    const float lp = MT_ONEPOLE(m->lp, m->a_piv, mm);                   /* :68 */
    const float m_hi = mm - lp;                                         /* :70 */
    const float m2_tilt = (lp * m->g_lo) + (m_hi * m->g_hi);            /* :73 */
    const float m2 = (mm * (1.0f - depth)) + (m2_tilt * depth);         /* :74 */
    const float env_t = base + m2;                                      /* :77 */
    const float env_tp = fmaxf(env_t, 0.05f * env);                     /* :80 */
    const float r0 = (env_tp + eps) / (env + eps);                      /* :82 */
    const float r0c = fminf(fmaxf(r0, 0.67f), 1.5f);                    /* :85 */
    const float rs = 1.0f + MT_ONEPOLE(m->rs, a_ratio, r0c - 1.0f);     /* :88 */
    const float rdb = 20.0f * log10f(fmaxf(rs, 1e-12f));                /* :92 */
    const float mean = MT_ONEPOLE(m->mean, m->a_trim, rdb);             /* :93 */
    const float trim = powf(10.0f, (-mean) / 20.0f);                    /* :94 */
    const float yL = xL * rs, yR = xR * rs;                             /* :97-98 */
    *oL = ((xL * (1.0f - m->mix)) + (yL * m->mix)) * trim;              /* :100 */
    *oR = ((xR * (1.0f - m->mix)) + (yR * m->mix)) * trim;              /* :101 */
}

/* coefficient rows for the product's host half to be checked against (tests only) */
ZO_API void zo_modtilt_coefs(double srate, float tilt_db, float pivot_hz, float mix, float* out8)
{
    ModTilt m; mt_init(&m, srate); mt_slow(&m, tilt_db, pivot_hz, mix);
    out8[0] = m.a_env; out8[1] = m.a_base; out8[2] = m.a_trim; out8[3] = m.a_piv;
    out8[4] = m.g_hi; out8[5] = m.g_lo; out8[6] = m.mix; out8[7] = m.fs;
}

typedef struct Job {
    const ZoSliders* sl; double srate; const float* params; int n_param_blocks;
    const float* in; float* out; double* out_d; int n_streams, block; int64_t n_frames, frame_stride; int s0, s1;
} Job;

static void* mt_worker(void* arg)
{
    Job* j = (Job*)arg;
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    const int nsl = j->sl->n;
    for (int s = j->s0; s < j->s1; ++s) {
        ModTilt m; mt_init(&m, j->srate);
        const float* xL = j->in + (size_t)s * 2 * j->frame_stride; const float* xR = xL + j->frame_stride;
        float* yL = j->out + (size_t)s * 2 * j->frame_stride; float* yR = yL + j->frame_stride;
        double* dL = j->out_d ? j->out_d + (size_t)s * 2 * j->frame_stride : NULL;
        int64_t pos = 0; int b = 0;
        while (pos < j->n_frames) {
            const int pb = j->n_param_blocks > 1 ? (b < j->n_param_blocks ? b : j->n_param_blocks - 1) : 0;
            const float* hv = j->params + ((size_t)pb * j->n_streams + s) * nsl;
            mt_slow(&m, zone_value(j->sl, 0, hv[0]), zone_value(j->sl, 1, hv[1]), zone_value(j->sl, 2, hv[2]));
            const int n = (int)((j->n_frames - pos) < j->block ? (j->n_frames - pos) : j->block);
            for (int i = 0; i < n; ++i) {
                float oL, oR;
                mt_frame(&m, xL[pos + i], xR[pos + i], &oL, &oR);
                yL[pos + i] = oL; yR[pos + i] = oR;
                if (dL) { dL[pos + i] = (double)oL; dL[j->frame_stride + pos + i] = (double)oR; }
            }
            pos += n; ++b;
        }
    }
    return NULL;
}

ZO_API int zo_modtilt_render(const ZoSliders* sl, double srate, const float* params, int n_param_blocks,
                             const float* in, float* out, double* out_d, int n_streams, int64_t n_frames,
                             int64_t frame_stride, int block, int n_threads)
{
    if (sl->n != 3) return -2;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_streams) n_threads = n_streams > 0 ? n_streams : 1;
    if (n_threads > 512) n_threads = 512;
    pthread_t th[512]; Job jobs[512];
    for (int t = 0; t < n_threads; ++t) {
        Job* j = &jobs[t];
        j->sl = sl; j->srate = srate; j->params = params; j->n_param_blocks = n_param_blocks;
        j->in = in; j->out = out; j->out_d = out_d; j->n_streams = n_streams; j->block = block;
        j->n_frames = n_frames; j->frame_stride = frame_stride;
        j->s0 = (int)((int64_t)n_streams * t / n_threads);
        j->s1 = (int)((int64_t)n_streams * (t + 1) / n_threads);
        if (n_threads == 1) { mt_worker(j); return 0; }
        if (pthread_create(&th[t], NULL, mt_worker, j) != 0) return -1;
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
    return 0;
}


/* ================================================================================================
 * Shared pieces of the other four Faust leaves.  Library definitions written out (recalled from
 * faustlibraries, to be re-verified against Faust 2.81.2 should a copy ever be available -- PARITY UNPINNED):
 *   si.smooth(s)      = *(1 - s) : + ~ *(s)                       y = x (1 - s) + y' s
 *   si.smoo           = si.smooth(1 - 44.1 / ma.SR)               (signals.lib >= 1.3; older libraries: 0.999)
 *   ba.tau2pole(t)    = exp(-1 / (t ma.SR))
 *   si.onePoleSwitching(att, rel, x): y = (1 - c) x + c y', c = x > y' ? tau2pole(att) : tau2pole(rel)
 *   an.amp_follower_ar(att, rel) = abs : si.onePoleSwitching(att, rel)
 *   fi.tf22t(b0,b1,b2,a1,a2): transposed direct form II   y = b0 x + s1; s1 = b1 x - a1 y + s2; s2 = b2 x - a2 y
 *   fi.fir(c0..cN)    = sum_k c_k x@k (summed left to right)      de.delay(n, d) = @(min(n, max(0, d)))
 *   ba.db2linear(x)   = pow(10, x / 20)      ba.if(c, t, e) / ba.selectn: selectors
 *   no.noise          = (+(12345) ~ *(1103515245)) / 2147483647.0   (int32 wrap-around)
 * ============================================================================================== */
static float fs_of(double srate) { return fminf(192000.0f, fmaxf(1.0f, (float)(int)srate)); }
static float clampf(float x, float lo, float hi) { return fmaxf(lo, fminf(hi, x)); }
static float db2lin(float db) { return powf(10.0f, db / 20.0f); }
#define SMOOTH(state, pole, x) ((state) = ((x) * (1.0f - (pole))) + ((state) * (pole)))
typedef struct { float b0, b1, b2, a1, a2; } Bq;
typedef struct { float s1, s2; } BqZ;
static float tf22t(const Bq* c, BqZ* z, float x)
{
    const float y = (c->b0 * x) + z->s1;
    z->s1 = ((c->b1 * x) - (c->a1 * y)) + z->s2;
    z->s2 = (c->b2 * x) - (c->a2 * y);
    return y;
}

/* ================================================================================================
 * RED  (plugins/Dynamics/RED/src/Reverb Expanding Downwards (RED).dsp) -- 6 in / 6 out: wet L/R, ch3, ch4, ref L/R
 * ============================================================================================== */
typedef struct {
    float fs;
    float thr_db, ratio, knee_db, maxduck, pole_rms, pole_tgt, pole_grace, pole_hold, pole_dry, c_atk, c_rel, c_rel_in;   /* slider rate */
    float wet_env2, ref_env2, offA_s, tgt_db, dryA_s, tgt_hold_s, gr_norm, gr_fast;                                      /* state */
} Red;
static float ms2pole(float ms, float fs) { return expf(-1.0f / (fs * (ms / 1000.0f))); }                     /* :24 */
static float smoothstep01(float x) { const float x1 = clampf(x, 0.0f, 1.0f); return (x1 * x1) * (3.0f - (2.0f * x1)); }   /* :17 */
static void red_init(void* st, double srate) { Red* m = (Red*)st; memset(m, 0, sizeof *m); m->fs = fs_of(srate); }
static void red_slow(void* st, const float* z)        /* z: amount_dB, sens_pct, rel_ms_ui (:8-10) */
{
    Red* m = (Red*)st;
    const float sens = z[1] / 100.0f, rel_ms = z[2];
    m->maxduck = z[0];
    m->thr_db = 18.0f - (sens * 21.0f); m->ratio = 1.2f + (sens * 3.0f); m->knee_db = 10.0f - (sens * 6.0f);    /* :49-51 */
    const float grace_ms = clampf(rel_ms * 0.25f, 60.0f, 200.0f);                                              /* :54 */
    m->pole_rms = ms2pole(35.0f, m->fs); m->pole_tgt = ms2pole(25.0f, m->fs); m->pole_grace = ms2pole(grace_ms, m->fs);
    m->pole_hold = ms2pole(80.0f, m->fs); m->pole_dry = ms2pole(10.0f, m->fs);
    m->c_atk = expf(-1.0f / ((12.0f / 1000.0f) * m->fs));                                                      /* ba.tau2pole(att_s) */
    m->c_rel = expf(-1.0f / ((rel_ms / 1000.0f) * m->fs));
    m->c_rel_in = expf(-1.0f / ((70.0f / 1000.0f) * m->fs));
}
static void red_frame(void* st, const float* in, float* out)
{
    Red* m = (Red*)st;
    const float eps = 1e-12f, floor_lin = db2lin(-80.0f), dry_on_lin = db2lin(-50.0f), ref_off_lin = db2lin(-60.0f);
    const float wetL = in[0], wetR = in[1], refL = in[4], refR = in[5];
    const float wet_p = 0.5f * ((wetL * wetL) + (wetR * wetR)), ref_p = 0.5f * ((refL * refL) + (refR * refR));   /* :68-69 */
    SMOOTH(m->wet_env2, m->pole_rms, wet_p); SMOOTH(m->ref_env2, m->pole_rms, ref_p);                       /* :72-73 */
    const float Ey = fmaxf(sqrtf(fmaxf(m->wet_env2, 0.0f)), floor_lin), Ex = fmaxf(sqrtf(fmaxf(m->ref_env2, 0.0f)), floor_lin);
    const float dryA = Ex > dry_on_lin ? 1.0f : 0.0f, offA = Ex <= ref_off_lin ? 1.0f : 0.0f;                 /* :78-79 */
    SMOOTH(m->offA_s, m->pole_grace, offA);                                                                  /* :83 */
    const float tail_w = (1.0f - offA) + (offA * smoothstep01(m->offA_s));                                   /* :84 */
    const float rdB = 20.0f * log10f(fmaxf((Ey + eps) / (Ex + eps), 1e-30f));                                /* :87 */
    const float over = rdB - m->thr_db;
    const float over_eff = over <= 0.0f ? 0.0f : over * smoothstep01(clampf(over / fmaxf(m->knee_db, 0.001f), 0.0f, 1.0f));   /* :91-94 */
    const float tgt0 = over_eff > 0.0f ? fminf(m->maxduck, over_eff * m->ratio) : 0.0f;                        /* :97 */
    const float tgt1 = tgt0 * tail_w;
    SMOOTH(m->tgt_db, m->pole_tgt, tgt1);                                                                    /* :101 */
    SMOOTH(m->dryA_s, m->pole_dry, dryA);                                                                    /* :107 */
    SMOOTH(m->tgt_hold_s, m->pole_hold, m->tgt_db);
    const float tgt_hold = fmaxf(m->tgt_db, m->tgt_hold_s);                                                  /* :112 */
    const float tgt_pin = ((1.0f - dryA) * tgt_hold) + (dryA * m->tgt_db);                                   /* :113 */
    const float x = fabsf(tgt_pin);
    { const float c = x > m->gr_norm ? m->c_atk : m->c_rel; m->gr_norm = ((1.0f - c) * x) + (c * m->gr_norm); }        /* :120 */
    { const float c = x > m->gr_fast ? m->c_atk : m->c_rel_in; m->gr_fast = ((1.0f - c) * x) + (c * m->gr_fast); }      /* :121 */
    const float gr_db = ((1.0f - m->dryA_s) * m->gr_norm) + (m->dryA_s * m->gr_fast);                         /* :124 */
    const float g = db2lin(-gr_db);                                                                          /* :127 */
    out[0] = wetL * g; out[1] = wetR * g; out[2] = in[2]; out[3] = in[3]; out[4] = refL; out[5] = refR;      /* :43 */
}

/* ================================================================================================
 * ClickBeGoneSG  (plugins/Restoration/ClickBeGoneSG/src/Click-Be-Gone (SG).dsp)
 * ============================================================================================== */
typedef struct {
    float fs;
    float a, env_rel, base_a, ratio_thr, err_thr, mix_max, relHold, range; int mode, monitor;     /* slider rate */
    float hL[32], hR[32]; int pos;                                                                /* L@k, R@k: h[(pos - k) & 31] */
    float hpL, hpR, env, base, holdEnv;
} Cbg;
static void cbg_init(void* st, double srate) { Cbg* m = (Cbg*)st; memset(m, 0, sizeof *m); m->fs = fs_of(srate); }
static void cbg_slow(void* st, const float* z)        /* z: Amount %, Sensitivity %, HPF Hz, Mode, Monitor (:11-15) */
{
    Cbg* m = (Cbg*)st;
    static const float ratio_mul[3] = {1.12f, 1.00f, 0.92f}, err_mul[3] = {1.18f, 1.00f, 0.90f}, mix_mul[3] = {0.85f, 1.00f, 1.08f},
                       hold_mul[3] = {0.75f, 1.00f, 1.35f}, base_mul[3] = {0.85f, 1.00f, 1.10f}, env_mul[3] = {0.85f, 1.00f, 1.10f};
    const float amount = z[0] / 100.0f, sens = z[1] / 100.0f, eps = 1e-12f;
    const int md = (int)z[3] < 0 ? 0 : ((int)z[3] > 2 ? 2 : (int)z[3]);
    m->mode = md; m->monitor = z[4] != 0.0f;
    const float env_rel_ms = (30.0f - (20.0f * sens)) * env_mul[md], base_ms = (300.0f - (180.0f * sens)) * base_mul[md];   /* :28-35 */
    m->ratio_thr = (6.0f - (4.0f * sens)) * ratio_mul[md]; m->err_thr = (0.25f - (0.17f * sens)) * err_mul[md];            /* :20-21,37-38 */
    m->mix_max = fminf((0.60f + (0.32f * amount)) * mix_mul[md], 0.96f);                                                   /* :40-41 */
    const float holdN = fmaxf((8.0f + (amount * 32.0f)) * hold_mul[md], 4.0f);                                             /* :43-44 */
    m->env_rel = expf(-1000.0f / (m->fs * env_rel_ms)); m->base_a = 1.0f - expf(-1000.0f / (m->fs * base_ms));             /* :46-47 */
    m->a = expf((-2.0f * 3.14159265358979323846f * z[2]) / m->fs);                                                         /* :49 */
    m->relHold = expf(logf(1e-3f) / (holdN + eps));                                                                        /* :102 */
    m->range = m->err_thr * 3.0f;                                                                                          /* :108 */
}
#define CBG_AT(h, k) ((h)[(m->pos - (k)) & 31])
static float cbg_sg(const Cbg* m, const float* h, int which)     /* :53-59: Savitzky-Golay predictors on delayed taps, summed left to right */
{
    static const float c11[11] = {-36, 9, 44, 69, 84, 89, 84, 69, 44, 9, -36};
    static const float c15[15] = {-78, -13, 42, 87, 122, 147, 162, 167, 162, 147, 122, 87, 42, -13, -78};
    static const float c21[21] = {-171, -76, 9, 84, 149, 204, 249, 284, 309, 324, 329, 324, 309, 284, 249, 204, 149, 84, 9, -76, -171};
    static const float c31[31] = {-406, -261, -126, -1, 114, 219, 314, 399, 474, 539, 594, 639, 674, 699, 714, 719, 714, 699, 674, 639, 594, 539,
                                  474, 399, 314, 219, 114, -1, -126, -261, -406};
    const float* c; int n, d0; float den;
    switch (which) { case 0: c = c11; n = 11; d0 = 20; den = 429.0f; break; case 1: c = c15; n = 15; d0 = 22; den = 1105.0f; break;
                     case 2: c = c21; n = 21; d0 = 25; den = 3059.0f; break; default: c = c31; n = 31; d0 = 30; den = 9889.0f; break; }
    float acc = c[0] * CBG_AT(h, d0);
    for (int k = 1; k < n; ++k) acc = acc + (c[k] * CBG_AT(h, d0 - k));
    return acc / den;
}
static void cbg_frame(void* st, const float* in, float* out)
{
    Cbg* m = (Cbg*)st;
    const float eps = 1e-12f, T = 1e-3f;
    const float L1 = CBG_AT(m->hL, 0), R1 = CBG_AT(m->hR, 0);              /* x@1: the newest stored sample */
    m->pos = (m->pos + 1) & 31;
    m->hL[m->pos] = in[0]; m->hR[m->pos] = in[1];
    m->hpL = ((in[0] - L1) * m->a) + (m->hpL * m->a);                       /* :67  hpf_jsfx */
    m->hpR = ((in[1] - R1) * m->a) + (m->hpR * m->a);
    const float ehf = fmaxf(fabsf(m->hpL), fabsf(m->hpR));                  /* :76 */
    m->env = fmaxf(m->env * m->env_rel, ehf);                               /* :77  max ~ *(env_rel) */
    m->base = (m->env * m->base_a) + (m->base * (1.0f - m->base_a));        /* :78-79 */
    const float ratio = m->env / (m->base + eps);
    const float xC_L = CBG_AT(m->hL, 15), xC_R = CBG_AT(m->hR, 15);         /* :83-84 */
    const float small_L = cbg_sg(m, m->hL, m->mode), small_R = cbg_sg(m, m->hR, m->mode);           /* :86-87 */
    const float large_L = cbg_sg(m, m->hL, m->mode + 1), large_R = cbg_sg(m, m->hR, m->mode + 1);   /* :88-89 */
    const float eA = fmaxf(fabsf(xC_L - small_L), fabsf(xC_R - small_R)) / (fmaxf(fabsf(small_L), fabsf(small_R)) + 1e-6f);
    const float eB = fmaxf(fabsf(xC_L - large_L), fabsf(xC_R - large_R)) / (fmaxf(fabsf(large_L), fabsf(large_R)) + 1e-6f);
    const int useA = eA <= eB;
    const float pred_L = useA ? small_L : large_L, pred_R = useA ? small_R : large_R, e_norm = useA ? eA : eB;
    const float trig = (float)((ratio > m->ratio_thr) * (e_norm > m->err_thr));                       /* :100 */
    m->holdEnv = fmaxf(m->holdEnv * m->relHold, trig);                                                /* :103-104 */
    const int active = m->holdEnv > T;
    const float mix_base = active ? clampf((e_norm - m->err_thr) / (m->range + eps), 0.0f, 1.0f) : 0.0f;
    const float mix = mix_base * m->mix_max;
    const float outL = (xC_L * (1.0f - mix)) + (pred_L * mix), outR = (xC_R * (1.0f - mix)) + (pred_R * mix);
    out[0] = m->monitor ? outL - xC_L : outL; out[1] = m->monitor ? outR - xC_R : outR;
}

/* ================================================================================================
 * VAR  (plugins/Restoration/VAR/src/Vocal Air Recovery (VAR).dsp)
 * ============================================================================================== */
typedef struct {
    float fs, smoo_p, detSmooth_a, hfLvl_a, c_atk, c_rel, airBase;
    Bq det, hf, air;
    float amount_t, sens_t, floorLin;                        /* slider rate: zone / 100 (smoothed per sample below), db2linear(floor) */
    float amount, sens;                                      /* si.smoo states */
    BqZ detL, detR, hf1L, hf2L, hf1R, hf2R, airZ;
    float hfLvl, sm1L, sm2L, sm1R, sm2R, s1L, s2L, s1R, s2R, env;
    int32_t noise;
} Var;
static float safeFc(float fc, float fs) { return fminf(fc, 0.45f * fs); }                                        /* :12 */
static void rbj(float fc, float Q, float fs, int kind, Bq* c)      /* :15-76: kind 0 = HP, 1 = LP, 2 = BP (constant skirt, b0 = sw/2) */
{
    const float f = safeFc(fc, fs), q = fmaxf(0.001f, Q);
    const float w0 = ((2.0f * 3.14159265358979323846f) * f) / fs, cw = cosf(w0), sw = sinf(w0), alpha = sw / (2.0f * q);
    float bb0, bb1, bb2;
    if (kind == 0) { bb0 = (1.0f + cw) / 2.0f; bb1 = -(1.0f + cw); bb2 = (1.0f + cw) / 2.0f; }
    else if (kind == 1) { bb0 = (1.0f - cw) / 2.0f; bb1 = 1.0f - cw; bb2 = (1.0f - cw) / 2.0f; }
    else { bb0 = sw / 2.0f; bb1 = 0.0f; bb2 = -sw / 2.0f; }
    const float aa0 = 1.0f + alpha, aa1 = -2.0f * cw, aa2 = 1.0f - alpha;
    c->b0 = bb0 / aa0; c->b1 = bb1 / aa0; c->b2 = bb2 / aa0; c->a1 = aa1 / aa0; c->a2 = aa2 / aa0;
}
static void var_init(void* st, double srate)
{
    Var* m = (Var*)st; memset(m, 0, sizeof *m);
    m->fs = fs_of(srate);
    m->smoo_p = 1.0f - (44.1f / m->fs);
    m->detSmooth_a = expf(((-2.0f * 3.14159265358979323846f) * safeFc(8500.0f, m->fs)) / m->fs);                /* :104 */
    m->hfLvl_a = expf(-1.0f / (m->fs * 0.14f));                                                               /* :123 */
    m->c_atk = expf(-1.0f / (0.0025f * m->fs)); m->c_rel = expf(-1.0f / (0.080f * m->fs));                     /* :110-111 */
    m->airBase = db2lin(-34.0f);                                                                              /* :97 */
    rbj(9500.0f, 1.0f, m->fs, 2, &m->det); rbj(11500.0f, 0.707f, m->fs, 0, &m->hf); rbj(16000.0f, 1.2f, m->fs, 2, &m->air);
}
static void var_slow(void* st, const float* z) { Var* m = (Var*)st; m->amount_t = z[0] / 100.0f; m->sens_t = z[1] / 100.0f; m->floorLin = db2lin(z[2]); }
#define ONEPOLE_EXP(state, a, x) ((state) = ((x) * (1.0f - (a))) + ((state) * (a)))           /* :78 */
static void var_frame(void* st, const float* in, float* out)
{
    Var* m = (Var*)st;
    const float eps = 1e-12f, inL = in[0], inR = in[1];
    SMOOTH(m->amount, m->smoo_p, m->amount_t); SMOOTH(m->sens, m->smoo_p, m->sens_t);                          /* :88-89 */
    const float maxExp_lin = db2lin(5.0f * m->amount), airMix = 0.25f * m->amount, thrN = 0.18f - (0.13f * m->sens);   /* :95-100 */
    const float detL = tf22t(&m->det, &m->detL, inL), detR = tf22t(&m->det, &m->detR, inR);                  /* :117-118 */
    const float hfAbs = 0.5f * (fabsf(detL) + fabsf(detR));
    ONEPOLE_EXP(m->hfLvl, m->hfLvl_a, hfAbs);                                                               /* :122-124 */
    const float gu = fminf(1.0f, fmaxf(0.0f, ((m->hfLvl / (m->floorLin + eps)) - 1.0f) / (2.0f - 1.0f)));
    const float gate = (gu * gu) * (3.0f - (2.0f * gu));                                                     /* :126 */
    ONEPOLE_EXP(m->sm1L, m->detSmooth_a, detL); const float p2L = m->sm2L; ONEPOLE_EXP(m->sm2L, m->detSmooth_a, m->sm1L);   /* :130 */
    ONEPOLE_EXP(m->sm1R, m->detSmooth_a, detR); const float p2R = m->sm2R; ONEPOLE_EXP(m->sm2R, m->detSmooth_a, m->sm1R);
    (void)p2L; (void)p2R;
    const float s0L = m->sm2L, s1L = m->s1L, s2L = m->s2L, s0R = m->sm2R, s1R = m->s1R, s2R = m->s2R;          /* :134-142 */
    const float lapL = (s0L - (2.0f * s1L)) + s2L, lapR = (s0R - (2.0f * s1R)) + s2R;
    const float denomL = ((fabsf(s0L) + (2.0f * fabsf(s1L))) + fabsf(s2L)) + eps, denomR = ((fabsf(s0R) + (2.0f * fabsf(s1R))) + fabsf(s2R)) + eps;
    m->s2L = s1L; m->s1L = s0L; m->s2R = s1R; m->s1R = s0R;
    const float curvN = 0.5f * ((fabsf(lapL) / denomL) + (fabsf(lapR) / denomR));                             /* :147-149 */
    { const float c = curvN > m->env ? m->c_atk : m->c_rel; m->env = ((1.0f - c) * curvN) + (c * m->env); }   /* :152 */
    const float u = fmaxf(0.0f, (m->env / thrN) - 1.0f);
    const float t = (u / (1.0f + u)) * gate;                                                                 /* :155-156 */
    const float t2 = powf(fmaxf(eps, t), 1.8f);                                                              /* :157 */
    const float g = 1.0f + (t * (maxExp_lin - 1.0f));                                                        /* :160 */
    const float hfL = tf22t(&m->hf, &m->hf2L, tf22t(&m->hf, &m->hf1L, inL)), hfR = tf22t(&m->hf, &m->hf2R, tf22t(&m->hf, &m->hf1R, inR));   /* :163-164 */
    m->noise = (int32_t)((uint32_t)m->noise * 1103515245u + 12345u);
    const float n = (float)m->noise / 2147483647.0f;                                                         /* :170-171: nL == nR */
    const float air = tf22t(&m->air, &m->airZ, n);
    const float airGain = (t2 * m->airBase) * airMix;                                                        /* :176 */
    out[0] = (inL + (hfL * (g - 1.0f))) + (air * airGain);                                                   /* :180 (outGain = 1) */
    out[1] = (inR + (hfR * (g - 1.0f))) + (air * airGain);
}

/* ================================================================================================
 * GTS  (plugins/Dynamics/GTS/src/Gaussian Transient Shaper (GTS).dsp) -- 257-tap Gaussian FIR whose sigma is a smoothed
 * slider, i.e. the kernel is rebuilt every sample; latency 128 frames
 * ============================================================================================== */
typedef struct {
    float fs, smoo_p;
    float t[5];                       /* slider rate targets: sigma ms, attack dB, sustain dB, mix, db2linear(out gain dB) */
    float s[5];                       /* the five si.smoo states (both channels share the widgets, hence the values) */
    float hL[512], hR[512]; int pos;  /* x@k = h[(pos - k) & 511] */
} Gts;
static void gts_init(void* st, double srate) { Gts* m = (Gts*)st; memset(m, 0, sizeof *m); m->fs = fs_of(srate); m->smoo_p = 1.0f - (44.1f / m->fs); }
static void gts_slow(void* st, const float* z) { Gts* m = (Gts*)st; m->t[0] = z[0]; m->t[1] = z[1]; m->t[2] = z[2]; m->t[3] = z[3]; m->t[4] = db2lin(z[4]); }
static void gts_frame(void* st, const float* in, float* out)
{
    Gts* m = (Gts*)st;
    for (int k = 0; k < 5; ++k) SMOOTH(m->s[k], m->smoo_p, m->t[k]);                                         /* :52-64 */
    m->pos = (m->pos + 1) & 511;
    m->hL[m->pos] = in[0]; m->hR[m->pos] = in[1];
    const float sigma = fmaxf(0.25f, (m->s[0] * m->fs) * 0.001f);                                            /* :68-69 */
    float g[129];
    for (int i = 0; i <= 128; ++i) { const float r = (float)i / sigma; g[i] = expf(-0.5f * (r * r)); }       /* :26 */
    float sumRest = 0.0f;
    for (int i = 0; i < 128; ++i) sumRest = sumRest + g[i + 1];                                              /* :30 */
    const float norm = 1.0f / ((g[0] + (2.0f * sumRest)) + 1e-20f);                                          /* :31 */
    const float aGain = db2lin(m->s[1]), sGain = db2lin(m->s[2]), mix = m->s[3], og = m->s[4];
    for (int c = 0; c < 2; ++c) {
        const float* h = c ? m->hR : m->hL;
        float sustain = (norm * g[128]) * h[m->pos & 511];
        for (int k = 1; k <= 256; ++k) sustain = sustain + ((norm * g[k < 128 ? 128 - k : k - 128]) * h[(m->pos - k) & 511]);   /* :35,40 */
        const float xAligned = h[(m->pos - 128) & 511];                                                     /* :80 */
        const float attack = xAligned - sustain;
        const float shaped = (aGain * attack) + (sGain * sustain);                                          /* :88 */
        out[c] = ((mix * shaped) + ((1.0f - mix) * xAligned)) * og;                                         /* :91 */
    }
}

/* ---- generic driver: init(sr), zones per compute() call, frame by frame ---------------------------------------------------- */
typedef struct LeafOps {
    const char* key; int n_in, n_out, n_sliders; size_t state_bytes;
    void (*init)(void*, double); void (*slow)(void*, const float*); void (*frame)(void*, const float*, float*);
} LeafOps;
static const LeafOps kLeaves[] = {
    {"RED", 6, 6, 3, sizeof(Red), red_init, red_slow, red_frame},
    {"ClickBeGoneSG", 2, 2, 5, sizeof(Cbg), cbg_init, cbg_slow, cbg_frame},
    {"VAR", 2, 2, 3, sizeof(Var), var_init, var_slow, var_frame},
    {"GTS", 2, 2, 5, sizeof(Gts), gts_init, gts_slow, gts_frame},
};
typedef struct LJob { const LeafOps* ops; Job j; } LJob;
static void* leaf_worker(void* arg)
{
    LJob* lj = (LJob*)arg; Job* j = &lj->j; const LeafOps* o = lj->ops;
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    void* st = malloc(o->state_bytes);
    for (int s = j->s0; s < j->s1; ++s) {
        o->init(st, j->srate);
        const float* x = j->in + (size_t)s * o->n_in * j->frame_stride;
        float* y = j->out + (size_t)s * o->n_out * j->frame_stride;
        double* d = j->out_d ? j->out_d + (size_t)s * o->n_out * j->frame_stride : NULL;
        int64_t pos = 0; int b = 0;
        while (pos < j->n_frames) {
            const int pb = j->n_param_blocks > 1 ? (b < j->n_param_blocks ? b : j->n_param_blocks - 1) : 0;
            const float* hv = j->params + ((size_t)pb * j->n_streams + s) * o->n_sliders;
            float z[8];
            for (int k = 0; k < o->n_sliders; ++k) z[k] = zone_value(j->sl, k, hv[k]);
            o->slow(st, z);
            const int n = (int)((j->n_frames - pos) < j->block ? (j->n_frames - pos) : j->block);
            for (int i = 0; i < n; ++i) {
                float fi[8], fo[8];
                for (int c = 0; c < o->n_in; ++c) fi[c] = x[(size_t)c * j->frame_stride + pos + i];
                o->frame(st, fi, fo);
                for (int c = 0; c < o->n_out; ++c) {
                    y[(size_t)c * j->frame_stride + pos + i] = fo[c];
                    if (d) d[(size_t)c * j->frame_stride + pos + i] = (double)fo[c];
                }
            }
            pos += n; ++b;
        }
    }
    free(st);
    return NULL;
}
static int leaf_render(const LeafOps* o, const ZoSliders* sl, double srate, const float* params, int n_param_blocks, const float* in, float* out,
                       double* out_d, int n_streams, int64_t n_frames, int64_t frame_stride, int block, int n_threads)
{
    if (sl->n != o->n_sliders) return -2;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_streams) n_threads = n_streams > 0 ? n_streams : 1;
    if (n_threads > 512) n_threads = 512;
    pthread_t th[512]; static __thread LJob jobs[512];
    for (int t = 0; t < n_threads; ++t) {
        LJob* lj = &jobs[t]; Job* j = &lj->j;
        lj->ops = o;
        j->sl = sl; j->srate = srate; j->params = params; j->n_param_blocks = n_param_blocks;
        j->in = in; j->out = out; j->out_d = out_d; j->n_streams = n_streams; j->block = block;
        j->n_frames = n_frames; j->frame_stride = frame_stride;
        j->s0 = (int)((int64_t)n_streams * t / n_threads);
        j->s1 = (int)((int64_t)n_streams * (t + 1) / n_threads);
        if (n_threads == 1) { leaf_worker(lj); return 0; }
        if (pthread_create(&th[t], NULL, leaf_worker, lj) != 0) return -1;
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
    return 0;
}
#define ZO_LEAF(fn, idx)                                                                                                              \
    ZO_API int fn(const ZoSliders* sl, double srate, const float* params, int n_param_blocks, const float* in, float* out, double* out_d, \
                  int n_streams, int64_t n_frames, int64_t frame_stride, int block, int n_threads)                                    \
    { return leaf_render(&kLeaves[idx], sl, srate, params, n_param_blocks, in, out, out_d, n_streams, n_frames, frame_stride, block, n_threads); }
ZO_LEAF(zo_red_render, 0)
ZO_LEAF(zo_clickbegonesg_render, 1)
ZO_LEAF(zo_var_render, 2)
ZO_LEAF(zo_gts_render, 3)
