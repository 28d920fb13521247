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
