/*
 * oracle/za_oracle_port.c -- TEST INFRASTRUCTURE ONLY: plain-C CPU restatement of the hot path.
 *
 * Sequential, one state per stream, double arithmetic in the reference's expression order
 * (compile with -ffp-contract=off: the reference's AOT targets generic x86-64 without FMA and
 * without fast-math, dsp_jsfx_aot.py:6211-6228).  PINNED: tests/test_oracle.py checks every
 * function here bit-for-bit (doubles before the float store) against the reference's own compiled
 * object (oracle/_ref, built from /root/reference by oracle/build_ref.py) and against the golden
 * vectors in tests/golden/ that were generated from that object.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference when
 * oracle/_ref is absent) may load this library.  The product never does.
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

typedef struct ZoSliders {           /* same layout as ZrefSliders in ref_shim.c */
    int32_t n; int32_t index0[64]; int32_t alias[64];
    float vmin[64], vmax[64], vstep[64]; int32_t is_choice[64];
} ZoSliders;

/* hostParameterToJsfxSliderValue, src/JSFXJuceProcessor.cpp:5556-5597 */
ZO_API double zo_quantise(const ZoSliders* sl, int j, float host_value)
{
    double v = (double)host_value;
    const double mn = (double)sl->vmin[j], mx = (double)sl->vmax[j];
    if (sl->is_choice[j]) v = mn + (double)(int64_t)llround(v) * (double)sl->vstep[j];
    v = v < mn ? mn : (v > mx ? mx : v);
    if (!sl->is_choice[j]) {
        const double step = (double)(sl->vstep[j] > 0.0f ? sl->vstep[j] : 0.0f);
        if (step > 0.0) {
            v = mn + (double)llround((v - mn) / step) * step;
            v = v < mn ? mn : (v > mx ? mx : v);
        }
    }
    return v;
}

/* ================================================================================================
 * cascade8  (zorakaudio-experimental-plugins_b200/plugins/Workloads/cascade8/src/cascade8.jsfx;
 *            section forms: ATTACK.jsfx:185-219, Spectral Stabilizer.jsfx:73-111, VAR.dsp:15-76,
 *            ERB Tilt.jsfx:151,397, EasyExpander.jsfx:121,186)
 * ============================================================================================== */
typedef struct { double b0, b1, b2, a1, a2; } Bq;
typedef struct {
    Bq s[6]; double a7, c8, mix;
    double z[6][2]; double lp7, lp8;
} Cascade8;

static double jmin(double a, double b) { return a < b ? a : b; }   /* min: olt select, dsp_jsfx_aot.py:5223-5232 */
static double jmax(double a, double b) { return a > b ? a : b; }

enum { BQ_HP, BQ_LP, BQ_BP_PEAK, BQ_BP_SKIRT };
static const int kKinds[6] = { BQ_HP, BQ_LP, BQ_BP_PEAK, BQ_BP_SKIRT, BQ_LP, BQ_HP };

static void c8_slider(Cascade8* c, const double* sl /* slider1..15 */, double srate)
{
    const double two_pi = 6.283185307179586;
    for (int k = 0; k < 6; ++k) {
        const double f = jmin(jmax(sl[2 * k], 10), srate * 0.45), q = jmax(sl[2 * k + 1], 0.1);
        const double w0 = two_pi * f / srate, cw = cos(w0), sw = sin(w0), alpha = sw / (2.0 * q);
        const double aa0 = 1.0 + alpha;
        Bq* b = &c->s[k];
        switch (kKinds[k]) {
        case BQ_HP: b->b0 = ((1.0 + cw) / 2.0) / aa0; b->b1 = (-(1.0 + cw)) / aa0; b->b2 = ((1.0 + cw) / 2.0) / aa0; break;
        case BQ_LP: b->b0 = ((1.0 - cw) / 2.0) / aa0; b->b1 = (1.0 - cw) / aa0;    b->b2 = ((1.0 - cw) / 2.0) / aa0; break;
        case BQ_BP_PEAK:  b->b0 = alpha / aa0;      b->b1 = 0.0; b->b2 = (-alpha) / aa0;    break;
        case BQ_BP_SKIRT: b->b0 = (sw / 2.0) / aa0; b->b1 = 0.0; b->b2 = (-sw / 2.0) / aa0; break;
        }
        b->a1 = (-2.0 * cw) / aa0; b->a2 = (1.0 - alpha) / aa0;
    }
    c->a7 = exp(-two_pi * jmin(sl[12], srate * 0.45) / srate);
    c->c8 = 1.0 - exp(-two_pi * jmin(sl[13], srate * 0.45) / srate);
    c->mix = sl[14];
}

static inline double bq(const Bq* b, double* z, double x)
{
    const double y = b->b0 * x + z[0];
    z[0] = b->b1 * x - b->a1 * y + z[1];
    z[1] = b->b2 * x - b->a2 * y;
    return y;
}

static inline double c8_sample(Cascade8* c, double x)
{
    const double y1 = bq(&c->s[0], c->z[0], x);
    const double y2 = bq(&c->s[1], c->z[1], y1);
    const double y3 = bq(&c->s[2], c->z[2], y2);
    const double u3 = y2 + c->mix * y3;
    const double y4 = bq(&c->s[3], c->z[3], u3);
    const double u4 = u3 + c->mix * y4;
    const double y5 = bq(&c->s[4], c->z[4], u4);
    const double y6 = bq(&c->s[5], c->z[5], y5);
    c->lp7 = (1 - c->a7) * y6 + c->a7 * c->lp7;
    c->lp8 += c->c8 * (c->lp7 - c->lp8);
    return c->lp8;
}

/* exported for coefficient-level checks: coef[6*5 + 3] = b0,b1,b2,a1,a2 per section, a7, c8, mix */
ZO_API void zo_cascade8_coefs(const double* sliders15, double srate, double* coef33)
{
    Cascade8 c; memset(&c, 0, sizeof c);
    c8_slider(&c, sliders15, srate);
    memcpy(coef33, c.s, sizeof c.s);
    coef33[30] = c.a7; coef33[31] = c.c8; coef33[32] = c.mix;
}

/* ---- batch render, same contract as zref_render (ref_shim.c) ------------------------------ */
typedef struct Job {
    const ZoSliders* sl; double srate; const float* params; int n_param_blocks;
    const float* in; float* out; double* out_d; int n_streams, block; int64_t n_frames, frame_stride; int s0, s1;
} Job;

static void* c8_worker(void* arg)
{
    Job* j = (Job*)arg;
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    const int ns = j->sl->n;
    for (int s = j->s0; s < j->s1; ++s) {
        Cascade8 c; memset(&c, 0, sizeof c);
        double cur[64], last[64]; int last_valid = 0;
        const float* x = j->in + (size_t)s * j->frame_stride;
        float* y = j->out + (size_t)s * j->frame_stride;
        double* yd = j->out_d ? j->out_d + (size_t)s * j->frame_stride : NULL;
        int64_t pos = 0; int b = 0;
        /* prepareToPlay: push sliders, @init (zero state), @slider -- :3297-3318 */
        while (pos < j->n_frames || b == 0) {
            const int pb = j->n_param_blocks > 1 ? (b < j->n_param_blocks ? b : j->n_param_blocks - 1) : 0;
            const float* hv = j->params + ((size_t)pb * j->n_streams + s) * ns;
            int changed = 0;
            for (int k = 0; k < ns; ++k) {
                cur[k] = zo_quantise(j->sl, k, hv[k]);
                if (!last_valid || cur[k] != last[k]) { changed = 1; last[k] = cur[k]; }
            }
            last_valid = 1;
            if (changed) c8_slider(&c, cur, j->srate);
            if (pos >= j->n_frames) break;
            const int n = (int)((j->n_frames - pos) < j->block ? (j->n_frames - pos) : j->block);
            for (int i = 0; i < n; ++i) {
                const double v = c8_sample(&c, (double)x[pos + i]);
                if (yd) yd[pos + i] = v;
                y[pos + i] = (float)v;
            }
            pos += n; ++b;
        }
    }
    return NULL;
}

ZO_API int zo_cascade8_render(const ZoSliders* sl, double srate, const float* params, int n_param_blocks,
                              const float* in, float* out, double* out_d, int n_streams, int64_t n_frames,
                              int64_t frame_stride, int block, int n_threads)
{
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
        if (n_threads == 1) { c8_worker(j); return 0; }
        if (pthread_create(&th[t], NULL, c8_worker, j) != 0) return -1;
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
    return 0;
}
