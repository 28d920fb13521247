/*
 * oracle/ref_shim.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * JUCE-free runner around ONE object produced by the reference's own compiler
 * (/root/reference/dsp_jsfx_aot.py --opt 2).  It is compiled once per plugin by
 * oracle/build_ref.py together with that plugin's generated JSFXDSP.o/.h into
 * oracle/_ref/libref_<key>.so.  It restates only the thin host glue the object needs:
 *
 *   - jsfx_ensure_mem            src/JSFXJuceProcessor.cpp:1452-1491 (grow x1.5+64 from 1024, zero-fill)
 *   - prepareToPlay call order   src/JSFXJuceProcessor.cpp:3250-3318 (memset keeping mem, srate,
 *                                push sliders, jsfx_init, alias re-apply, jsfx_slider)
 *   - processBlock call order    src/JSFXJuceProcessor.cpp:3536-3562,3733 (push sliders, jsfx_slider on
 *                                change, jsfx_process_block) and the --correctness-check loop
 *                                3605-3658 (jsfx_block + per-sample jsfx_sample) used for the double tap
 *   - parameter quantiser        src/JSFXJuceProcessor.cpp:5556-5597
 *   - FTZ/DAZ                    juce::ScopedNoDenormals at :3437
 *
 * The DSP arithmetic itself is the reference's generated code, untouched.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>
#include <pmmintrin.h>

#ifndef ZREF_HEADER
#error "compile with -DZREF_HEADER=\"<generated header>\""
#endif
#include ZREF_HEADER

#define ZREF_API __attribute__((visibility("default")))

/* ---- runtime hooks called from generated code ------------------------------------------- */
void jsfx_ensure_mem(DSPJSFX_State* st, int64_t needed)
{
    if (!st) return;
    if (needed <= st->memN) return;
    int64_t newN = st->memN > 0 ? st->memN : 1024;
    while (newN < needed) newN = (newN * 3) / 2 + 64;
    size_t oldBytes = (size_t)st->memN * sizeof(double);
    size_t newBytes = (size_t)newN * sizeof(double);
    void* p = realloc(st->mem, newBytes);
    if (!p) return;
    st->mem = (double*)p;
    if (newBytes > oldBytes) memset((uint8_t*)p + oldBytes, 0, newBytes - oldBytes);
    st->memN = newN;
}
/* src/JSFXJuceProcessor.cpp:2448-2476 */
int jsfx_sliderchange(DSPJSFX_State* st, double sliderMask)
{
    const int64_t mask = (sliderMask > 0.0) ? (int64_t)llround(sliderMask) : 0;
    if (!st || mask <= 0) return 0;
    st->pendingSliderChangeMask |= mask;
    return 1;
}
int jsfx_slider_automate(DSPJSFX_State* st, double sliderMask, double endTouch)
{
    const int64_t mask = (sliderMask > 0.0) ? (int64_t)llround(sliderMask) : 0;
    if (!st || mask <= 0) return 0;
    st->pendingSliderChangeMask |= mask;
    if (endTouch != 0.0) st->pendingSliderAutomateEndMask |= mask; else st->pendingSliderAutomateMask |= mask;
    return 1;
}

/* ---- slider layout passed in by the caller (parsed from the .jsfx text by the test host) -- */
typedef struct ZrefSliders {
    int32_t n;            /* number of declared (numeric) sliders            */
    int32_t index0[64];   /* slider index (0-based) of column j              */
    int32_t alias[64];    /* vars[] index mirrored by column j, or -1        */
    float   vmin[64], vmax[64], vstep[64];
    int32_t is_choice[64];
} ZrefSliders;

typedef struct Zref {
    DSPJSFX_State st;
    ZrefSliders   sl;
    double        last[64];
    int           last_valid;
} Zref;

static void ftz_on(void)
{
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
}

/* src/JSFXJuceProcessor.cpp:5556-5597 */
static double quantise(const ZrefSliders* sl, int j, float host_value)
{
    double v = (double)host_value;
    const double mn = (double)sl->vmin[j], mx = (double)sl->vmax[j];
    if (sl->is_choice[j]) {
        const int64_t idx = (int64_t)llround(v);
        v = mn + (double)idx * (double)sl->vstep[j];
    }
    v = v < mn ? mn : (v > mx ? mx : v);
    if (!sl->is_choice[j]) {
        const double step = (double)(sl->vstep[j] > 0.0f ? sl->vstep[j] : 0.0f);
        if (step > 0.0) {
            const double q = (double)llround((v - mn) / step);
            v = mn + q * step;
            v = v < mn ? mn : (v > mx ? mx : v);
        }
    }
    return v;
}

/* src/JSFXJuceProcessor.cpp:9286-9357 (GUI preview / internal-pending branches are host-only) */
static int push_sliders(Zref* z, const float* host_values)
{
    int changed = 0;
    const int cap = (int)(sizeof(z->st.vars) / sizeof(z->st.vars[0]));
    for (int j = 0; j < z->sl.n; ++j) {
        const int i = z->sl.index0[j];
        const double v = quantise(&z->sl, j, host_values[j]);
        if (!z->last_valid || v != z->last[i]) { changed = 1; z->last[i] = v; }
        z->st.sliders[i] = v;
        const int a = z->sl.alias[j];
        if (a >= 0 && a < cap) z->st.vars[a] = v;
    }
    z->last_valid = 1;
    return changed;
}

ZREF_API int zref_in_channels(void)  { return DSPJSFX_INPUT_CHANNELS; }
ZREF_API int zref_out_channels(void) { return DSPJSFX_OUTPUT_CHANNELS; }
ZREF_API int zref_var_cap(void)      { return DSPJSFX_VARS_COUNT; }
ZREF_API int zref_var_index(const char* name)
{
    for (int i = 0; i < DSPJSFX_VARS_COUNT; ++i)
        if (DSPJSFX_VARS[i].name && strcmp(DSPJSFX_VARS[i].name, name) == 0) return DSPJSFX_VARS[i].index;
    return -1;
}

ZREF_API void* zref_new(const ZrefSliders* sl)
{
    Zref* z = (Zref*)calloc(1, sizeof(Zref));
    if (z && sl) z->sl = *sl;
    return z;
}
ZREF_API void zref_free(void* h)
{
    Zref* z = (Zref*)h;
    if (!z) return;
    free(z->st.mem);
    free(z);
}

/* prepareToPlay order, src/JSFXJuceProcessor.cpp:3250-3318 */
ZREF_API void zref_prepare(void* h, double srate, const float* host_values)
{
    Zref* z = (Zref*)h;
    ftz_on();
    double* memKeep = z->st.mem; int64_t memNKeep = z->st.memN;
    memset(&z->st, 0, sizeof(z->st));
    z->st.mem = memKeep; z->st.memN = memNKeep;
    z->st.srate = srate; z->st.currentSampleRate = srate;
    z->last_valid = 0;
    (void)push_sliders(z, host_values);
    jsfx_init(&z->st);
    const int cap = (int)(sizeof(z->st.vars) / sizeof(z->st.vars[0]));
    for (int j = 0; j < z->sl.n; ++j) {
        const int a = z->sl.alias[j];
        if (a >= 0 && a < cap) z->st.vars[a] = z->st.sliders[z->sl.index0[j]];
    }
    jsfx_slider(&z->st);
}

/* consumeDspSliderChanges (src/JSFXJuceProcessor.cpp:5665-5738) after every block: sliders the DSP changed itself
 * (sliderchange / slider_automate) become the host's view of them (lastSliders), the pending masks are cleared. */
static void consume_dsp_slider_changes(Zref* z)
{
    DSPJSFX_State* st = &z->st;
    const uint64_t m = (uint64_t)(st->pendingSliderChangeMask > 0 ? st->pendingSliderChangeMask : 0) |
                       (uint64_t)(st->pendingSliderAutomateMask > 0 ? st->pendingSliderAutomateMask : 0) |
                       (uint64_t)(st->pendingSliderAutomateEndMask > 0 ? st->pendingSliderAutomateEndMask : 0);
    for (int i = 0; i < 64; ++i)
        if (m & ((uint64_t)1 << i)) z->last[i] = st->sliders[i];
    st->pendingSliderChangeMask = 0; st->pendingSliderAutomateMask = 0; st->pendingSliderAutomateEndMask = 0;
}

/* processBlock, normal build: :3536-3562 then jsfx_process_block :3733.  Returns "sliders changed". */
ZREF_API int zref_process_block(void* h, const float* host_values, const float* const* in, float* const* out,
                                int nch, int n)
{
    Zref* z = (Zref*)h;
    ftz_on();
    int changed = 0;
    if (host_values) { changed = push_sliders(z, host_values); if (changed) jsfx_slider(&z->st); }
    z->st.currentBlockSize = n; z->st.samplesblock = (double)n;
    jsfx_process_block(&z->st, in, out, nch, n);
    consume_dsp_slider_changes(z);
    return changed;
}

/* processBlock, --correctness-check build order (:3605-3658), minus the shadow VM: same state
 * evolution, but st.spl[] is visible as double before the (float) store (dsp_jsfx_aot.py:5866-5892). */
ZREF_API int zref_process_block_f64(void* h, const float* host_values, const float* const* in,
                                    float* const* out, double* const* out_d, int nch, int n)
{
    Zref* z = (Zref*)h;
    ftz_on();
    int changed = 0;
    if (host_values) { changed = push_sliders(z, host_values); if (changed) jsfx_slider(&z->st); }
    DSPJSFX_State* st = &z->st;
    st->currentBlockSize = n; st->samplesblock = (double)n; st->currentSampleRate = st->srate;
    jsfx_block(st);
    if (st->pendingSliderChangeMask | st->pendingSliderAutomateMask | st->pendingSliderAutomateEndMask)
        jsfx_slider(st);                                   /* dsp_jsfx_aot.py:5791-5804 */
    if (nch > 64) nch = 64;
    for (int i = 0; i < n; ++i) {
        for (int c = 0; c < nch; ++c) st->spl[c] = (double)in[c][i];
        jsfx_sample(st);
        for (int c = 0; c < nch; ++c) {
            if (out_d) out_d[c][i] = st->spl[c];
            if (out)   out[c][i]   = (float)st->spl[c];
        }
    }
    consume_dsp_slider_changes(z);
    return changed;
}

ZREF_API double* zref_vars(void* h)   { return ((Zref*)h)->st.vars; }
ZREF_API double* zref_sliders(void* h){ return ((Zref*)h)->st.sliders; }
ZREF_API double* zref_mem(void* h, int64_t* n) { Zref* z = (Zref*)h; if (n) *n = z->st.memN; return z->st.mem; }
ZREF_API uint32_t* zref_rand_state(void* h, uint32_t* idx) { Zref* z = (Zref*)h; if (idx) *idx = z->st.randIndex; return z->st.randMT; }

/* ---- multithreaded batch render: one DSPJSFX_State per stream, static partition over threads
 *      (BASELINE.md section 3).  Layouts:  in/out [stream][ch][frame] float32 (frame_stride apart),
 *      params [n_param_blocks][n_streams][n_sliders] float32 (n_param_blocks==1 -> static).     */
typedef struct Job {
    const ZrefSliders* sl; double srate;
    const float* params; int n_param_blocks;
    const float* in; float* out; double* out_d;
    int n_streams, n_in, n_out, block; int64_t n_frames, frame_stride;
    int s0, s1;
    double* vars_out;      /* optional [n_streams][DSPJSFX_VARS_COUNT]: st.vars after the last frame */
} Job;

static void* worker(void* arg)
{
    Job* j = (Job*)arg;
    const int nch = j->n_in > j->n_out ? j->n_in : j->n_out;
    const int ns = j->sl->n;
    float* zero = (float*)calloc((size_t)j->block, sizeof(float));
    float* scratch = (float*)malloc((size_t)j->block * sizeof(float));
    for (int s = j->s0; s < j->s1; ++s) {
        Zref* z = (Zref*)zref_new(j->sl);
        zref_prepare(z, j->srate, j->params + (size_t)s * ns);
        const float* inp[64]; float* outp[64]; double* outd[64];
        int64_t pos = 0; int b = 0;
        while (pos < j->n_frames) {
            const int n = (int)((j->n_frames - pos) < j->block ? (j->n_frames - pos) : j->block);
            for (int c = 0; c < nch; ++c) {
                inp[c]  = c < j->n_in  ? j->in  + ((size_t)s * j->n_in  + c) * j->frame_stride + pos : zero;
                outp[c] = c < j->n_out ? j->out + ((size_t)s * j->n_out + c) * j->frame_stride + pos : scratch;
                outd[c] = (j->out_d && c < j->n_out) ? j->out_d + ((size_t)s * j->n_out + c) * j->frame_stride + pos : NULL;
            }
            const int pb = j->n_param_blocks > 1 ? (b < j->n_param_blocks ? b : j->n_param_blocks - 1) : 0;
            const float* hv = j->params + ((size_t)pb * j->n_streams + s) * ns;
            if (j->out_d) {
                /* tap path: per-channel NULL entries are allowed for scratch channels */
                double* dd[64]; for (int c = 0; c < nch; ++c) dd[c] = outd[c];
                Zref* zz = z; DSPJSFX_State* st = &zz->st;
                ftz_on();
                if (push_sliders(zz, hv)) jsfx_slider(st);
                st->currentBlockSize = n; st->samplesblock = (double)n; st->currentSampleRate = st->srate;
                jsfx_block(st);
                if (st->pendingSliderChangeMask | st->pendingSliderAutomateMask | st->pendingSliderAutomateEndMask)
                    jsfx_slider(st);
                for (int i = 0; i < n; ++i) {
                    for (int c = 0; c < nch; ++c) st->spl[c] = (double)inp[c][i];
                    jsfx_sample(st);
                    for (int c = 0; c < nch; ++c) {
                        if (dd[c]) dd[c][i] = st->spl[c];
                        outp[c][i] = (float)st->spl[c];
                    }
                }
                consume_dsp_slider_changes(zz);
            } else {
                zref_process_block(z, hv, inp, outp, nch, n);
            }
            pos += n; ++b;
        }
        if (j->vars_out) memcpy(j->vars_out + (size_t)s * DSPJSFX_VARS_COUNT, z->st.vars, sizeof(double) * DSPJSFX_VARS_COUNT);
        zref_free(z);
    }
    free(zero); free(scratch);
    return NULL;
}

ZREF_API int zref_render_vars(const ZrefSliders* sl, double srate, const float* params, int n_param_blocks,
                              const float* in, float* out, double* out_d, int n_streams, int64_t n_frames,
                              int64_t frame_stride, int block, int n_threads, double* vars_out);

ZREF_API int zref_render(const ZrefSliders* sl, double srate, const float* params, int n_param_blocks,
                         const float* in, float* out, double* out_d, int n_streams, int64_t n_frames,
                         int64_t frame_stride, int block, int n_threads)
{
    return zref_render_vars(sl, srate, params, n_param_blocks, in, out, out_d, n_streams, n_frames, frame_stride, block,
                            n_threads, NULL);
}

ZREF_API int zref_render_vars(const ZrefSliders* sl, double srate, const float* params, int n_param_blocks,
                              const float* in, float* out, double* out_d, int n_streams, int64_t n_frames,
                              int64_t frame_stride, int block, int n_threads, double* vars_out)
{
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_streams) n_threads = n_streams > 0 ? n_streams : 1;
    if (n_threads > 512) n_threads = 512;
    pthread_t th[512]; Job jobs[512];
    for (int t = 0; t < n_threads; ++t) {
        Job* j = &jobs[t];
        j->sl = sl; j->srate = srate; j->params = params; j->n_param_blocks = n_param_blocks;
        j->in = in; j->out = out; j->out_d = out_d; j->n_streams = n_streams;
        j->n_in = DSPJSFX_INPUT_CHANNELS; j->n_out = DSPJSFX_OUTPUT_CHANNELS;
        j->block = block; j->n_frames = n_frames; j->frame_stride = frame_stride; j->vars_out = vars_out;
        j->s0 = (int)((int64_t)n_streams * t / n_threads);
        j->s1 = (int)((int64_t)n_streams * (t + 1) / n_threads);
        if (n_threads == 1) { worker(j); return 0; }
        if (pthread_create(&th[t], NULL, worker, j) != 0) return -1;
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
    return 0;
}
