/* oracle/ref_fft_hooks.c -- TEST INFRASTRUCTURE: the FFT runtime hooks the reference's generated DSP calls,
 * restated JUCE-free over the reference's OWN WDL FFT (src/WDL/fft.c, compiled where it lies through
 * -DZREF_WDL_FFT_C="..." with WDL_FFT_REALSIZE=8 as in cmake/plugin/CMakeLists.txt:224).
 *
 * Follows src/JSFXJuceProcessor.cpp:1074-1447 (strict mode, ZA_JSFX_FFT_LEGACY_IN_ORDER=0):
 *   size rule 16..32768 power of two (:1090-1091,1126-1129), region inside one 65536-double page
 *   (:1131-1158), index rounding v +- 1e-5 (:1121-1124), heap growth through jsfx_ensure_mem,
 *   fft/ifft on WDL's permuted order (:1267-1302), fft_real/ifft_real (:1304-1339),
 *   fft_permute / fft_ipermute through WDL_fft_permute_tab (:1232-1265, 1417-1447),
 *   convolve_c complex multiply (:1341-1383), memcpy = memmove on mem[] (:1385-1415).
 * Invalid arguments return 0.0 and do nothing, like the reference. */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include ZREF_HEADER
#include ZREF_WDL_FFT_C

void jsfx_ensure_mem(DSPJSFX_State* st, int64_t needed);

static int64_t round_index(double v) { return (int64_t)(v + (v >= 0.0 ? 1.0e-5 : -1.0e-5)); }
static int pow2(int64_t n) { return n > 0 && (n & (n - 1)) == 0; }
static int size_ok(int64_t n) { return n >= 16 && n <= 32768 && pow2(n); }
static int in_page(int64_t base, int64_t span)
{
    if (base < 0 || span <= 0) return 0;
    const int64_t last = base + span - 1;
    if (last < base) return 0;
    return (base / 65536) == (last / 65536);
}
static pthread_once_t g_wdl_once = PTHREAD_ONCE_INIT;
static void wdl_init_once(void) { pthread_once(&g_wdl_once, WDL_fft_init); }      /* std::call_once in the reference (:1150-1154) */

static int complex_region(DSPJSFX_State* st, double baseD, double countD, int64_t* base, int* count)
{
    if (!st) return 0;
    const int64_t n = round_index(countD);
    int64_t start = round_index(baseD);
    if (start < 0) start = 0;
    if (n <= 0 || n > 65536 / 2) return 0;
    if (!in_page(start, 2 * n)) return 0;
    const int64_t needed = start + 2 * n;
    jsfx_ensure_mem(st, needed);
    if (!st->mem || needed > st->memN) return 0;
    *base = start; *count = (int)n;
    return 1;
}
static int fft_region(DSPJSFX_State* st, double baseD, double sizeD, int64_t* base, int* N)
{
    const int64_t size = round_index(sizeD);
    if (!size_ok(size)) return 0;
    if (!complex_region(st, baseD, (double)size, base, N)) return 0;
    wdl_init_once();
    return 1;
}
static int real_region(DSPJSFX_State* st, double baseD, double sizeD, int64_t* base, int* N)
{
    if (!st) return 0;
    const int64_t size = round_index(sizeD);
    int64_t start = round_index(baseD);
    if (start < 0) start = 0;
    if (!size_ok(size)) return 0;
    if (!in_page(start, size)) return 0;
    const int64_t needed = start + size;
    jsfx_ensure_mem(st, needed);
    if (!st->mem || needed > st->memN) return 0;
    wdl_init_once();
    *base = start; *N = (int)size;
    return 1;
}

static __thread WDL_FFT_COMPLEX* g_scratch = NULL;
static __thread int g_scratch_cap = 0;
static WDL_FFT_COMPLEX* scratch(int n)
{
    if (n > g_scratch_cap) {
        void* p = realloc(g_scratch, (size_t)n * sizeof(WDL_FFT_COMPLEX));
        if (!p) return NULL;
        g_scratch = (WDL_FFT_COMPLEX*)p; g_scratch_cap = n;
    }
    return g_scratch;
}

double jsfx_fft(DSPJSFX_State* st, double b, double s)
{
    int64_t base; int N;
    if (!fft_region(st, b, s, &base, &N)) return 0.0;
    WDL_fft((WDL_FFT_COMPLEX*)(st->mem + base), N, 0);
    return 0.0;
}
double jsfx_ifft(DSPJSFX_State* st, double b, double s)
{
    int64_t base; int N;
    if (!fft_region(st, b, s, &base, &N)) return 0.0;
    WDL_fft((WDL_FFT_COMPLEX*)(st->mem + base), N, 1);
    return 0.0;
}
double jsfx_fft_real(DSPJSFX_State* st, double b, double s)
{
    int64_t base; int N;
    if (!real_region(st, b, s, &base, &N)) return 0.0;
    WDL_real_fft(st->mem + base, N, 0);
    return 0.0;
}
double jsfx_ifft_real(DSPJSFX_State* st, double b, double s)
{
    int64_t base; int N;
    if (!real_region(st, b, s, &base, &N)) return 0.0;
    WDL_real_fft(st->mem + base, N, 1);
    return 0.0;
}
double jsfx_fft_permute(DSPJSFX_State* st, double b, double s)
{
    int64_t base; int N;
    if (!fft_region(st, b, s, &base, &N)) return 0.0;
    WDL_FFT_COMPLEX* sc = scratch(N);
    WDL_FFT_COMPLEX* buf = (WDL_FFT_COMPLEX*)(st->mem + base);
    const int* perm = WDL_fft_permute_tab(N);
    if (!sc || !perm) return 0.0;
    for (int i = 0; i < N; ++i) sc[i] = buf[perm[i]];
    memcpy(buf, sc, (size_t)N * sizeof(WDL_FFT_COMPLEX));
    return 0.0;
}
double jsfx_fft_ipermute(DSPJSFX_State* st, double b, double s)
{
    int64_t base; int N;
    if (!fft_region(st, b, s, &base, &N)) return 0.0;
    WDL_FFT_COMPLEX* sc = scratch(N);
    WDL_FFT_COMPLEX* buf = (WDL_FFT_COMPLEX*)(st->mem + base);
    const int* perm = WDL_fft_permute_tab(N);
    if (!sc || !perm) return 0.0;
    for (int i = 0; i < N; ++i) sc[perm[i]] = buf[i];
    memcpy(buf, sc, (size_t)N * sizeof(WDL_FFT_COMPLEX));
    return 0.0;
}
double jsfx_convolve_c(DSPJSFX_State* st, double d, double s, double n)
{
    int64_t db, sb; int N, M;
    if (!complex_region(st, d, n, &db, &N)) return 0.0;
    if (!complex_region(st, s, n, &sb, &M)) return 0.0;
    if (M != N) return 0.0;
    WDL_FFT_COMPLEX* dest = (WDL_FFT_COMPLEX*)(st->mem + db);
    WDL_FFT_COMPLEX* src = (WDL_FFT_COMPLEX*)(st->mem + sb);
    const int overlaps = (db < sb + 2 * (int64_t)M) && (sb < db + 2 * (int64_t)N);
    if (overlaps && db != sb) {
        WDL_FFT_COMPLEX* sc = scratch(N);
        if (!sc) return 0.0;
        memcpy(sc, src, (size_t)N * sizeof(WDL_FFT_COMPLEX));
        src = sc;
    }
    for (int i = 0; i < N; ++i) {
        const double ar = dest[i].re, ai = dest[i].im, br = src[i].re, bi = src[i].im;
        dest[i].re = ar * br - ai * bi;
        dest[i].im = ar * bi + ai * br;
    }
    return 0.0;
}
double jsfx_memcpy(DSPJSFX_State* st, double d, double s, double l)
{
    if (!st) return 0.0;
    int64_t dest = round_index(d), src = round_index(s), len = round_index(l);
    if (dest < 0) dest = 0;
    if (src < 0) src = 0;
    if (len <= 0) return 0.0;
    const int64_t de = dest + len, se = src + len;
    if (de < dest || se < src) return 0.0;
    const int64_t needed = de > se ? de : se;
    jsfx_ensure_mem(st, needed);
    if (!st->mem || needed > st->memN) return 0.0;
    memmove(st->mem + dest, st->mem + src, (size_t)len * sizeof(double));
    return 0.0;
}
