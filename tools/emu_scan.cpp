// tools/emu_scan.cpp -- host emulation of the time-parallel scan kernels: csrc/modtilt_scan_core.h compiled by g++ with
// lanes.h's 32-float "warp" type, i.e. the SAME tile source the CUDA kernel runs (k_modtilt_scan.cu), one warp = one
// stream, 256-frame tiles, tile-to-tile carries, per-host-block coefficient rows.  tests/test_faust_port.py checks it
// against the float restatement of the .dsp without a GPU (error budget of the re-associated chunk start states, block
// boundaries, state hand-over to the stream-per-thread tail path).  Test infrastructure only.
//   g++ -O2 -std=c++17 -ffp-contract=off -mfma -shared -fPIC -I<csrc> -I/usr/local/cuda/include tools/emu_scan.cpp -o libemu_scan.so
#include <type_traits>
#include <xmmintrin.h>
#include <pmmintrin.h>
#include <vector>
#include "modtilt_scan_core.h"

using namespace za;
using namespace za::mtscan;

namespace {
struct HostIO {
    const float* xl; const float* xr; float* yl; float* yr;       // this tile's 256 frames
    void load(LF (&l)[kC], LF (&r)[kC]) const
    {
        for (int i = 0; i < kC; ++i)
            for (int lane = 0; lane < 32; ++lane) { l[i].v[lane] = xl[lane * kC + i]; r[i].v[lane] = xr[lane * kC + i]; }
    }
    void store(const LF (&l)[kC], const LF (&r)[kC]) const
    {
        for (int i = 0; i < kC; ++i)
            for (int lane = 0; lane < 32; ++lane) { yl[lane * kC + i] = l[i].v[lane]; yr[lane * kC + i] = r[i].v[lane]; }
    }
};
}

// coef [n_blocks][8]: a_piv na_piv g_lo g_hi mix nmix - -; fixed3 = a_env, a_base, a_trim; state5 in/out.
// Renders n_frames (a multiple of block_len, block_len a multiple of 256) of one stream.
extern "C" __attribute__((visibility("default")))
void emu_modtilt_scan(const float* coef, int n_blocks, int block_len, const float* fixed3, float* state5, const float* in, float* out,
                      long n_frames, long stride, int fast)
{
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    Consts K{};
    K.env = make_stage(fixed3[0], 1.0f - fixed3[0]);
    K.base = make_stage(fixed3[1], 1.0f - fixed3[1]);
    K.ratio = make_stage(0.05f, 1.0f - 0.05f);
    K.trim = make_stage(fixed3[2], 1.0f - fixed3[2]);
    LF q_env, q_base, q_ratio, q_trim, q_piv;
    power_row(K.env.na, q_env.v); power_row(K.base.na, q_base.v); power_row(K.ratio.na, q_ratio.v); power_row(K.trim.na, q_trim.v);
    Carry cy{state5[0], state5[1], state5[2], state5[3], state5[4]};
    for (long t = 0; t * kTile < n_frames; ++t) {
        const long b = t * kTile / block_len;
        const float* c8 = coef + (size_t)(b < n_blocks ? b : n_blocks - 1) * 8;
        MTCoef c{c8[0], c8[1], c8[2], c8[3], c8[4], c8[5], 0.f, 0.f};
        power_row(c.na_piv, q_piv.v);
        HostIO io{in + t * kTile, in + stride + t * kTile, out + t * kTile, out + stride + t * kTile};
        if (fast) tile<true>(cy, K, q_env, q_base, q_ratio, q_trim, q_piv, c, io);
        else tile<false>(cy, K, q_env, q_base, q_ratio, q_trim, q_piv, c, io);
    }
    state5[0] = cy.env2; state5[1] = cy.base; state5[2] = cy.lp; state5[3] = cy.rs; state5[4] = cy.mean;
}

// ---- ERB Tilt: csrc/erbtilt_core.h on the host (same tile / prologue source as k_erbtilt.cu) ---------------------------------------
#include "erbtilt_args.h"
#include "erbtilt_tables.h"

namespace {
template <class T, int C>
struct ErbHostIO {
    const float* inL; const float* inR; float* outL; float* outR; double* tapL; double* tapR; int m;     // at the tile's first frame
    void load(za::lanes::LV<T> (&l)[C], za::lanes::LV<T> (&r)[C]) const
    {
        for (int i = 0; i < C; ++i)
            for (int lane = 0; lane < 32; ++lane) {
                const bool live = lane < m;
                l[i].v[lane] = live ? (T)inL[lane * C + i] : T(0); r[i].v[lane] = live ? (T)inR[lane * C + i] : T(0);
            }
    }
    void store(const za::lanes::LV<T> (&l)[C], const za::lanes::LV<T> (&r)[C]) const
    {
        for (int i = 0; i < C; ++i)
            for (int lane = 0; lane < m; ++lane) {
                outL[lane * C + i] = (float)l[i].v[lane]; outR[lane * C + i] = (float)r[i].v[lane];
                if (tapL) { tapL[lane * C + i] = (double)l[i].v[lane]; tapR[lane * C + i] = (double)r[i].v[lane]; }
            }
    }
};

template <class T, int CM>
void erb_render(double srate, const double* q, int n_rows, int block_len, double* state, const float* in, float* out, double* tap,
                long n_frames, long stride)
{
    using za::erb::Init; using za::erb::Tables; using za::erb::BlockVars; using za::erb::Block; using za::erb::kNB; using za::erb::kNBm1;
    using za::erb::build_filterbank; using za::erb::build_tables; using za::erb::block_prologue;
    const Init init = build_filterbank(srate);
    static Tables<T, CM> tbm; static Tables<T, 1> tbt;
    build_tables<T, CM>(init, tbm); build_tables<T, 1>(init, tbt);
    za::erb::Carry<T> cy{};
    auto ld = [&](za::lanes::LV<T>& v, int slot0, int n) { for (int l = 0; l < 32; ++l) v.v[l] = l < n ? (T)state[slot0 + l] : T(0); };
    ld(cy.lpL, za::kErbLpL, kNBm1); ld(cy.lpR, za::kErbLpR, kNBm1); ld(cy.envF, za::kErbEnvF, kNB); ld(cy.envS, za::kErbEnvS, kNB);
    ld(cy.preE, za::kErbPreE, kNB); ld(cy.postE, za::kErbPostE, kNB);
    BlockVars bv{}; bv.global_gain = state[za::kErbGlobalGain];
    Block<T> bk{};
    long pos = 0;
    for (int blk = 0; pos < n_frames; ++blk) {
        const int nb = (int)std::min<long>(block_len, n_frames - pos);
        const int row = blk < n_rows ? blk : n_rows - 1;
        const double alpha = std::exp(-(double)nb / (init.gain_tau * srate + 0.000001));
        block_prologue<T>(init.bc, q + (size_t)row * 4, alpha, cy, bv, bk);
        int f = 0;
        for (; f + 32 * CM <= nb; f += 32 * CM) {
            const long g0 = pos + f;
            ErbHostIO<T, CM> io{in + g0, in + stride + g0, out + g0, out + stride + g0, tap ? tap + g0 : nullptr, tap ? tap + stride + g0 : nullptr, 32};
            za::erb::tile<T, CM, false>(tbm, cy, bk, 32, io);
        }
        for (; f < nb; f += 32) {
            const int m = std::min(32, nb - f);
            const long g0 = pos + f;
            ErbHostIO<T, 1> io{in + g0, in + stride + g0, out + g0, out + stride + g0, tap ? tap + g0 : nullptr, tap ? tap + stride + g0 : nullptr, m};
            za::erb::tile<T, 1, false>(tbt, cy, bk, m, io);
        }
        pos += nb;
    }
    auto st = [&](const za::lanes::LV<T>& v, int slot0, int n) { for (int l = 0; l < n; ++l) state[slot0 + l] = (double)v.v[l]; };
    st(cy.lpL, za::kErbLpL, kNBm1); st(cy.lpR, za::kErbLpR, kNBm1); st(cy.envF, za::kErbEnvF, kNB); st(cy.envS, za::kErbEnvS, kNB);
    st(cy.preE, za::kErbPreE, kNB); st(cy.postE, za::kErbPostE, kNB);
    state[za::kErbGlobalGain] = bv.global_gain;
}
}

// q [n_rows][4] quantised sliders; state [kErbSlots] doubles (zero + global_gain = 1 after prepare); f64 = 1: double arithmetic.
extern "C" __attribute__((visibility("default")))
void emu_erbtilt(double srate, const double* q, int n_rows, int block_len, double* state, const float* in, float* out, double* tap,
                 long n_frames, long stride, int f64)
{
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    if (f64) erb_render<double, za::kErbChunkF64>(srate, q, n_rows, block_len, state, in, out, tap, n_frames, stride);
    else erb_render<float, za::kErbChunkF32>(srate, q, n_rows, block_len, state, in, out, tap, n_frames, stride);
}

// ---- Spectral Stabilizer: csrc/specstab_core.h on the host (same tile / prologue source as k_specstab.cu) ------------------------
#include "specstab_tables.h"

namespace {
template <class T, int CM>
void sstab_render(double srate, const double* q, int n_rows, int block_len, double* state /* [10][12] carries */, const float* in, float* out,
                  double* tap, long n_frames, long stride)
{
    namespace S = za::sstab;
    const S::Init init = S::build_init(srate);
    static S::Tables<T, CM> tbm; static S::Tables<T, 1> tbt;
    S::build_tables<T, CM>(init, tbm); S::build_tables<T, 1>(init, tbt);
    const std::vector<T> pwA = S::power_table<T>(init.k.atk_a, block_len), pwR = S::power_table<T>(init.k.rel_a, block_len);
    S::Carry<T> cy{};
    za::lanes::LV<T>* cv[10] = {&cy.z1L, &cy.z2L, &cy.z1R, &cy.z2R, &cy.env, &cy.elog, &cy.d1, &cy.d2, &cy.mot, &cy.gs};
    for (int v = 0; v < 10; ++v) for (int l = 0; l < 32; ++l) cv[v]->v[l] = l < S::kBands ? (T)state[v * S::kBands + l] : T(0);
    S::Prolog pv{}; S::Block<T> bk{};
    long pos = 0;
    for (int blk = 0; pos < n_frames; ++blk) {
        const int nb = (int)std::min<long>(block_len, n_frames - pos);
        const int row = blk < n_rows ? blk : n_rows - 1;
        S::block_prologue<T>(init.k, q + (size_t)row * 3, cy, pv, bk);
        int f = 0;
        for (; f + 32 * CM <= nb; f += 32 * CM) {
            const long g0 = pos + f;
            ErbHostIO<T, CM> io{in + g0, in + stride + g0, out + g0, out + stride + g0, tap ? tap + g0 : nullptr, tap ? tap + stride + g0 : nullptr, 32};
            S::tile<T, CM, false>(tbm, pwA.data(), pwR.data(), cy, bk, f, 32, io);
        }
        for (; f < nb; f += 32) {
            const int m = std::min(32, nb - f);
            const long g0 = pos + f;
            ErbHostIO<T, 1> io{in + g0, in + stride + g0, out + g0, out + stride + g0, tap ? tap + g0 : nullptr, tap ? tap + stride + g0 : nullptr, m};
            S::tile<T, 1, false>(tbt, pwA.data(), pwR.data(), cy, bk, f, m, io);
        }
        S::block_epilogue<T>(cy, bk, pwA[nb], pwR[nb]);
        pos += nb;
    }
    for (int v = 0; v < 10; ++v) for (int l = 0; l < S::kBands; ++l) state[v * S::kBands + l] = (double)cv[v]->v[l];
}
}

// q [n_rows][3] quantised sliders; state [10][12] doubles (z1L z2L z1R z2R env elog d1 d2 mot gs; gs = 1 after @init).
extern "C" __attribute__((visibility("default")))
void emu_specstab(double srate, const double* q, int n_rows, int block_len, double* state, const float* in, float* out, double* tap,
                  long n_frames, long stride, int f64)
{
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    if (f64) sstab_render<double, 4>(srate, q, n_rows, block_len, state, in, out, tap, n_frames, stride);
    else sstab_render<float, 8>(srate, q, n_rows, block_len, state, in, out, tap, n_frames, stride);
}
