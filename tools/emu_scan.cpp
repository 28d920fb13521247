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
