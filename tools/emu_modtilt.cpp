// tools/emu_modtilt.cpp -- host emulation of the ModTilt kernel's per-frame arithmetic (csrc/modtilt_core.h compiled by
// g++: same source text as the device code, libm in place of the MUFU approximations).  Used by tests/test_faust_port.py
// to check, without a GPU, that the kernel's expression order is the restatement's.  Test infrastructure only.
//   g++ -O2 -ffp-contract=off -mfma -shared -fPIC -I<csrc> -I/usr/local/cuda/include tools/emu_modtilt.cpp -o libemu_modtilt.so
#include <xmmintrin.h>
#include <pmmintrin.h>
#include "modtilt_core.h"

extern "C" __attribute__((visibility("default")))
void emu_modtilt_render(const float* coef8 /* [ns][8]: a_env a_base a_trim a_piv g_hi g_lo mix fs */, const float* in, float* out,
                        int n_streams, long n_frames, int use_run16)
{
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
    for (int s = 0; s < n_streams; ++s) {
        const float* c8 = coef8 + (size_t)s * 8;
        const za::MTK k{c8[0], 1.0f - c8[0], c8[1], 1.0f - c8[1], c8[2], 1.0f - c8[2]};
        za::MTCoef c{c8[3], 1.0f - c8[3], c8[5], c8[4], c8[6], 1.0f - c8[6], 0.f, 0.f};
        za::MTRegs r{0.f, 0.f, 0.f, 0.f, 0.f};
        const float* xl = in + (size_t)s * 2 * n_frames; const float* xr = xl + n_frames;
        float* yl = out + (size_t)s * 2 * n_frames; float* yr = yl + n_frames;
        long i = 0;
        if (use_run16)
            for (; i + 16 <= n_frames; i += 16) {
                for (int j = 0; j < 16; ++j) { yl[i + j] = xl[i + j]; yr[i + j] = xr[i + j]; }
                za::mt_run16(r, k, c, yl + i, yr + i);
            }
        for (; i < n_frames; ++i) za::mt_frame(r, k, c, xl[i], xr[i], yl[i], yr[i]);
    }
}
