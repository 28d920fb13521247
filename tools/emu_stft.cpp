// emu_stft.cpp -- host emulation of k_stft.cu's thread decomposition (stft_core.h phases run for t = 0..127 in
// turn, phase boundaries = the kernel's barriers).  Development check of the index arithmetic against the
// reference build before GPU time is spent; not part of the product.
//   g++ -O2 -ffp-contract=off -shared -fPIC -o tools/libemu_stft.so tools/emu_stft.cpp
#include <cmath>
#include <cstring>
#include <vector>
#include "../zorakaudio-experimental-plugins_b200/csrc/stft_core.h"
using namespace za::stft;

template <typename R>
static void run(const float* in, float* out, double* tap, int64_t n, const double* cf)
{
    static Tables<R> tb;
    std::vector<double> win(kN), syn(kN);
    for (int i = 0; i < kN; ++i) win[i] = 0.5 - 0.5 * std::cos(6.283185307179586 * i / (kN - 1));
    for (int i = 0; i < kN; ++i) {
        double sum = 0; int idx = i;
        for (int j = 0; j < 4; ++j) { sum += win[idx] * win[idx]; idx += kHop; if (idx >= kN) idx -= kN; }
        syn[i] = (sum > 1e-12 ? win[i] / sum : win[i]) * (1.0 / kN);
    }
    for (int i = 0; i < kN; ++i) { tb.win[i] = (R)win[i]; tb.syn[i] = (R)syn[i]; }
    for (int q = 1; q < 16; ++q) {
        for (int t = 0; t < 128; ++t) { double a = 6.283185307179586 * (t * q) / 2048.0; tb.tw1[q - 1][t] = {(R)std::cos(a), (R)-std::sin(a)}; }
        for (int j = 0; j < 8; ++j) { double a = 6.283185307179586 * (j * q) / 128.0; tb.tw2[q - 1][j] = {(R)std::cos(a), (R)-std::sin(a)}; }
    }
    static StreamSmem<R> S;
    std::memset(&S, 0, sizeof S);
    std::vector<Regs<R>> g(kT);
    std::memset(g.data(), 0, sizeof(Regs<R>) * kT);
    Coef c{cf[0], cf[1], cf[2], cf[3], cf[4], cf[5]};
    const R mixv = (R)c.mixv;
    const float* inL = in; const float* inR = in + n;
    float* outL = out; float* outR = out + n;
    int64_t hop = 0;
    for (int64_t pos = 0; pos < n; pos += kHop, ++hop) {
        const int len = (int)std::min<int64_t>(kHop, n - pos);
        const int slot = (int)(hop & 3) * kHop;
        for (int i = 0; i < len; ++i) {
            const int o = slot + i;
            const R dryL = (R)S.xin[0][o], dryR = (R)S.xin[1][o];
            S.xin[0][o] = inL[pos + i]; S.xin[1][o] = inR[pos + i];
            const R rawL = S.acc[0][o], rawR = S.acc[1][o];
            S.acc[0][o] = 0; S.acc[1][o] = 0;
            const R o0 = dryL + mixv * (rawL - dryL), o1 = dryR + mixv * (rawR - dryR);
            outL[pos + i] = (float)o0; outR[pos + i] = (float)o1;
            if (tap) { tap[pos + i] = (double)o0; tap[n + pos + i] = (double)o1; }
        }
        if (len < kHop) break;
        for (int t = 0; t < kT; ++t) fwd1<R>(t, g[t], S, tb, hop);
        for (int t = 0; t < kT; ++t) fwd2<R>(t, g[t], S, tb);
        for (int t = 0; t < kT; ++t) fwd3<R>(t, g[t], S);
        for (int t = 0; t < kT; ++t) t0_permute<R>(t == 0, g[t]);
        for (int t = 0; t < kT; ++t) mags_u<R>(t, g[t], S);
        for (int t = 0; t < kT; ++t) gains_u<R>(t, g[t], S, c);
        for (int t = 0; t < kT; ++t) { apply_gains_u<R>(t, g[t], S); t0_unpermute<R>(t == 0, g[t]); }
        for (int t = 0; t < kT; ++t) inv3<R>(t, g[t], S);
        for (int t = 0; t < kT; ++t) inv2<R>(t, g[t], S, tb);
        for (int t = 0; t < kT; ++t) inv1_ola<R>(t, g[t], S, tb, hop);
    }
}

extern "C" void emu_stft(const float* in, float* out, double* tap, long long n, const double* coef6, int f64)
{
    if (f64) run<double>(in, out, tap, n, coef6); else run<float>(in, out, tap, n, coef6);
}
