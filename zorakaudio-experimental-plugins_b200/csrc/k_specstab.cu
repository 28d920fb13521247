// k_specstab.cu -- DoSG Spectral Stabilizer (plugins/Spectral/SpectralStabilizer/src/Spectral Stabilizer.jsfx) on the B200: ONE
// WARP PER STREAM, time-parallel scans (specstab_core.h): 24 two-state biquad scans, 12 energy scans, closed-form gain
// smoothing and an end-state-only motion EMA per tile; @slider / @block on the device with lane = band.  The translated
// stream-per-thread form of this leaf (csrc/gen/g_SpectralStabilizer.cu, kept as "SpectralStabilizer.translated") ran at
// 0.22 % of HBM and, living off the L1 for its mem[] heap, was the longest kernel of every mixed-catalog step.
// Same launch shape as k_erbtilt.cu: CTA = 4 independent warps, no shared memory, direct 16-byte global I/O, per host
// block 32 x C frame tiles then 32-frame (C = 1) tiles, the last one partial -- any block and call length.
#include "specstab_args.h"
#include "scan_io.cuh"

namespace za {

namespace {

using namespace sstab;
constexpr int kWarps = 4;

template <class T, int CM, bool TAP>
__global__ void __launch_bounds__(32 * kWarps, 4) specstab_kernel(const SsArgs a)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * kWarps + warp;
    if (s >= a.n_streams) return;
    const size_t ns = (size_t)a.n_streams;
    const Tables<T, CM>& tbm = *reinterpret_cast<const Tables<T, CM>*>(a.tb_main);
    const Tables<T, 1>& tbt = *reinterpret_cast<const Tables<T, 1>*>(a.tb_tail);
    const T* pwA = reinterpret_cast<const T*>(a.pwA);
    const T* pwR = reinterpret_cast<const T*>(a.pwR);
    constexpr bool FAST = sizeof(T) == 4;                // float mode: MUFU log (criterion: -100 dBFS); fp64: libdevice log

    auto ld = [&](int v) { return lane < kBands ? (T)a.state[(size_t)(kSsCarry + v * kBands + lane) * ns + s] : T(0); };
    Carry<T> cy{ld(0), ld(1), ld(2), ld(3), ld(4), ld(5), ld(6), ld(7), ld(8), ld(9)};
    Prolog pv{};
    Block<T> bk{};

    const float* in0 = a.in + (size_t)s * 2 * a.row_stride;
    float* out0 = a.out + (size_t)s * 2 * a.row_stride;
    double* tap0 = TAP ? a.tap + (size_t)s * 2 * a.row_stride : nullptr;
    const bool base_vec = (a.row_stride % 4) == 0 && ((uintptr_t)a.in & 15) == 0 && ((uintptr_t)a.out & 15) == 0;

    int64_t pos = 0;
    for (int blk = 0; pos < a.n_frames; ++blk) {
        const int nb = (int)min((int64_t)a.block_len, a.n_frames - pos);
        const int row = blk < a.n_rows ? blk : a.n_rows - 1;
        block_prologue<T>(a.k, a.q + ((size_t)row * ns + s) * 3, cy, pv, bk);
        int f = 0;
        for (; f + 32 * CM <= nb; f += 32 * CM) {
            const int64_t g0 = pos + f + lane * CM;
            ScanIO<T, CM, TAP> io{in0 + g0, in0 + a.row_stride + g0, out0 + g0, out0 + a.row_stride + g0,
                                  TAP ? tap0 + g0 : nullptr, TAP ? tap0 + a.row_stride + g0 : nullptr, true, base_vec && ((pos + f) & 3) == 0};
            tile<T, CM, FAST>(tbm, pwA, pwR, cy, bk, f, 32, io);
        }
        for (; f < nb; f += 32) {
            const int m = min(32, nb - f);
            const int64_t g0 = pos + f + lane;
            ScanIO<T, 1, TAP> io{in0 + g0, in0 + a.row_stride + g0, out0 + g0, out0 + a.row_stride + g0,
                                 TAP ? tap0 + g0 : nullptr, TAP ? tap0 + a.row_stride + g0 : nullptr, lane < m, false};
            tile<T, 1, FAST>(tbt, pwA, pwR, cy, bk, f, m, io);
        }
        block_epilogue<T>(cy, bk, pwA[nb], pwR[nb]);
        pos += nb;
    }

    auto st = [&](int v, T x) { if (lane < kBands) a.state[(size_t)(kSsCarry + v * kBands + lane) * ns + s] = (double)x; };
    st(0, cy.z1L); st(1, cy.z2L); st(2, cy.z1R); st(3, cy.z2R); st(4, cy.env); st(5, cy.elog); st(6, cy.d1); st(7, cy.d2); st(8, cy.mot); st(9, cy.gs);
    if (lane < kTaps) { a.state[(size_t)(kSsGn + lane) * ns + s] = pv.gn; a.state[(size_t)(kSsGw + lane) * ns + s] = pv.gw; }
    if (lane < kBands) { a.state[(size_t)(kSsRes + lane) * ns + s] = pv.res; a.state[(size_t)(kSsGt + lane) * ns + s] = pv.gt; }
    if (lane == 0) {
        a.state[(size_t)(kSsScalars + 0) * ns + s] = pv.depth; a.state[(size_t)(kSsScalars + 1) * ns + s] = pv.mb;
        a.state[(size_t)(kSsScalars + 2) * ns + s] = pv.low_min_gain; a.state[(size_t)(kSsScalars + 3) * ns + s] = pv.action_amt;
    }
}

}  // namespace

bool specstab_launch(const SsArgs& a, int precision, cudaStream_t stream)
{
    if (a.n_streams <= 0 || a.n_frames <= 0) return true;
    const unsigned grid = (unsigned)((a.n_streams + kWarps - 1) / kWarps);
    if (precision == 0) specstab_kernel<float, kSsChunkF32, false><<<grid, 32 * kWarps, 0, stream>>>(a);
    else if (a.tap) specstab_kernel<double, kSsChunkF64, true><<<grid, 32 * kWarps, 0, stream>>>(a);
    else specstab_kernel<double, kSsChunkF64, false><<<grid, 32 * kWarps, 0, stream>>>(a);
    return cuda_ok(cudaGetLastError(), "specstab launch");
}

}  // namespace za
