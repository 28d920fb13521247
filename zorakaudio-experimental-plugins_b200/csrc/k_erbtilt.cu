// k_erbtilt.cu -- ERB Tilt (plugins/Spectral/ERBTilt/src/ERB Tilt.jsfx) on the B200: ONE WARP PER STREAM, time-parallel
// scans (erbtilt_core.h).  4096 stereo streams are 4096 warps; the translated stream-per-thread form of this leaf
// (csrc/gen/g_ERBTilt.cu, still built as "ERBTilt.translated" for cross-checks) had 128 warps' worth of threads for 592
// schedulers and ran at 0.24 % of HBM.
//   * CTA = 4 independent warps, no block-wide barrier, no shared memory: a lane reads and writes its own C consecutive
//     frames straight from / to global memory as 16-byte vectors (C = 8 in float: two LDG.128 per channel and tile whose
//     second halves hit L1; C = 4 in double: fully coalesced), the scans exchange chunk end states by shuffle;
//   * per host block: @slider + @block in double on the device (erb::block_prologue), then 32 x C frame tiles, then
//     32-frame tiles (C = 1) for what is left of the block, the last one partial -- any block length and call length;
//   * state ([slot][stream] doubles) persists across calls: 30 low-pass states, 64 follower / energy states, global_gain.
// The kernel is FP32/FP64-issue bound by construction (~660 thread-operations per frame): its roofline is the FMA pipe,
// the HBM fraction is reported beside it.
#include "erbtilt_args.h"
#include "scan_io.cuh"

namespace za {

namespace {

using namespace erb;

constexpr int kWarps = 4;

template <class T, int CM, bool TAP>
__global__ void __launch_bounds__(32 * kWarps, 4) erbtilt_kernel(const ErbArgs a)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * kWarps + warp;
    if (s >= a.n_streams) return;                        // warps are independent
    const size_t ns = (size_t)a.n_streams;
    const Tables<T, CM>& tbm = *reinterpret_cast<const Tables<T, CM>*>(a.tb_main);
    const Tables<T, 1>& tbt = *reinterpret_cast<const Tables<T, 1>*>(a.tb_tail);
    constexpr bool FAST = sizeof(T) == 4;                // float mode: MUFU-based division (criterion: -100 dBFS); fp64: IEEE

    auto ld = [&](int slot0, int n) { return lane < n ? (T)a.state[(size_t)(slot0 + lane) * ns + s] : T(0); };
    Carry<T> cy{ld(kErbLpL, kNBm1), ld(kErbLpR, kNBm1), ld(kErbEnvF, kNB), ld(kErbEnvS, kNB), ld(kErbPreE, kNB), ld(kErbPostE, kNB)};
    BlockVars bv{};
    bv.global_gain = a.state[(size_t)kErbGlobalGain * ns + s];
    Block<T> bk{};

    const float* in0 = a.in + (size_t)s * 2 * a.row_stride;
    float* out0 = a.out + (size_t)s * 2 * a.row_stride;
    double* tap0 = TAP ? a.tap + (size_t)s * 2 * a.row_stride : nullptr;
    const bool base_vec = (a.row_stride % 4) == 0 && ((uintptr_t)a.in & 15) == 0 && ((uintptr_t)a.out & 15) == 0;

    int64_t pos = 0;
    for (int blk = 0; pos < a.n_frames; ++blk) {
        const int nb = (int)min((int64_t)a.block_len, a.n_frames - pos);
        const int row = blk < a.n_rows ? blk : a.n_rows - 1;
        block_prologue<T>(a.bc, a.q + ((size_t)row * ns + s) * 4, nb == a.block_len ? a.alpha_full : a.alpha_last, cy, bv, bk);
        int f = 0;
        for (; f + 32 * CM <= nb; f += 32 * CM) {
            const int64_t g0 = pos + f + lane * CM;
            ScanIO<T, CM, TAP> io{in0 + g0, in0 + a.row_stride + g0, out0 + g0, out0 + a.row_stride + g0,
                                 TAP ? tap0 + g0 : nullptr, TAP ? tap0 + a.row_stride + g0 : nullptr, true, base_vec && ((pos + f) & 3) == 0};
            tile<T, CM, FAST>(tbm, cy, bk, 32, io);
        }
        for (; f < nb; f += 32) {
            const int m = min(32, nb - f);
            const int64_t g0 = pos + f + lane;
            ScanIO<T, 1, TAP> io{in0 + g0, in0 + a.row_stride + g0, out0 + g0, out0 + a.row_stride + g0,
                                TAP ? tap0 + g0 : nullptr, TAP ? tap0 + a.row_stride + g0 : nullptr, lane < m, false};
            tile<T, 1, FAST>(tbt, cy, bk, m, io);
        }
        pos += nb;
    }

    auto st = [&](int slot0, int n, T v) { if (lane < n) a.state[(size_t)(slot0 + lane) * ns + s] = (double)v; };
    st(kErbLpL, kNBm1, cy.lpL); st(kErbLpR, kNBm1, cy.lpR); st(kErbEnvF, kNB, cy.envF); st(kErbEnvS, kNB, cy.envS);
    st(kErbPreE, kNB, cy.preE); st(kErbPostE, kNB, cy.postE);
    if (lane == 0) {
        const double v[12] = {bv.global_gain, bv.tilt_db, bv.pivot_hz, bv.erb_pivot, bv.slope, bv.comp_mix, bv.guard, bv.Lpre, bv.Lpost,
                              bv.target, bv.blend, bv.alpha};
#pragma unroll
        for (int k = 0; k < 12; ++k) a.state[(size_t)(kErbGlobalGain + k) * ns + s] = v[k];
    }
    if (lane < kNB) a.state[(size_t)(kErbGBand + lane) * ns + s] = (double)bk.g;
}

}  // namespace

bool erbtilt_launch(const ErbArgs& a, int precision, cudaStream_t stream)
{
    if (a.n_streams <= 0 || a.n_frames <= 0) return true;
    const unsigned grid = (unsigned)((a.n_streams + kWarps - 1) / kWarps);
    if (precision == 0) erbtilt_kernel<float, kErbChunkF32, false><<<grid, 32 * kWarps, 0, stream>>>(a);
    else if (a.tap) erbtilt_kernel<double, kErbChunkF64, true><<<grid, 32 * kWarps, 0, stream>>>(a);
    else erbtilt_kernel<double, kErbChunkF64, false><<<grid, 32 * kWarps, 0, stream>>>(a);
    return cuda_ok(cudaGetLastError(), "erbtilt launch");
}

}  // namespace za
