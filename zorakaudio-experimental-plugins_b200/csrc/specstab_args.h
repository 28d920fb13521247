// specstab_args.h -- launch arguments of the Spectral Stabilizer scan kernel (k_specstab.cu), shared with its host half.
#pragma once
#include "za_internal.h"
#include "specstab_core.h"

namespace za {
struct SsArgs {
    const float* in; float* out; double* tap;      // [stream][2][row_stride]
    int64_t row_stride, n_frames;
    int n_streams, block_len, n_rows;
    const double* q;                               // device [n_rows][n_streams][3] quantised slider values
    double* state;                                 // [kSsSlots][n_streams]
    const void* tb_main; const void* tb_tail;      // sstab::Tables<T, chunk> / <T, 1> (device)
    const void* pwA; const void* pwR;              // atk_a^n / rel_a^n, n = 0 .. block_len (+ slack), T
    sstab::Consts k;
};
bool specstab_launch(const SsArgs& a, int precision, cudaStream_t stream);
}  // namespace za
