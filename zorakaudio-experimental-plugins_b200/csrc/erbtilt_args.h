// erbtilt_args.h -- launch arguments of the ERB Tilt scan kernel (k_erbtilt.cu), shared with its host half (p_erbtilt.cu).
#pragma once
#include "za_internal.h"
#include "erbtilt_core.h"

namespace za {
struct ErbArgs {
    const float* in; float* out; double* tap;      // [stream][2][row_stride]
    int64_t row_stride, n_frames;
    int n_streams, block_len, n_rows;
    const double* q;                               // device [n_rows][n_streams][4] quantised slider values
    double* state;                                 // [kErbSlots][n_streams]
    const void* tb_main; const void* tb_tail;      // erb::Tables<T, chunk> / erb::Tables<T, 1> (device)
    erb::BlockConsts bc;
    double alpha_full, alpha_last;                 // exp(-samplesblock / (gain_tau * srate + 1e-6)) for a full / the short last block (host libm)
};
bool erbtilt_launch(const ErbArgs& a, int precision, cudaStream_t stream);
}  // namespace za
