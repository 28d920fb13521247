// scan_io.cuh -- global-memory tile I/O of the warp-per-stream scan kernels (k_erbtilt.cu, k_specstab.cu): a lane reads and
// writes its own C consecutive frames of both channels, as 16-byte vectors when the rows are aligned (C = 8 in float: two
// LDG.128 per channel whose second halves hit L1; C = 4 in double: fully coalesced), scalar otherwise; lanes beyond the
// valid part of a partial tile read zeros and store nothing.  in == out is legal: all loads of a tile precede its stores.
#pragma once
#include <cuda_runtime.h>

namespace za {

template <class T, int C, bool TAP>
struct ScanIO {
    const float* inL; const float* inR; float* outL; float* outR; double* tapL; double* tapR;   // at this lane's first frame
    bool live, vec;
    __device__ __forceinline__ void load(T (&l)[C], T (&r)[C]) const
    {
        if (!live) {
#pragma unroll
            for (int i = 0; i < C; ++i) { l[i] = T(0); r[i] = T(0); }
            return;
        }
        if (C % 4 == 0 && vec) {
#pragma unroll
            for (int j = 0; j < C / 4; ++j) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(inL) + j), b = __ldg(reinterpret_cast<const float4*>(inR) + j);
                l[4 * j] = (T)a.x; l[4 * j + 1] = (T)a.y; l[4 * j + 2] = (T)a.z; l[4 * j + 3] = (T)a.w;
                r[4 * j] = (T)b.x; r[4 * j + 1] = (T)b.y; r[4 * j + 2] = (T)b.z; r[4 * j + 3] = (T)b.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < C; ++i) { l[i] = (T)__ldg(inL + i); r[i] = (T)__ldg(inR + i); }
        }
    }
    __device__ __forceinline__ void store(const T (&l)[C], const T (&r)[C]) const
    {
        if (!live) return;
        if (TAP) {
#pragma unroll
            for (int i = 0; i < C; ++i) { tapL[i] = (double)l[i]; tapR[i] = (double)r[i]; }
        }
        if (C % 4 == 0 && vec) {
#pragma unroll
            for (int j = 0; j < C / 4; ++j) {
                reinterpret_cast<float4*>(outL)[j] = make_float4((float)l[4 * j], (float)l[4 * j + 1], (float)l[4 * j + 2], (float)l[4 * j + 3]);
                reinterpret_cast<float4*>(outR)[j] = make_float4((float)r[4 * j], (float)r[4 * j + 1], (float)r[4 * j + 2], (float)r[4 * j + 3]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < C; ++i) { outL[i] = (float)l[i]; outR[i] = (float)r[i]; }
        }
    }
};


}  // namespace za
