// lanes.h -- "one warp, written once": the value type of warp-cooperative scan code.
//
// The time-parallel kernels of the linear-recurrence leaves (k_modtilt_scan.cu, k_erbtilt.cu, ...) give ONE stream to a
// whole warp: lane l owns C consecutive frames of a 32 x C frame tile and the lanes exchange chunk end states with
// shuffles.  To check that logic (tiles, carries, host-block boundaries, error budget) on a machine without a GPU, the
// per-tile code is written against the small vocabulary below: on the device `LF` is a float and the functions are the
// rounding-pinned intrinsics; on the host `LF` is 32 floats -- one warp -- and every function is the same operation
// applied lane by lane, `shfl_up` a shift of that array.  The SAME source text then runs in the kernel and in the
// CPU tests (tools/emu_scan.cpp), like modtilt_core.h / stft_core.h do for their kernels.
#pragma once

#ifdef __CUDACC__
#include <cuda_runtime.h>
namespace za {
namespace lanes {

using LF = float;
using LB = bool;
#define ZA_LN __device__ __forceinline__
ZA_LN LF mul(LF a, LF b) { return __fmul_rn(a, b); }
ZA_LN LF add(LF a, LF b) { return __fadd_rn(a, b); }
ZA_LN LF sub(LF a, LF b) { return __fsub_rn(a, b); }
ZA_LN LF fma_rn(LF a, LF b, LF c) { return __fmaf_rn(a, b, c); }
ZA_LN LF max_(LF a, LF b) { return fmaxf(a, b); }
ZA_LN LF min_(LF a, LF b) { return fminf(a, b); }
ZA_LN LF abs_(LF a) { return fabsf(a); }
ZA_LN LF neg(LF a) { return -a; }
ZA_LN LF rcp_approx(LF b) { float r; asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(b)); return r; }
ZA_LN LF rsqrt_approx(LF x) { float y; asm("rsqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
ZA_LN LF lg2_approx(LF x) { float y; asm("lg2.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }     // MUFU.LG2, abs err <= 2^-22 near 1
ZA_LN LF ex2_approx(LF x) { float y; asm("ex2.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }     // MUFU.EX2, 2 ulp
ZA_LN LF log10_(LF x) { return log10f(x); }
ZA_LN LF exp10_(LF x) { return exp10f(x); }
ZA_LN LF log_(LF x) { return logf(x); }
ZA_LN LF exp_(LF x) { return expf(x); }
ZA_LN LF pow_(LF x, LF y) { return powf(x, y); }
ZA_LN LB gt(LF a, LF b) { return a > b; }
ZA_LN LB ge(LF a, LF b) { return a >= b; }
ZA_LN LF select(LB m, LF a, LF b) { return m ? a : b; }
ZA_LN LF splat(float v) { return v; }
ZA_LN LF shfl_up(LF v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }          // lanes < d keep their own value
ZA_LN LF shfl_idx(LF v, int l) { return __shfl_sync(0xffffffffu, v, l); }
ZA_LN LB lane_ge(int d) { return (int)(threadIdx.x & 31) >= d; }
ZA_LN float lane_value(LF v, int l) { return __shfl_sync(0xffffffffu, v, l); }      // a uniform scalar out of lane l

}  // namespace lanes
}  // namespace za
#else
#include <cmath>
#include <type_traits>
namespace za {
namespace lanes {

struct LF { float v[32]; };
struct LB { bool v[32]; };
#define ZA_LN inline
#define ZA_LANES_OP2(name, expr)                                                                          \
    ZA_LN LF name(const LF& a, const LF& b) { LF r; for (int l = 0; l < 32; ++l) { const float x = a.v[l], y = b.v[l]; r.v[l] = (expr); } return r; } \
    ZA_LN LF name(const LF& a, float y) { LF r; for (int l = 0; l < 32; ++l) { const float x = a.v[l]; r.v[l] = (expr); } return r; }                \
    ZA_LN LF name(float x, const LF& b) { LF r; for (int l = 0; l < 32; ++l) { const float y = b.v[l]; r.v[l] = (expr); } return r; }
#define ZA_LANES_OP1(name, expr) \
    ZA_LN LF name(const LF& a) { LF r; for (int l = 0; l < 32; ++l) { const float x = a.v[l]; r.v[l] = (expr); } return r; }
ZA_LANES_OP2(mul, x * y)
ZA_LANES_OP2(add, x + y)
ZA_LANES_OP2(sub, x - y)
ZA_LANES_OP2(max_, std::fmax(x, y))
ZA_LANES_OP2(min_, std::fmin(x, y))
ZA_LANES_OP2(pow_, std::pow(x, y))
ZA_LANES_OP1(abs_, std::fabs(x))
ZA_LANES_OP1(neg, -x)
ZA_LANES_OP1(rcp_approx, 1.0f / x)
ZA_LANES_OP1(rsqrt_approx, (float)(1.0 / std::sqrt((double)x)))
ZA_LANES_OP1(lg2_approx, std::log2(x))
ZA_LANES_OP1(ex2_approx, std::exp2(x))
ZA_LANES_OP1(log10_, std::log10(x))
ZA_LANES_OP1(exp10_, std::pow(10.0f, x))
ZA_LANES_OP1(log_, std::log(x))
ZA_LANES_OP1(exp_, std::exp(x))
#undef ZA_LANES_OP1
#undef ZA_LANES_OP2
template <class A, class B, class Cc> ZA_LN LF fma_rn(const A& a, const B& b, const Cc& c)
{
    auto at = [](const auto& q, int l) -> float { if constexpr (std::is_same_v<std::decay_t<decltype(q)>, LF>) return q.v[l]; else return (float)q; };
    LF r;
    for (int l = 0; l < 32; ++l) r.v[l] = std::fmaf(at(a, l), at(b, l), at(c, l));
    return r;
}
ZA_LN LB gt(const LF& a, float b) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = a.v[l] > b; return r; }
ZA_LN LB gt(const LF& a, const LF& b) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = a.v[l] > b.v[l]; return r; }
ZA_LN LB ge(const LF& a, const LF& b) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = a.v[l] >= b.v[l]; return r; }
ZA_LN LF select(const LB& m, const LF& a, const LF& b) { LF r; for (int l = 0; l < 32; ++l) r.v[l] = m.v[l] ? a.v[l] : b.v[l]; return r; }
ZA_LN LF select(const LB& m, const LF& a, float b) { LF r; for (int l = 0; l < 32; ++l) r.v[l] = m.v[l] ? a.v[l] : b; return r; }
ZA_LN LF splat(float v) { LF r; for (int l = 0; l < 32; ++l) r.v[l] = v; return r; }
ZA_LN LF shfl_up(const LF& a, int d) { LF r; for (int l = 0; l < 32; ++l) r.v[l] = l >= d ? a.v[l - d] : a.v[l]; return r; }
ZA_LN LF shfl_idx(const LF& a, int i) { return splat(a.v[i]); }
ZA_LN LB lane_ge(int d) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = l >= d; return r; }
ZA_LN float lane_value(const LF& a, int l) { return a.v[l]; }

}  // namespace lanes
}  // namespace za
#endif
