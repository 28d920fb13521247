// lanes.h -- "one warp, written once": the value types of warp-cooperative scan code.
//
// The time-parallel kernels of the linear-recurrence leaves (k_modtilt_scan.cu, k_erbtilt.cu, ...) give ONE stream to a
// whole warp: lane l owns C consecutive frames of a 32 x C frame tile and the lanes exchange chunk end states with
// shuffles.  To check that logic (tiles, carries, host-block boundaries, error budget) on a machine without a GPU, the
// per-tile code is written against the small vocabulary below: on the device a lane value `LV<T>` is a plain T (float or
// double) and the functions are the rounding-pinned intrinsics; on the host `LV<T>` is 32 values -- one warp -- and every
// function is the same operation applied lane by lane, `shfl_up` a shift of that array.  The SAME source text then runs
// in the kernel and in the CPU tests (tools/emu_scan.cpp), like modtilt_core.h / stft_core.h do for their kernels.
#pragma once

#ifdef __CUDACC__
#include <cuda_runtime.h>
namespace za {
namespace lanes {

template <class T> using LV = T;
using LF = float;
using LD = double;
using LB = bool;
#define ZA_LN __device__ __forceinline__
// float: plain operators, so nvcc contracts a * b + c into one FFMA (the float mode only has to null below -100 dBFS; FADD +
// FMUL pairs were a quarter of the scan kernels' issue slots).  double: the _rn intrinsics, never contracted (1e-12 parity).
ZA_LN float mul(float a, float b) { return a * b; }
ZA_LN float add(float a, float b) { return a + b; }
ZA_LN float sub(float a, float b) { return a - b; }
ZA_LN float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }
ZA_LN float div_(float a, float b) { return __fdiv_rn(a, b); }
ZA_LN float div_fast_(float a, float b) { return __fdividef(a, b); }
ZA_LN float sqrt_(float a) { return __fsqrt_rn(a); }
ZA_LN double mul(double a, double b) { return __dmul_rn(a, b); }
ZA_LN double add(double a, double b) { return __dadd_rn(a, b); }
ZA_LN double sub(double a, double b) { return __dsub_rn(a, b); }
ZA_LN double fma_rn(double a, double b, double c) { return __fma_rn(a, b, c); }
ZA_LN double div_(double a, double b) { return __ddiv_rn(a, b); }
ZA_LN double div_fast_(double a, double b) { return __ddiv_rn(a, b); }
ZA_LN double sqrt_(double a) { return __dsqrt_rn(a); }
ZA_LN float max_(float a, float b) { return fmaxf(a, b); }
ZA_LN float min_(float a, float b) { return fminf(a, b); }
ZA_LN float abs_(float a) { return fabsf(a); }
ZA_LN double max_(double a, double b) { return fmax(a, b); }
ZA_LN double min_(double a, double b) { return fmin(a, b); }
ZA_LN double abs_(double a) { return fabs(a); }
ZA_LN float neg(float a) { return -a; }
ZA_LN double neg(double a) { return -a; }
ZA_LN LF rcp_approx(LF b) { float r; asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(b)); return r; }
ZA_LN LF rsqrt_approx(LF x) { float y; asm("rsqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
ZA_LN LF lg2_approx(LF x) { float y; asm("lg2.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }     // MUFU.LG2, abs err <= 2^-22 near 1
ZA_LN LF ex2_approx(LF x) { float y; asm("ex2.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }     // MUFU.EX2, 2 ulp
ZA_LN LF log10_(LF x) { return log10f(x); }
ZA_LN LF exp10_(LF x) { return exp10f(x); }
ZA_LN float log_(float x) { return logf(x); }
ZA_LN float exp_(float x) { return expf(x); }
ZA_LN double log_(double x) { return log(x); }
ZA_LN double exp_(double x) { return exp(x); }
ZA_LN LF pow_(LF x, LF y) { return powf(x, y); }
ZA_LN bool gt(float a, float b) { return a > b; }
ZA_LN bool ge(float a, float b) { return a >= b; }
ZA_LN bool lt(float a, float b) { return a < b; }
ZA_LN bool gt(double a, double b) { return a > b; }
ZA_LN bool ge(double a, double b) { return a >= b; }
ZA_LN bool lt(double a, double b) { return a < b; }
ZA_LN float select(bool m, float a, float b) { return m ? a : b; }
ZA_LN double select(bool m, double a, double b) { return m ? a : b; }
ZA_LN LF splat(float v) { return v; }
ZA_LN float shfl_up(float v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }          // lanes < d keep their own value
ZA_LN double shfl_up(double v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
ZA_LN float shfl_down(float v, int d) { return __shfl_down_sync(0xffffffffu, v, d); }      // lanes >= 32 - d keep their own value
ZA_LN double shfl_down(double v, int d) { return __shfl_down_sync(0xffffffffu, v, d); }
ZA_LN LF shfl_idx(LF v, int l) { return __shfl_sync(0xffffffffu, v, l); }
ZA_LN LB lane_ge(int d) { return (int)(threadIdx.x & 31) >= d; }
ZA_LN LB lane_lt(int d) { return (int)(threadIdx.x & 31) < d; }
ZA_LN LB lane_eq(int d) { return (int)(threadIdx.x & 31) == d; }
ZA_LN float lane_value(float v, int l) { return __shfl_sync(0xffffffffu, v, l); }          // a uniform scalar out of lane l
ZA_LN double lane_value(double v, int l) { return __shfl_sync(0xffffffffu, v, l); }
ZA_LN void lane_set(float& v, int l, float x) { if ((int)(threadIdx.x & 31) == l) v = x; }  // write lane l of a lane-distributed table
ZA_LN void lane_set(double& v, int l, double x) { if ((int)(threadIdx.x & 31) == l) v = x; }
template <class T> ZA_LN T lane_index_as() { return (T)(threadIdx.x & 31); }
template <class T> ZA_LN T lane_zero() { return T(0); }
template <class T> ZA_LN T lane_load(const T* table) { return table[threadIdx.x & 31]; }                 // table[lane]
template <class T> ZA_LN T lane_load_rev(const T* table, int m)                                          // lane < m ? table[m - 1 - lane] : 0
{
    const int l = (int)(threadIdx.x & 31);
    return l < m ? table[m - 1 - l] : T(0);
}
template <class T, class F> ZA_LN T lane_generate(F f) { return f((int)(threadIdx.x & 31)); }              // f(lane)
template <class F> ZA_LN float lane_gather(float v, F idx) { return __shfl_sync(0xffffffffu, v, idx((int)(threadIdx.x & 31))); }    // v[idx(lane)]
template <class F> ZA_LN double lane_gather(double v, F idx) { return __shfl_sync(0xffffffffu, v, idx((int)(threadIdx.x & 31))); }
ZA_LN bool lane_any(bool p) { return __any_sync(0xffffffffu, p); }
template <class To, class From> ZA_LN To lane_cast(From v) { return (To)v; }

}  // namespace lanes
}  // namespace za
#else
#include <cmath>
#include <type_traits>
namespace za {
namespace lanes {

template <class T> struct LV { T v[32]; };
using LF = LV<float>;
using LD = LV<double>;
struct LB { bool v[32]; };
#define ZA_LN inline
template <class X> struct is_lv : std::false_type {};
template <class T> struct is_lv<LV<T>> : std::true_type {};
template <class T, class Q> ZA_LN T lane_at(const Q& q, int l) { if constexpr (is_lv<Q>::value) return q.v[l]; else return (T)q; }
#define ZA_LANES_OP2(name, expr)                                                                                                         \
    template <class T> ZA_LN LV<T> name(const LV<T>& a, const LV<T>& b) { LV<T> r; for (int l = 0; l < 32; ++l) { const T x = a.v[l], y = b.v[l]; r.v[l] = (expr); } return r; } \
    template <class T, class S, class = std::enable_if_t<std::is_arithmetic_v<S>>> ZA_LN LV<T> name(const LV<T>& a, S yy) { LV<T> r; const T y = (T)yy; for (int l = 0; l < 32; ++l) { const T x = a.v[l]; r.v[l] = (expr); } return r; } \
    template <class T, class S, class = std::enable_if_t<std::is_arithmetic_v<S>>> ZA_LN LV<T> name(S xx, const LV<T>& b) { LV<T> r; const T x = (T)xx; for (int l = 0; l < 32; ++l) { const T y = b.v[l]; r.v[l] = (expr); } return r; }
#define ZA_LANES_OP1(name, expr) \
    template <class T> ZA_LN LV<T> name(const LV<T>& a) { LV<T> r; for (int l = 0; l < 32; ++l) { const T x = a.v[l]; r.v[l] = (expr); } return r; }
ZA_LANES_OP2(mul, x * y)
ZA_LANES_OP2(add, x + y)
ZA_LANES_OP2(sub, x - y)
ZA_LANES_OP2(div_, x / y)
ZA_LANES_OP2(div_fast_, x / y)
ZA_LANES_OP2(max_, std::fmax(x, y))
ZA_LANES_OP2(min_, std::fmin(x, y))
ZA_LANES_OP2(pow_, std::pow(x, y))
ZA_LANES_OP1(abs_, std::fabs(x))
ZA_LANES_OP1(neg, -x)
ZA_LANES_OP1(sqrt_, std::sqrt(x))
ZA_LANES_OP1(rcp_approx, T(1) / x)
ZA_LANES_OP1(rsqrt_approx, (T)(1.0 / std::sqrt((double)x)))
ZA_LANES_OP1(lg2_approx, std::log2(x))
ZA_LANES_OP1(ex2_approx, std::exp2(x))
ZA_LANES_OP1(log10_, std::log10(x))
ZA_LANES_OP1(exp10_, std::pow(T(10), x))
ZA_LANES_OP1(log_, std::log(x))
ZA_LANES_OP1(exp_, std::exp(x))
#undef ZA_LANES_OP1
#undef ZA_LANES_OP2
// scalar (uniform) forms of the same vocabulary, so that code written for the device's plain-T lane values also compiles
// where a quantity is uniform on the host
ZA_LN float mul(float a, float b) { return a * b; }
ZA_LN float add(float a, float b) { return a + b; }
ZA_LN float sub(float a, float b) { return a - b; }
ZA_LN double mul(double a, double b) { return a * b; }
ZA_LN double add(double a, double b) { return a + b; }
ZA_LN double sub(double a, double b) { return a - b; }
ZA_LN float div_(float a, float b) { return a / b; }
ZA_LN double div_(double a, double b) { return a / b; }
ZA_LN float sqrt_(float a) { return std::sqrt(a); }
ZA_LN double sqrt_(double a) { return std::sqrt(a); }
ZA_LN float exp_(float a) { return std::exp(a); }
ZA_LN double exp_(double a) { return std::exp(a); }
ZA_LN float log_(float a) { return std::log(a); }
ZA_LN double log_(double a) { return std::log(a); }
template <class A, class B, class Cc> struct lv_of { using type = std::conditional_t<is_lv<A>::value, A, std::conditional_t<is_lv<B>::value, B, Cc>>; };
template <class A, class B, class Cc, class = std::enable_if_t<is_lv<A>::value || is_lv<B>::value || is_lv<Cc>::value>>
ZA_LN typename lv_of<A, B, Cc>::type fma_rn(const A& a, const B& b, const Cc& c)
{
    using V = typename lv_of<A, B, Cc>::type;
    using T = std::decay_t<decltype(V{}.v[0])>;
    V r;
    for (int l = 0; l < 32; ++l) r.v[l] = std::fma(lane_at<T>(a, l), lane_at<T>(b, l), lane_at<T>(c, l));
    return r;
}
template <class To, class From> ZA_LN LV<To> lane_cast(const LV<From>& a) { LV<To> r; for (int l = 0; l < 32; ++l) r.v[l] = (To)a.v[l]; return r; }
ZA_LN float neg(float a) { return -a; }
ZA_LN double neg(double a) { return -a; }
ZA_LN float abs_(float a) { return std::fabs(a); }
ZA_LN double abs_(double a) { return std::fabs(a); }
ZA_LN float fma_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
ZA_LN double fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
#define ZA_LANES_CMP(name, op)                                                                                                             \
    template <class T> ZA_LN LB name(const LV<T>& a, const LV<T>& b) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = a.v[l] op b.v[l]; return r; } \
    template <class T, class S, class = std::enable_if_t<std::is_arithmetic_v<S>>> ZA_LN LB name(const LV<T>& a, S b) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = a.v[l] op (T)b; return r; }
ZA_LANES_CMP(gt, >)
ZA_LANES_CMP(ge, >=)
ZA_LANES_CMP(lt, <)
#undef ZA_LANES_CMP
template <class T> ZA_LN LV<T> select(const LB& m, const LV<T>& a, const LV<T>& b) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = m.v[l] ? a.v[l] : b.v[l]; return r; }
template <class T, class S, class = std::enable_if_t<std::is_arithmetic_v<S>>> ZA_LN LV<T> select(const LB& m, const LV<T>& a, S b) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = m.v[l] ? a.v[l] : (T)b; return r; }
template <class T, class S, class = std::enable_if_t<std::is_arithmetic_v<S>>> ZA_LN LV<T> select(const LB& m, S a, const LV<T>& b) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = m.v[l] ? (T)a : b.v[l]; return r; }
ZA_LN LF splat(float v) { LF r; for (int l = 0; l < 32; ++l) r.v[l] = v; return r; }
template <class T> ZA_LN LV<T> shfl_up(const LV<T>& a, int d) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = l >= d ? a.v[l - d] : a.v[l]; return r; }
template <class T> ZA_LN LV<T> shfl_down(const LV<T>& a, int d) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = l + d < 32 ? a.v[l + d] : a.v[l]; return r; }
ZA_LN LF shfl_idx(const LF& a, int i) { return splat(a.v[i]); }
ZA_LN LB lane_ge(int d) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = l >= d; return r; }
ZA_LN LB lane_lt(int d) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = l < d; return r; }
ZA_LN LB lane_eq(int d) { LB r; for (int l = 0; l < 32; ++l) r.v[l] = l == d; return r; }
template <class T> ZA_LN T lane_value(const LV<T>& a, int l) { return a.v[l]; }
template <class T> ZA_LN void lane_set(LV<T>& a, int l, T x) { a.v[l] = x; }
template <class V> ZA_LN V lane_index_as() { V r; for (int l = 0; l < 32; ++l) r.v[l] = l; return r; }
template <class T> ZA_LN LV<T> lane_zero() { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = T(0); return r; }
template <class T> ZA_LN LV<T> lane_load(const T* table) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = table[l]; return r; }
template <class T> ZA_LN LV<T> lane_load_rev(const T* table, int m) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = l < m ? table[m - 1 - l] : T(0); return r; }
template <class T, class F> ZA_LN LV<T> lane_generate(F f) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = f(l); return r; }
template <class T, class F> ZA_LN LV<T> lane_gather(const LV<T>& a, F idx) { LV<T> r; for (int l = 0; l < 32; ++l) r.v[l] = a.v[idx(l) & 31]; return r; }

}  // namespace lanes
}  // namespace za
#endif
