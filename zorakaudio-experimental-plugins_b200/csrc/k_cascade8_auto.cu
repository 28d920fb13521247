// k_cascade8_auto.cu -- cascade8 with per-block automation: ONE WARP PER STREAM, time-parallel scans whose coefficient
// tables are rebuilt ON THE DEVICE whenever a host block brings new coefficients.
//
// The throughput kernels of this plugin (k_cascade8.cu, k_cascade8_x2.cu) read 5.6 - 6 KB of scan tables per stream that a
// table kernel builds from the @slider coefficients, so a render whose sliders move was cut into one launch (plus two table
// launches) per host block in which anything changed: ~5 600 launches per 60 s, slower than the CPU.  This kernel serves those
// renders in one launch per call: per stream and host block it takes the block's 36 coefficients (host @slider with the host
// libm, [block][stream][36] rows), and if they differ from the previous block's it recomputes, in double, what the scans need:
// for each of the six TDF-II biquads A = [[-a1, 1], [-a2, 0]], (A^C)^(2^k) for the Kogge-Stone steps and (A^C)^lane for the
// carry -- all read off the section's impulse response, which six lanes compute in double-double arithmetic (see
// impulse_response below) -- and the same powers of the two one-pole coefficients.  That costs about as much as rendering the
// block -- against a launch per block it is three orders of magnitude.
// The tiles are the ones the other scan kernels use (specstab_core.h biquad_tile, erbtilt_core.h lp_tile / ema_tile): zero-state
// pass, warp scan, pass B in the script's own operation order (cascade8.jsfx:83-101), so fp64 meets the 1e-12 criterion like the
// static kernels; 32 x 4 frame tiles and 32 x 1 tail tiles, any block and call length; double arithmetic for both precisions.
// State: the plugin's own rows ([stream][16] doubles: z11 .. z62, lp7, lp8), shared with the static kernels.
#include "za_internal.h"
#include "specstab_core.h"
#include "erbtilt_core.h"

namespace za {

namespace {

using namespace lanes;

constexpr int kWarps = 4;

struct C8AutoArgs {
    const float* in; float* out; double* tap;      // mono rows [stream][row_stride]
    int64_t row_stride, n_frames;
    int n_streams, block_len, n_rows;
    const double* coef;                            // [n_rows][n_streams][kC8Coefs]
    double* state;                                 // [n_streams][kC8State]
};

template <class T> struct SecTab { T c5[5]; T M[5][4]; T Q[32][4]; };
template <class T> struct PoleTab { T P[5]; T Q[32]; };
template <class T> struct WarpTab { SecTab<T> bq[6]; PoleTab<T> p7, p8; T a7, oma7, c8, mix; };

// ---- double-double helpers (error-free transformations; intrinsics, so nothing is contracted or reassociated) -------------
struct DD { double hi, lo; };
__device__ __forceinline__ DD fast_two_sum(double a, double b) { const double s = __dadd_rn(a, b); return DD{s, __dsub_rn(b, __dsub_rn(s, a))}; }
__device__ __forceinline__ DD two_sum(double a, double b)
{
    const double s = __dadd_rn(a, b), bb = __dsub_rn(s, a);
    return DD{s, __dadd_rn(__dsub_rn(a, __dsub_rn(s, bb)), __dsub_rn(b, bb))};
}
__device__ __forceinline__ DD dd_mul_d(DD x, double c)               // (hi, lo) * c
{
    const double p = __dmul_rn(x.hi, c), e = __fma_rn(x.hi, c, -p);
    return fast_two_sum(p, __fma_rn(x.lo, c, e));
}
__device__ __forceinline__ DD dd_add(DD x, DD y)
{
    const DD s = two_sum(x.hi, y.hi);
    return fast_two_sum(s.hi, __dadd_rn(s.lo, __dadd_rn(x.lo, y.lo)));
}

// Powers of a TDF-II section's state matrix A = [[-a1, 1], [-a2, 0]] are its impulse response in disguise:
//   A^n = [[h_n, h_{n-1}], [-a2 h_{n-1}, -a2 h_{n-2}]],  h_{-1} = 0, h_0 = 1, h_n = -a1 h_{n-1} - a2 h_{n-2}.
// For poles close to the unit circle (a 40 Hz high-pass at 48 kHz) the entries are O(n) numbers that come out of O(n^2)
// cancellations: computed in plain double (by squaring or by the recurrence) they carry ~1e-12 relative error, which the
// carry propagation hands straight to the output (the static kernels' host tables are made in long double for the same
// reason).  So the recurrence runs in double-double here when PRECISE (fp64 mode), one section per lane, 32 C + 1 steps, and
// every table entry -- for the C-frame tiles and for the 1-frame tail tiles alike -- is read off that one sequence.
template <bool PRECISE>
__device__ __forceinline__ void impulse_response(double a1, double a2, double* h /* h[n + 1] = h_n, n = -1 .. N */, int N)
{
    h[0] = 0.0; h[1] = 1.0;
    if (PRECISE) {
        DD p2{0.0, 0.0}, p1{1.0, 0.0};
        for (int n = 1; n <= N; ++n) {
            const DD t = dd_add(dd_mul_d(p1, -a1), dd_mul_d(p2, -a2));
            h[n + 1] = t.hi;
            p2 = p1; p1 = t;
        }
    } else {
        double p2 = 0.0, p1 = 1.0;
        for (int n = 1; n <= N; ++n) {
            const double t = __fma_rn(-a1, p1, __dmul_rn(-a2, p2));
            h[n + 1] = t;
            p2 = p1; p1 = t;
        }
    }
}

// tables of chunk length C for the coefficient row c, from the impulse responses hs[section][.] (all lanes call)
template <class T>
__device__ void build_tables(WarpTab<T>& wt, const double* c, const double* hs, int hstride, int C, int lane)
{
#pragma unroll 1
    for (int k = 0; k < 6; ++k) {
        const double* b = c + 5 * k;
        const double* h = hs + k * hstride + 1;                            // h[n] = h_n, h[-1] = 0
        const double na2 = -b[4];
        SecTab<T>& st = wt.bq[k];
        if (lane < 5) {                                                    // (A^C)^(2^lane) = A^(C << lane)
            const int n = C << lane;
            st.M[lane][0] = (T)h[n]; st.M[lane][1] = (T)h[n - 1]; st.M[lane][2] = (T)(na2 * h[n - 1]); st.M[lane][3] = (T)(na2 * h[n - 2]);
            st.c5[lane] = (T)b[lane];
        }
        const int n = C * lane;                                            // (A^C)^lane
        if (n == 0) { st.Q[0][0] = T(1); st.Q[0][1] = T(0); st.Q[0][2] = T(0); st.Q[0][3] = T(1); }
        else { st.Q[lane][0] = (T)h[n]; st.Q[lane][1] = (T)h[n - 1]; st.Q[lane][2] = (T)(na2 * h[n - 1]); st.Q[lane][3] = (T)(na2 * h[n - 2]); }
    }
    auto pole = [&](PoleTab<T>& pt, double a) {
        double p = a;
        for (int r = 1; r < C; r <<= 1) p *= p;
        double q = 1.0;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            if (lane == 0) pt.P[j] = (T)p;
            if ((lane >> j) & 1) q *= p;
            p *= p;
        }
        pt.Q[lane] = (T)q;
    };
    pole(wt.p7, c[30]);
    pole(wt.p8, 1.0 - c[32]);            // lp8 += c8 (x - lp8): pole 1 - c8
    if (lane == 0) { wt.a7 = (T)c[30]; wt.oma7 = (T)c[31]; wt.c8 = (T)c[32]; wt.mix = (T)c[33]; }
    __syncwarp();
}

// one tile of 32 x C frames of one stream (cascade8.jsfx:83-101); m = valid lanes
template <class T, int C, bool TAP>
__device__ __forceinline__ void tile8(const WarpTab<T>& wt, T (&z)[12], T& lp7, T& lp8, const float* in, float* out, double* tap, int m, int lane)
{
    T x[C], y[C], u[C];
    const bool live = lane < m;
    const bool vec = C % 4 == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    if (live) {
        if (vec) {
#pragma unroll
            for (int j = 0; j < C / 4; ++j) {
                const float4 v = __ldg(reinterpret_cast<const float4*>(in) + j);
                x[4 * j] = (T)v.x; x[4 * j + 1] = (T)v.y; x[4 * j + 2] = (T)v.z; x[4 * j + 3] = (T)v.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < C; ++i) x[i] = (T)__ldg(in + i);
        }
    } else {
#pragma unroll
        for (int i = 0; i < C; ++i) x[i] = T(0);
    }
    const T mix = wt.mix;
    sstab::biquad_tile<T, C>(x, y, wt.bq[0].c5, wt.bq[0].M, wt.bq[0].Q, z[0], z[1], m);        // S1
    sstab::biquad_tile<T, C>(y, x, wt.bq[1].c5, wt.bq[1].M, wt.bq[1].Q, z[2], z[3], m);        // S2: x = y2
    sstab::biquad_tile<T, C>(x, y, wt.bq[2].c5, wt.bq[2].M, wt.bq[2].Q, z[4], z[5], m);        // S3: y = y3
#pragma unroll
    for (int i = 0; i < C; ++i) u[i] = add(x[i], mul(mix, y[i]));                              // u3 = y2 + mix * y3
    sstab::biquad_tile<T, C>(u, y, wt.bq[3].c5, wt.bq[3].M, wt.bq[3].Q, z[6], z[7], m);        // S4: y = y4
#pragma unroll
    for (int i = 0; i < C; ++i) u[i] = add(u[i], mul(mix, y[i]));                              // u4 = u3 + mix * y4
    sstab::biquad_tile<T, C>(u, y, wt.bq[4].c5, wt.bq[4].M, wt.bq[4].Q, z[8], z[9], m);        // S5
    sstab::biquad_tile<T, C>(y, x, wt.bq[5].c5, wt.bq[5].M, wt.bq[5].Q, z[10], z[11], m);      // S6: x = y6
#pragma unroll
    for (int i = 0; i < C; ++i) x[i] = mul(wt.oma7, x[i]);                                     // lp7 = (1 - a7) * y6 + a7 * lp7
    lp7 = erb::lp_tile<T, C>(x, wt.a7, wt.p7.P, wt.p7.Q[lane], lp7, m);
    lp8 = erb::ema_tile<T, C>(x, y, wt.c8, wt.p8.P, wt.p8.Q[lane], lp8, m);                    // lp8 += c8 * (lp7 - lp8)
    if (live) {
        if (TAP) {
#pragma unroll
            for (int i = 0; i < C; ++i) tap[i] = (double)y[i];
        }
        if (vec) {
#pragma unroll
            for (int j = 0; j < C / 4; ++j)
                reinterpret_cast<float4*>(out)[j] = make_float4((float)y[4 * j], (float)y[4 * j + 1], (float)y[4 * j + 2], (float)y[4 * j + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < C; ++i) out[i] = (float)y[i];
        }
    }
}

template <class T, int CM, bool TAP>
__global__ void __launch_bounds__(32 * kWarps) cascade8_auto_kernel(const C8AutoArgs a)
{
    extern __shared__ __align__(16) unsigned char c8a_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * kWarps + warp;
    if (s >= a.n_streams) return;                        // warps are independent
    constexpr int kH = 32 * CM + 2;                      // impulse-response samples kept per section: h_-1 .. h_(32 CM)
    unsigned char* wbase = c8a_smem + (size_t)warp * (2 * sizeof(WarpTab<T>) + 6 * kH * sizeof(double));
    WarpTab<T>& tm = reinterpret_cast<WarpTab<T>*>(wbase)[0];            // chunk length CM
    WarpTab<T>& tt = reinterpret_cast<WarpTab<T>*>(wbase)[1];            // chunk length 1 (tail tiles)
    double* hs = reinterpret_cast<double*>(wbase + 2 * sizeof(WarpTab<T>));
    double* st = a.state + (size_t)s * kC8State;
    T z[12], lp7, lp8;
#pragma unroll
    for (int k = 0; k < 12; ++k) z[k] = (T)st[k];
    lp7 = (T)st[12]; lp8 = (T)st[13];
    const float* in = a.in + (size_t)s * a.row_stride;
    float* out = a.out + (size_t)s * a.row_stride;
    double* tap = TAP ? a.tap + (size_t)s * a.row_stride : nullptr;
    const size_t ns = (size_t)a.n_streams;
    double cprev0 = 0.0, cprev1 = 0.0;                    // the coefficients the tables in shared memory were built from
    bool have_m = false, have_t = false;
    int64_t pos = 0;
    for (int blk = 0; pos < a.n_frames; ++blk) {
        const int nb = (int)min((int64_t)a.block_len, a.n_frames - pos);
        const double* c = a.coef + ((size_t)(blk < a.n_rows ? blk : a.n_rows - 1) * ns + s) * kC8Coefs;
        // new coefficients?  (bitwise: the rows of a block with unchanged sliders are copies of the previous block's)
        static_assert(kC8Coefs == 36, "cascade8 coefficient row: lanes 0 .. 31 watch c[lane], lanes 0 .. 3 also c[32 + lane]");
        const double c0 = c[lane], c1 = lane < 4 ? c[32 + lane] : 0.0;
        const bool changed = __any_sync(0xffffffffu, blk == 0 || c0 != cprev0 || c1 != cprev1);
        cprev0 = c0; cprev1 = c1;
        if (changed) {
            have_m = false; have_t = false;
            if (lane < 6) impulse_response<sizeof(T) == 8>(c[5 * lane + 3], c[5 * lane + 4], hs + lane * kH, 32 * CM);
            __syncwarp();
        }
        int f = 0;
        if (nb >= 32 * CM) {
            if (!have_m) { build_tables<T>(tm, c, hs, kH, CM, lane); have_m = true; }
            for (; f + 32 * CM <= nb; f += 32 * CM)
                tile8<T, CM, TAP>(tm, z, lp7, lp8, in + pos + f + lane * CM, out + pos + f + lane * CM, TAP ? tap + pos + f + lane * CM : nullptr, 32, lane);
        }
        if (f < nb) {
            if (!have_t) { build_tables<T>(tt, c, hs, kH, 1, lane); have_t = true; }
            for (; f < nb; f += 32)
                tile8<T, 1, TAP>(tt, z, lp7, lp8, in + pos + f + lane, out + pos + f + lane, TAP ? tap + pos + f + lane : nullptr, min(32, nb - f), lane);
        }
        pos += nb;
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 12; ++k) st[k] = (double)z[k];
        st[12] = (double)lp7; st[13] = (double)lp8;
    }
}

}  // namespace

size_t cascade8_auto_smem(int precision)
{
    (void)precision;
    return (size_t)kWarps * (2 * sizeof(WarpTab<double>) + 6 * (32 * 4 + 2) * sizeof(double));
}

bool cascade8_auto_launch(const float* in, float* out, double* tap, int64_t row_stride, int64_t n_frames, int n_streams, int block_len,
                          int n_rows, const double* coef, double* state, int precision, cudaStream_t stream)
{
    if (n_frames <= 0 || n_streams <= 0) return true;
    const C8AutoArgs a{in, out, tap, row_stride, n_frames, n_streams, block_len, n_rows, coef, state};
    const size_t smem = cascade8_auto_smem(precision);
    const unsigned grid = (unsigned)((n_streams + kWarps - 1) / kWarps);
    auto go = [&](auto kern) {
        if (smem > 48 * 1024 && !cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "cascade8 auto smem attr")) return false;
        kern<<<grid, 32 * kWarps, smem, stream>>>(a);
        return cuda_ok(cudaGetLastError(), "cascade8_auto_kernel launch");
    };
    // Float handles render automated calls in double too: float tables of near-unit-circle poles (entries of A^256) cost the
    // float mode its -100 dBFS (measured -98.8 with a float instantiation of this kernel), and this path is not the throughput
    // mode.  Only the store is float.
    (void)precision;
    return tap ? go(cascade8_auto_kernel<double, 4, true>) : go(cascade8_auto_kernel<double, 4, false>);
}

}  // namespace za
