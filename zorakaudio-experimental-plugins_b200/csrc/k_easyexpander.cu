// k_easyexpander.cu -- nonlinear dynamics class: plugin "EasyExpander" (BASELINE config 3) on sm_100a.
//
// Replaces the generated @sample loop of plugins/Dynamics/EasyExpander/src/EasyExpander.jsfx:177-281
// (jsfx_process_block, dsp_jsfx_aot.py:5713-5905) for n_streams instances at once.
//
// Mapping.  Time is sequential inside a stream (hysteresis flag, switching-coefficient smoother), so
// parallelism comes from streams x detector bands x a short batch of frames:
//   * one warp = 4 streams; lane = (stream s = lane >> 3, j = lane & 7); all state in registers;
//   * frames go in batches of 8.  Phase 1 (lane = band j): the detector one-poles (replicated in
//     the 8 lanes of a stream) and the stream's j-th TPT-SVF band-pass + power follower run over
//     the 8 frames (EasyExpander.jsfx:177-219); the 8x8 weighted band powers are transposed
//     through shared memory so that Phase 2 (lane = frame j) sums them in the reference's band
//     order and takes sqrt/log for ITS frame -- the transcendental work is spread over all lanes;
//   * Phase 3 (replicated per stream): the 8 frames' hysteresis / gain computer / smoother
//     (:221-275) run in order on shuffled (det_lin, det_db); lane j keeps the gain after frame j;
//   * Phase 4 (lane = frame j): exp() and the two output multiplies (:277-281).
// fp64 mode keeps the reference's operation order and rounding (separate mul/add, true division,
// IEEE sqrt), so only libm's log/exp (<= 1 ulp either side) can differ from the compiled reference;
// the integer state (`expanding`) follows exactly.  fp32 mode computes in float.
// Per-block automation: derived @slider coefficients arrive as [block][stream] rows and are
// re-read at every host-block boundary (sliders are sampled once per block, :3536-3556).
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

template <typename R> __device__ __forceinline__ R div_(R a, R b);
template <> __device__ __forceinline__ double div_<double>(double a, double b) { return __ddiv_rn(a, b); }
template <> __device__ __forceinline__ float div_<float>(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ double sqrt_(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ float sqrt_(float a) { return sqrtf(a); }
__device__ __forceinline__ double log_(double a) { return log(a); }
__device__ __forceinline__ float log_(float a) { return __logf(a); }
__device__ __forceinline__ double exp_(double a) { return exp(a); }
__device__ __forceinline__ float exp_(float a) { return __expf(a); }
template <typename R> __device__ __forceinline__ R jmax(R a, R b) { return a > b ? a : b; }   // ogt select, dsp_jsfx_aot.py:5223-5232

template <typename R>
__device__ __forceinline__ R shfl_r(R v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// state layout in HBM: double [kEEState][n_streams] (structure of arrays, stream-minor)
//   0..7 ic1[b], 8..15 ic2[b], 16..23 env[b], 24 hpf_lp, 25 lpf_y, 26 det_slow, 27 expanding,
//   28 gain_db_s, 29 det_db_m, 30 gr_db_m
template <typename R, bool TAP>
__global__ void __launch_bounds__(32)
easyexpander_kernel(const EEArgs a)
{
    __shared__ R tr[4][8][9];                 // [stream][frame][band] (+1 pad): weighted band powers of one batch
    const int lane = threadIdx.x, sl = lane >> 3, j = lane & 7;
    const int s = blockIdx.x * 4 + sl;        // stream of this lane group
    const bool live = s < a.n_streams;
    const int sc = live ? s : a.n_streams - 1;
    const int gbase = lane & ~7;

    // band constants (@init, same for every stream at one sample rate)
    const R g = (R)a.k.g[j], kk = (R)a.k.k[j], den = (R)a.k.den[j], w = (R)a.k.w[j];
    const R band_c = (R)a.k.band_c, open_c = (R)a.k.open_c, close_fast_c = (R)a.k.close_fast_c;
    const R close_slow_c = (R)a.k.close_slow_c, slow_c = (R)a.k.slow_c;
    const R eps = (R)1e-12, db_per_ln = (R)8.685889638065037, ln_per_db = (R)0.115129254649702;
    const R knee_db = (R)6.0, hys_db = (R)2.0;

    double* st = a.state + sc;
    const size_t ns = (size_t)a.n_streams;
    R ic1 = (R)st[(0 + j) * ns], ic2 = (R)st[(8 + j) * ns], env = (R)st[(16 + j) * ns];
    R hpf_lp = (R)st[24 * ns], lpf_y = (R)st[25 * ns], det_slow = (R)st[26 * ns];
    int expanding = st[27 * ns] != 0.0;
    R gain_db_s = (R)st[28 * ns], det_db_m = (R)st[29 * ns], gr_db_m = (R)st[30 * ns];

    const float* inL = a.in + (size_t)sc * 2 * a.row_stride;
    const float* inR = inL + a.row_stride;
    float* outL = a.out + (size_t)sc * 2 * a.row_stride;
    float* outR = outL + a.row_stride;

    // derived @slider values of the current host block
    R thresh_db = 0, depth_db = 0, ratio_base = 1, hpf_c = 0, lpf_c = 0;
    int hpf_on = 0, lpf_on = 0;
    int64_t next_block_frame = 0;
    int blk = 0;

    const int64_t n = a.n_frames;
    float l_nx = 0.f, r_nx = 0.f;
    if (j < n) { l_nx = inL[j]; r_nx = inR[j]; }
    for (int64_t t0 = 0; t0 < n; t0 += 8) {
        if (t0 >= next_block_frame) {           // host-block boundary: sliders are sampled once per block
            const int row = blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1;
            const EECoef c = a.coef[(size_t)row * a.n_streams + sc];
            thresh_db = (R)c.thresh_db; depth_db = (R)c.depth_db; ratio_base = (R)c.ratio_base;
            hpf_c = (R)c.hpf_c; lpf_c = (R)c.lpf_c; hpf_on = c.hpf_on; lpf_on = c.lpf_on;
            ++blk; next_block_frame += a.block_len;
        }
        const float l_in = l_nx, r_in = r_nx;
        if (t0 + 8 + j < n) { l_nx = inL[t0 + 8 + j]; r_nx = inR[t0 + 8 + j]; }     // next batch, in flight during this one
        const int nv = (int)((n - t0) < 8 ? (n - t0) : 8);

        // ---- phase 1: lane = band j, frames in order -------------------------------------------
#pragma unroll
        for (int f = 0; f < 8; ++f) {
            const R l = (R)shfl_r(l_in, gbase + f), r = (R)shfl_r(r_in, gbase + f);
            if (f < nv) {
                R x_det = mul((R)0.5, add(l, r));
                if (hpf_on) { hpf_lp = add(hpf_lp, mul(hpf_c, sub(x_det, hpf_lp))); x_det = sub(x_det, hpf_lp); }
                if (lpf_on) { lpf_y = add(lpf_y, mul(lpf_c, sub(x_det, lpf_y))); x_det = lpf_y; }
                // svf_bp (:74-93)
                const R v1 = div_<R>(add(mul(g, sub(x_det, ic2)), ic1), den);
                const R v2 = add(ic2, mul(g, v1));
                ic1 = sub(mul((R)2.0, v1), ic1);
                ic2 = sub(mul((R)2.0, v2), ic2);
                const R p = mul(v1, v1);
                env = add(env, mul(band_c, sub(p, env)));
                tr[sl][f][j] = mul(w, env);
            }
        }
        __syncwarp();
        // ---- phase 2: lane = frame j: band sum in reference order, sqrt, log ------------------------
        R sum = (R)0.0;
#pragma unroll
        for (int b = 0; b < 8; ++b) sum = add(sum, tr[sl][j][b]);
        __syncwarp();
        const R det_lin_j = sqrt_(jmax(sum, eps));
        const R det_db_j = mul(db_per_ln, log_(jmax(det_lin_j, eps)));

        // ---- phase 3: replicated per stream, frames in order ---------------------------------------
        R gain_j = gain_db_s;
#pragma unroll
        for (int f = 0; f < 8; ++f) {
            const R det_lin = shfl_r(det_lin_j, gbase + f), det_db = shfl_r(det_db_j, gbase + f);
            if (f < nv) {
                det_db_m = add(det_db_m, mul((R)0.02, sub(det_db, det_db_m)));
                det_slow = add(det_slow, mul(slow_c, sub(det_lin, det_slow)));
                if (expanding == 0) { if (det_db < thresh_db) expanding = 1; }
                else                { if (det_db > add(thresh_db, hys_db)) expanding = 0; }
                R target_gr = (R)0.0;
                if (expanding) {
                    R d = sub(thresh_db, det_db);
                    if (d < (R)0.0) d = (R)0.0;
                    R ratio_eff;
                    if (d < knee_db) {
                        const R xk = div_<R>(d, knee_db);
                        const R sm = mul(mul(xk, xk), sub((R)3.0, mul((R)2.0, xk)));
                        ratio_eff = add((R)1.0, mul(sub(ratio_base, (R)1.0), sm));
                    } else {
                        ratio_eff = ratio_base;
                    }
                    target_gr = mul(d, sub(ratio_eff, (R)1.0));
                    if (target_gr > depth_db) target_gr = depth_db;
                }
                const R target_gain_db = -target_gr;
                gr_db_m = add(gr_db_m, mul((R)0.05, sub(target_gr, gr_db_m)));
                const R close_c = (det_lin < mul(det_slow, (R)0.92)) ? close_fast_c : close_slow_c;
                const R cc = (target_gain_db > gain_db_s) ? open_c : close_c;
                gain_db_s = add(gain_db_s, mul(cc, sub(target_gain_db, gain_db_s)));
                if (f == j) gain_j = gain_db_s;
            }
        }
        // ---- phase 4: lane = frame j: exp and the output multiplies ---------------------------------
        if (live && j < nv) {
            const R gain_lin = exp_(mul(gain_j, ln_per_db));
            const R o0 = mul((R)l_in, gain_lin), o1 = mul((R)r_in, gain_lin);
            outL[t0 + j] = (float)o0; outR[t0 + j] = (float)o1;
            if (TAP) {
                double* tp = a.tap + (size_t)s * 2 * a.row_stride + t0 + j;
                tp[0] = (double)o0; tp[a.row_stride] = (double)o1;
            }
        }
    }

    if (live) {
        st = a.state + s;
        st[(0 + j) * ns] = (double)ic1; st[(8 + j) * ns] = (double)ic2; st[(16 + j) * ns] = (double)env;
        if (j == 0) {
            st[24 * ns] = (double)hpf_lp; st[25 * ns] = (double)lpf_y; st[26 * ns] = (double)det_slow;
            st[27 * ns] = expanding ? 1.0 : 0.0; st[28 * ns] = (double)gain_db_s;
            st[29 * ns] = (double)det_db_m; st[30 * ns] = (double)gr_db_m;
        }
    }
}

bool easyexpander_launch(const EEArgs& a, int precision, cudaStream_t stream)
{
    if (a.n_frames <= 0 || a.n_streams <= 0) return true;
    const int blocks = (a.n_streams + 3) / 4;
    if (precision) {
        if (a.tap) easyexpander_kernel<double, true><<<blocks, 32, 0, stream>>>(a);
        else       easyexpander_kernel<double, false><<<blocks, 32, 0, stream>>>(a);
    } else {
        easyexpander_kernel<float, false><<<blocks, 32, 0, stream>>>(a);
    }
    return cuda_ok(cudaGetLastError(), "easyexpander_kernel launch");
}

}  // namespace za
