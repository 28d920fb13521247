// k_easyexpander.cu -- nonlinear dynamics class: plugin "EasyExpander" (BASELINE config 3) on sm_100a.
//
// Replaces the generated @sample loop of plugins/Dynamics/EasyExpander/src/EasyExpander.jsfx:177-281
// (jsfx_process_block, dsp_jsfx_aot.py:5713-5905) for n_streams instances at once.
//
// Time is sequential inside a stream, and 4096 streams x 8 detector bands is all the data parallelism there
// is (about one warp per scheduler), so the @sample body is cut into a software pipeline of stages that run
// CONCURRENTLY on different 32-frame batches, each stage with the thread mapping that suits it; one CTA
// (<= 32 streams) = 8 band warps + 1 filter warp + 1 gain warp, one __syncthreads per batch:
//   S0 band warps, lane = frame     coalesced 128-byte row loads of batch i -> shared memory
//   S1 filter warp, lane = stream   detector HPF/LPF one-poles (:180-194), in order over the batch (i-1)
//   S2 band warps, lane = (stream, band): the TPT-SVF band-pass + power follower of one band (:200-219), in
//      order over batch (i-2); the weighted band powers are transposed through (warp-private) shared memory;
//      then lane = frame: band sum in the reference's order, sqrt, log, and the MEMORYLESS half of the gain
//      computer (soft-knee target reduction as if `expanding` were set, :241-258)
//   S3 gain warp, lane = stream     what really is sequential (:224-275): detector meters, hysteresis flag,
//      switching-coefficient smoother, in order over batch (i-3): ~15 dependent operations per frame
//   S4 band warps, lane = frame     exp() of the smoothed gain, the two output multiplies, coalesced stores (i-4)
// Every value is computed by exactly the reference's sequence of operations (fp64: separate mul/add, true
// division, IEEE sqrt), only WHERE and WHEN differs, so the double tap matches the compiled reference to the
// last bits of libm's log/exp and the integer state (`expanding`) follows exactly.  fp32 mode computes in float.
// Per-block automation: derived @slider coefficients arrive as [block][stream] rows; every stage looks up the
// row of the host block its frame falls in (sliders are sampled once per block, JSFXJuceProcessor.cpp:3536-3556).
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

namespace {

constexpr int kF = 32;            // frames per batch
constexpr int kMaxS = 32;         // streams per CTA (one lane each in the filter / gain warps)
constexpr int kBandWarps = 8;
constexpr int kThreads = (kBandWarps + 2) * 32;
constexpr int kInBufs = 4;        // a batch's input lives from S0 to S4
constexpr int kPitchF = kF + 1;   // [stream][frame] rows read with lane = stream
constexpr int kInPitch = 2 * kPitchF + 1;
constexpr int kBpStream = kF * 9 + 8;

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

template <typename R> struct EESmem {
    float in[kInBufs][kMaxS * kInPitch];          // [buf][stream][ch][frame]
    R xd[2][kMaxS * kPitchF];                     // detector input after HPF/LPF
    R bp[kBandWarps][4 * kBpStream];              // [warp][stream][frame][band] weighted band powers
    R lin[2][kMaxS * kPitchF];                    // det_lin
    R db[2][kMaxS * kPitchF];                     // det_db
    R tg[2][kMaxS * kPitchF];                     // target reduction if expanding
    R gain[2][kMaxS * kPitchF];                   // gain_db_s after the frame
};

// host block of frame f (relative to the call), clamped to the coefficient rows there are
struct BlockMap {
    int block_len, n_rows;
    __device__ __forceinline__ int row(int64_t f) const
    {
        const int64_t b = f / block_len;
        return (int)(b < n_rows ? b : n_rows - 1);
    }
};

}  // namespace

// state layout in HBM: double [kEEState][n_streams] (structure of arrays, stream-minor)
//   0..7 ic1[b], 8..15 ic2[b], 16..23 env[b], 24 hpf_lp, 25 lpf_y, 26 det_slow, 27 expanding,
//   28 gain_db_s, 29 det_db_m, 30 gr_db_m
template <typename R, bool TAP>
__global__ void __launch_bounds__(kThreads, 1)
easyexpander_kernel(const EEArgs a, const int spc /* streams per CTA, multiple of 4, <= 32 */)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    EESmem<R>& sm = *reinterpret_cast<EESmem<R>*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s_base = blockIdx.x * spc;
    const int ns_cta = min(spc, a.n_streams - s_base);
    const size_t ns = (size_t)a.n_streams;
    const int64_t n = a.n_frames;
    const int64_t n_batches = (n + kF - 1) / kF;
    const BlockMap bm{a.block_len, a.n_coef_blocks};

    const R eps = (R)1e-12, db_per_ln = (R)8.685889638065037, ln_per_db = (R)0.115129254649702;
    const R knee_db = (R)6.0, hys_db = (R)2.0;

    if (warp < kBandWarps) {
        // ============================ band warps =====================================================
        const int sl = lane >> 3, j = lane & 7;
        const int ls = 4 * warp + sl;                          // stream of this lane group inside the CTA
        const bool act = 4 * warp < ns_cta;                    // warp has streams at all
        const bool live = ls < ns_cta;
        const int sg = s_base + (live ? ls : (ns_cta - 1));    // clamped global stream (loads stay in range)
        const R g = (R)a.k.g[j], den = (R)a.k.den[j], wj = (R)a.k.w[j], band_c = (R)a.k.band_c;
        double* st = a.state + sg;
        R ic1 = (R)st[(0 + j) * ns], ic2 = (R)st[(8 + j) * ns], env = (R)st[(16 + j) * ns];
        R* bp = sm.bp[warp];
        float hold[4][2];

        for (int64_t i = 0; i < n_batches + 4; ++i) {
            if (act) {
                // ---- S0 (issue): batch i, lane = frame ------------------------------------------------
                const int64_t f0 = i * kF + lane;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int s = s_base + min(4 * warp + q, ns_cta - 1);
                    const float* row = a.in + (size_t)s * 2 * a.row_stride;
                    const bool ok = i < n_batches && f0 < n;
                    hold[q][0] = ok ? row[f0] : 0.f;
                    hold[q][1] = ok ? row[a.row_stride + f0] : 0.f;
                }
                // ---- S2: batch i-2, lane = (stream, band), frames in order ------------------------------
                if (i >= 2 && i - 2 < n_batches) {
                    const int64_t b = i - 2;
                    const R* xd = sm.xd[b & 1] + ls * kPitchF;
                    const int nv = (int)min((int64_t)kF, n - b * kF);
#pragma unroll 8
                    for (int f = 0; f < kF; ++f) {
                        if (f < nv) {
                            const R x_det = xd[f];
                            // svf_bp (:74-93)
                            const R v1 = div_<R>(add(mul(g, sub(x_det, ic2)), ic1), den);
                            const R v2 = add(ic2, mul(g, v1));
                            ic1 = sub(mul((R)2.0, v1), ic1);
                            ic2 = sub(mul((R)2.0, v2), ic2);
                            const R p = mul(v1, v1);
                            env = add(env, mul(band_c, sub(p, env)));
                            bp[sl * kBpStream + f * 9 + j] = mul(wj, env);
                        }
                    }
                    __syncwarp();
                    // lane = frame: band sum in reference order, sqrt, log, memoryless gain computer
                    const int64_t fr = b * kF + lane;
                    const int row = bm.row(fr < n ? fr : n - 1);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int lq = 4 * warp + q;
                        const int s = s_base + min(lq, ns_cta - 1);
                        R sum = (R)0.0;
#pragma unroll
                        for (int bnd = 0; bnd < 8; ++bnd) sum = add(sum, bp[q * kBpStream + lane * 9 + bnd]);
                        const R det_lin = sqrt_(jmax(sum, eps));
                        const R det_db = mul(db_per_ln, log_(jmax(det_lin, eps)));
                        const EECoef& c = a.coef[(size_t)row * ns + s];
                        const R thresh_db = (R)c.thresh_db, depth_db = (R)c.depth_db, ratio_base = (R)c.ratio_base;
                        R d = sub(thresh_db, det_db);
                        if (d < (R)0.0) d = (R)0.0;
                        R ratio_eff;
                        if (d < knee_db) {
                            const R xk = div_<R>(d, knee_db);
                            const R smo = mul(mul(xk, xk), sub((R)3.0, mul((R)2.0, xk)));
                            ratio_eff = add((R)1.0, mul(sub(ratio_base, (R)1.0), smo));
                        } else {
                            ratio_eff = ratio_base;
                        }
                        R target_gr = mul(d, sub(ratio_eff, (R)1.0));
                        if (target_gr > depth_db) target_gr = depth_db;
                        const int o = lq * kPitchF + lane;
                        sm.lin[b & 1][o] = det_lin; sm.db[b & 1][o] = det_db; sm.tg[b & 1][o] = target_gr;
                    }
                    __syncwarp();
                }
                // ---- S4: batch i-4, lane = frame: exp and the output multiplies ----------------------------
                if (i >= 4) {
                    const int64_t b = i - 4;
                    const int64_t fr = b * kF + lane;
                    const float* ib = sm.in[b & (kInBufs - 1)];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int lq = 4 * warp + q;
                        if (lq < ns_cta && fr < n) {
                            const R gain_lin = exp_(mul(sm.gain[b & 1][lq * kPitchF + lane], ln_per_db));
                            const R o0 = mul((R)ib[lq * kInPitch + lane], gain_lin);
                            const R o1 = mul((R)ib[lq * kInPitch + kPitchF + lane], gain_lin);
                            float* orow = a.out + (size_t)(s_base + lq) * 2 * a.row_stride;
                            orow[fr] = (float)o0; orow[a.row_stride + fr] = (float)o1;
                            if (TAP) {
                                double* tp = a.tap + (size_t)(s_base + lq) * 2 * a.row_stride + fr;
                                tp[0] = (double)o0; tp[a.row_stride] = (double)o1;
                            }
                        }
                    }
                }
                // ---- S0 (commit): batch i -> shared memory (after S4 released this warp's rows of the slot) --
                {
                    float* ib = sm.in[i & (kInBufs - 1)];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        ib[(4 * warp + q) * kInPitch + lane] = hold[q][0];
                        ib[(4 * warp + q) * kInPitch + kPitchF + lane] = hold[q][1];
                    }
                }
            }
            __syncthreads();
        }
        if (live) {
            st = a.state + s_base + ls;
            st[(0 + j) * ns] = (double)ic1; st[(8 + j) * ns] = (double)ic2; st[(16 + j) * ns] = (double)env;
        }
    } else if (warp == kBandWarps) {
        // ============================ filter warp: lane = stream =========================================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        R hpf_lp = (R)st[24 * ns], lpf_y = (R)st[25 * ns];
        R hpf_c = 0, lpf_c = 0;
        int hpf_on = 0, lpf_on = 0;
        int64_t next_block_frame = 0;
        int blk = 0;
        for (int64_t i = 0; i < n_batches + 4; ++i) {
            if (live && i >= 1 && i - 1 < n_batches) {
                const int64_t b = i - 1;
                const float* ib = sm.in[b & (kInBufs - 1)] + lane * kInPitch;
                R* xd = sm.xd[b & 1] + lane * kPitchF;
                const int nv = (int)min((int64_t)kF, n - b * kF);
                int fb = (int)min((int64_t)kF, next_block_frame - b * kF);    // frame of the batch where the next host block starts
#pragma unroll 8
                for (int f = 0; f < kF; ++f) {
                    if (f < nv) {
                        if (f == fb) {                                 // host-block boundary: sliders sampled once per block
                            const EECoef& c = a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * ns + sg];
                            hpf_c = (R)c.hpf_c; lpf_c = (R)c.lpf_c; hpf_on = c.hpf_on; lpf_on = c.lpf_on;
                            ++blk; next_block_frame += a.block_len;
                            fb = (int)min((int64_t)kF, next_block_frame - b * kF);
                        }
                        R x_det = mul((R)0.5, add((R)ib[f], (R)ib[kPitchF + f]));
                        if (hpf_on) { hpf_lp = add(hpf_lp, mul(hpf_c, sub(x_det, hpf_lp))); x_det = sub(x_det, hpf_lp); }
                        if (lpf_on) { lpf_y = add(lpf_y, mul(lpf_c, sub(x_det, lpf_y))); x_det = lpf_y; }
                        xd[f] = x_det;
                    }
                }
            }
            __syncthreads();
        }
        if (live) { st[24 * ns] = (double)hpf_lp; st[25 * ns] = (double)lpf_y; }
    } else {
        // ============================ gain warp: lane = stream ============================================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        const R open_c = (R)a.k.open_c, close_fast_c = (R)a.k.close_fast_c, close_slow_c = (R)a.k.close_slow_c, slow_c = (R)a.k.slow_c;
        R det_slow = (R)st[26 * ns];
        int expanding = st[27 * ns] != 0.0;
        R gain_db_s = (R)st[28 * ns], det_db_m = (R)st[29 * ns], gr_db_m = (R)st[30 * ns];
        R thresh_db = 0, thresh_hi = 0;
        int64_t next_block_frame = 0;
        int blk = 0;
        for (int64_t i = 0; i < n_batches + 4; ++i) {
            if (live && i >= 3 && i - 3 < n_batches) {
                const int64_t b = i - 3;
                const int o = lane * kPitchF;
                const R* lin = sm.lin[b & 1] + o; const R* db = sm.db[b & 1] + o; const R* tg = sm.tg[b & 1] + o;
                R* gout = sm.gain[b & 1] + o;
                const int nv = (int)min((int64_t)kF, n - b * kF);
                int fb = (int)min((int64_t)kF, next_block_frame - b * kF);
#pragma unroll 8
                for (int f = 0; f < kF; ++f) {
                    if (f < nv) {
                        if (f == fb) {
                            const EECoef& c = a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * ns + sg];
                            thresh_db = (R)c.thresh_db; thresh_hi = add(thresh_db, hys_db);
                            ++blk; next_block_frame += a.block_len;
                            fb = (int)min((int64_t)kF, next_block_frame - b * kF);
                        }
                        const R det_lin = lin[f], det_db = db[f];
                        det_db_m = add(det_db_m, mul((R)0.02, sub(det_db, det_db_m)));
                        det_slow = add(det_slow, mul(slow_c, sub(det_lin, det_slow)));
                        if (expanding == 0) { if (det_db < thresh_db) expanding = 1; }
                        else                { if (det_db > thresh_hi) expanding = 0; }
                        const R target_gr = expanding ? tg[f] : (R)0.0;
                        const R target_gain_db = -target_gr;
                        gr_db_m = add(gr_db_m, mul((R)0.05, sub(target_gr, gr_db_m)));
                        const R close_c = (det_lin < mul(det_slow, (R)0.92)) ? close_fast_c : close_slow_c;
                        const R cc = (target_gain_db > gain_db_s) ? open_c : close_c;
                        gain_db_s = add(gain_db_s, mul(cc, sub(target_gain_db, gain_db_s)));
                        gout[f] = gain_db_s;
                    }
                }
            }
            __syncthreads();
        }
        if (live) {
            st[26 * ns] = (double)det_slow; st[27 * ns] = expanding ? 1.0 : 0.0; st[28 * ns] = (double)gain_db_s;
            st[29 * ns] = (double)det_db_m; st[30 * ns] = (double)gr_db_m;
        }
    }
}

template <typename R, bool TAP>
static bool launch_ee(const EEArgs& a, int n_sm, cudaStream_t stream)
{
    auto kern = easyexpander_kernel<R, TAP>;
    const size_t smem = sizeof(EESmem<R>);
    static bool attr_done = false;
    if (!attr_done) {
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "EasyExpander smem attr")) return false;
        attr_done = true;
    }
    // streams per CTA: spread the streams over all SMs, whole band warps (4 streams), at most one lane each
    int spc = (a.n_streams + n_sm - 1) / n_sm;
    spc = (spc + 3) / 4 * 4;
    if (spc > kMaxS) spc = kMaxS;
    if (spc < 4) spc = 4;
    const int blocks = (a.n_streams + spc - 1) / spc;
    kern<<<blocks, kThreads, smem, stream>>>(a, spc);
    return cuda_ok(cudaGetLastError(), "easyexpander_kernel launch");
}

bool easyexpander_launch(const EEArgs& a, int precision, int n_sm, cudaStream_t stream)
{
    if (a.n_frames <= 0 || a.n_streams <= 0) return true;
    if (precision) return a.tap ? launch_ee<double, true>(a, n_sm, stream) : launch_ee<double, false>(a, n_sm, stream);
    return launch_ee<float, false>(a, n_sm, stream);
}

}  // namespace za
