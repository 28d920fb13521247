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
constexpr int kCoefRing = 4;      // host blocks alive between the filter warp (first) and the gain warp (last) when block_len >= kF

template <typename R> __device__ __forceinline__ R div_(R a, R b);
template <> __device__ __forceinline__ double div_<double>(double a, double b) { return __ddiv_rn(a, b); }
template <> __device__ __forceinline__ float div_<float>(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ double sqrt_(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ float sqrt_(float a) { return a * rsqrtf(a); }    // a >= 1e-12 here; 2 ulp, no slow-path branch
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
    R cs[kCoefRing][3][kMaxS];                    // thresh_db, depth_db, ratio_base of the last host blocks (slot = block & 3)
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

// ---- stage bodies over one group of 8 frames; FULL = every frame of the group exists (no per-frame branches, so the
// ---- compiler can interleave the frames' off-chain work with the recurrence) ---------------------------------------
template <typename R> struct BandK { R g, den, wj, band_c, a1, a2, bx, g2; };

template <typename R, bool FULL>
__device__ __forceinline__ void band_group(const R* xd, R* bpw, int f0, int nv, const BandK<R>& c, R& ic1, R& ic2, R& env)
{
    R xv[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) xv[k] = xd[f0 + k];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (FULL || f0 + k < nv) {
            // svf_bp (:74-93)
            R v1;
            if (sizeof(R) == 4) {
                // float: the same recurrence in state-space form (shorter dependent chain)
                v1 = fma_(c.a1, ic1, fma_(c.a2, ic2, mul(c.bx, xv[k])));
                ic1 = fma_((R)2.0, v1, -ic1);
                ic2 = fma_(c.g2, v1, ic2);
            } else {
                // x / den as x * (1 / den): one rounding more than the script's division (<= 1 ulp per frame in a stable
                // recursion -- measured 1e-15 of level against the reference, criterion 1e-12) for a third of the FP64 work
                v1 = mul(add(mul(c.g, sub(xv[k], ic2)), ic1), c.a1);
                const R v2 = add(ic2, mul(c.g, v1));
                ic1 = sub(mul((R)2.0, v1), ic1);
                ic2 = sub(mul((R)2.0, v2), ic2);
            }
            const R p = mul(v1, v1);
            env = add(env, mul(c.band_c, sub(p, env)));
            bpw[(f0 + k) * 9] = mul(c.wj, env);
        }
    }
}

template <typename R> struct FiltS { R hpf_lp, lpf_y, hpf_c, lpf_c; int hpf_on, lpf_on; };

template <typename R, bool FULL>
__device__ __forceinline__ void filter_group(const float* ib, R* xd, int f0, int nv, FiltS<R>& s)
{
    float lv[8], rv[8];
    R xo[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { lv[k] = ib[f0 + k]; rv[k] = ib[kPitchF + f0 + k]; }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        R x_det = mul((R)0.5, add((R)lv[k], (R)rv[k]));
        if (FULL || f0 + k < nv) {                                  // state must not advance past the end of the call
            if (s.hpf_on) { s.hpf_lp = add(s.hpf_lp, mul(s.hpf_c, sub(x_det, s.hpf_lp))); x_det = sub(x_det, s.hpf_lp); }
            if (s.lpf_on) { s.lpf_y = add(s.lpf_y, mul(s.lpf_c, sub(x_det, s.lpf_y))); x_det = s.lpf_y; }
        }
        xo[k] = x_det;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) xd[f0 + k] = xo[k];
}

template <typename R> struct GainS { R det_slow, gain_db_s, det_db_m, gr_db_m, thresh_db, thresh_hi; int expanding; };
template <typename R> struct GainK { R open_c, close_fast_c, close_slow_c, slow_c; };

template <typename R>
__device__ __forceinline__ void gain_frame(R det_lin, R det_db, R tgv, const GainK<R>& c, GainS<R>& s)
{
    s.det_db_m = add(s.det_db_m, mul((R)0.02, sub(det_db, s.det_db_m)));
    s.det_slow = add(s.det_slow, mul(c.slow_c, sub(det_lin, s.det_slow)));
    // hysteresis flag (:230-234) as predicate logic: lanes are different streams, a branch here would serialise them
    const int below = det_db < s.thresh_db ? 1 : 0, above = det_db > s.thresh_hi ? 1 : 0;
    s.expanding = (s.expanding & (above ^ 1)) | ((s.expanding ^ 1) & below);
    const R target_gr = s.expanding ? tgv : (R)0.0;
    const R target_gain_db = -target_gr;
    s.gr_db_m = add(s.gr_db_m, mul((R)0.05, sub(target_gr, s.gr_db_m)));
    const R close_c = (det_lin < mul(s.det_slow, (R)0.92)) ? c.close_fast_c : c.close_slow_c;
    // gain_db_s += (target > gain ? open_c : close_c) * (target - gain): both candidates, then select
    const R dlt = sub(target_gain_db, s.gain_db_s);
    const R g_open = add(s.gain_db_s, mul(c.open_c, dlt)), g_close = add(s.gain_db_s, mul(close_c, dlt));
    s.gain_db_s = (target_gain_db > s.gain_db_s) ? g_open : g_close;
}

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
    const bool smem_coef = a.block_len >= kF;                  // else a batch may span more host blocks than the ring holds

    const R eps = (R)1e-12, db_per_ln = (R)8.685889638065037, ln_per_db = (R)0.115129254649702;
    const R knee_db = (R)6.0, hys_db = (R)2.0;

    if (warp < kBandWarps) {
        // ============================ band warps =====================================================
        const int sl = lane >> 3, j = lane & 7;
        const int ls = 4 * warp + sl;                          // stream of this lane group inside the CTA
        const bool act = 4 * warp < ns_cta;                    // warp has streams at all
        const bool live = ls < ns_cta;
        const int sg = s_base + (live ? ls : (ns_cta - 1));    // clamped global stream (loads stay in range)
        // float state-space form of svf_bp: v1 = a1 ic1 + a2 ic2 + bx x; ic1' = 2 v1 - ic1; ic2' = ic2 + 2 g v1
        const BandK<R> bk{(R)a.k.g[j], (R)a.k.den[j], (R)a.k.w[j], (R)a.k.band_c, (R)(1.0 / a.k.den[j]),
                          (R)(-a.k.g[j] / a.k.den[j]), (R)(a.k.g[j] / a.k.den[j]), (R)(2.0 * a.k.g[j])};
        double* st = a.state + sg;
        R ic1 = (R)st[(0 + j) * ns], ic2 = (R)st[(8 + j) * ns], env = (R)st[(16 + j) * ns];
        R* bp = sm.bp[warp];
        float hold[4][2];
        int64_t next2 = a.block_len;                          // first frame of host block blk2 + 1 (S2's own block tracking)
        int blk2 = 0;
        // lane = frame stages: rows of this warp's four streams (clamped so loads stay in range)
        const float* irow[4]; float* orow[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const size_t s = (size_t)(s_base + min(4 * warp + q, max(ns_cta - 1, 0)));
            irow[q] = a.in + s * 2 * a.row_stride + lane; orow[q] = a.out + s * 2 * a.row_stride + lane;
        }

        for (int64_t i = 0; i < n_batches + 4; ++i) {
            if (act) {
                // ---- S0 (issue): batch i, lane = frame ------------------------------------------------
                {
                    const int64_t fo = i * kF;
                    const bool ok = i < n_batches && fo + lane < n;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        hold[q][0] = ok ? irow[q][fo] : 0.f;
                        hold[q][1] = ok ? irow[q][a.row_stride + fo] : 0.f;
                    }
                }
                // ---- S2: batch i-2, lane = (stream, band), frames in order ------------------------------
                if (i >= 2 && i - 2 < n_batches) {
                    const int64_t b = i - 2;
                    const R* xd = sm.xd[b & 1] + ls * kPitchF;
                    const int nv = (int)min((int64_t)kF, n - b * kF);
                    R* bpw = bp + sl * kBpStream + j;
                    if (nv == kF) {
#pragma unroll 1
                        for (int f0 = 0; f0 < kF; f0 += 8) band_group<R, true>(xd, bpw, f0, nv, bk, ic1, ic2, env);
                    } else {
#pragma unroll 1
                        for (int f0 = 0; f0 < kF; f0 += 8) band_group<R, false>(xd, bpw, f0, nv, bk, ic1, ic2, env);
                    }
                    __syncwarp();
                    // lane = frame: band sum in reference order, sqrt, log, memoryless gain computer
                    const int64_t F0 = b * kF;
                    while (F0 >= next2) { ++blk2; next2 += a.block_len; }
                    const int64_t fr = F0 + lane;
                    const int blk = smem_coef ? (blk2 + (fr >= next2 ? 1 : 0)) : bm.row(fr < n ? fr : n - 1);
                    R det_lin[4], det_db[4], thr[4], dep[4], rat[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int lq = 4 * warp + q;
                        R sum = (R)0.0;
#pragma unroll
                        for (int bnd = 0; bnd < 8; ++bnd) sum = add(sum, bp[q * kBpStream + lane * 9 + bnd]);
                        det_lin[q] = sqrt_(jmax(sum, eps));
                        if (smem_coef) {
                            thr[q] = sm.cs[blk & (kCoefRing - 1)][0][lq]; dep[q] = sm.cs[blk & (kCoefRing - 1)][1][lq];
                            rat[q] = sm.cs[blk & (kCoefRing - 1)][2][lq];
                        } else {
                            const EECoef& c = a.coef[(size_t)blk * ns + s_base + min(lq, ns_cta - 1)];
                            thr[q] = (R)c.thresh_db; dep[q] = (R)c.depth_db; rat[q] = (R)c.ratio_base;
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) det_db[q] = mul(db_per_ln, log_(jmax(det_lin[q], eps)));
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        R d = sub(thr[q], det_db[q]);
                        if (d < (R)0.0) d = (R)0.0;
                        const R xk = div_<R>(d, knee_db);
                        const R smo = mul(mul(xk, xk), sub((R)3.0, mul((R)2.0, xk)));
                        const R ratio_eff = (d < knee_db) ? add((R)1.0, mul(sub(rat[q], (R)1.0), smo)) : rat[q];
                        R target_gr = mul(d, sub(ratio_eff, (R)1.0));
                        if (target_gr > dep[q]) target_gr = dep[q];
                        const int o = (4 * warp + q) * kPitchF + lane;
                        sm.lin[b & 1][o] = det_lin[q]; sm.db[b & 1][o] = det_db[q]; sm.tg[b & 1][o] = target_gr;
                    }
                    __syncwarp();
                }
                // ---- S4: batch i-4, lane = frame: exp and the output multiplies ----------------------------
                if (i >= 4) {
                    const int64_t b = i - 4;
                    const int64_t fo = b * kF;
                    const bool okf = fo + lane < n;
                    const float* ib = sm.in[b & (kInBufs - 1)];
                    R o0[4], o1[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int lq = 4 * warp + q;
                        const R gain_lin = exp_(mul(sm.gain[b & 1][lq * kPitchF + lane], ln_per_db));
                        o0[q] = mul((R)ib[lq * kInPitch + lane], gain_lin);
                        o1[q] = mul((R)ib[lq * kInPitch + kPitchF + lane], gain_lin);
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (okf && 4 * warp + q < ns_cta) {
                            orow[q][fo] = (float)o0[q]; orow[q][a.row_stride + fo] = (float)o1[q];
                            if (TAP) {
                                double* tp = a.tap + (orow[q] - a.out) + fo;
                                tp[0] = (double)o0[q]; tp[a.row_stride] = (double)o1[q];
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
        FiltS<R> fs{(R)st[24 * ns], (R)st[25 * ns], (R)0, (R)0, 0, 0};
        int64_t next_block_frame = 0;
        int blk = 0;
        for (int64_t i = 0; i < n_batches + 4; ++i) {
            if (live && i >= 1 && i - 1 < n_batches) {
                const int64_t b = i - 1;
                const float* ib = sm.in[b & (kInBufs - 1)] + lane * kInPitch;
                R* xd = sm.xd[b & 1] + lane * kPitchF;
                const int nv = (int)min((int64_t)kF, n - b * kF);
#pragma unroll 1
                for (int f0 = 0; f0 < kF; f0 += 8) {
                    // host-block boundaries inside this group? (sliders are sampled once per block)
                    const int64_t g0 = b * kF + f0;
                    if (next_block_frame >= g0 + 8 && f0 + 8 <= nv) {
                        filter_group<R, true>(ib, xd, f0, nv, fs);
                    } else {
                        // at most the group's frames individually, reloading the coefficients where a block starts
                        for (int k = 0; k < 8; ++k) {
                            const int f = f0 + k;
                            if (f < nv && b * kF + f >= next_block_frame) {
                                const EECoef& c = a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * ns + sg];
                                fs.hpf_c = (R)c.hpf_c; fs.lpf_c = (R)c.lpf_c; fs.hpf_on = c.hpf_on; fs.lpf_on = c.lpf_on;
                                sm.cs[blk & (kCoefRing - 1)][0][lane] = (R)c.thresh_db;
                                sm.cs[blk & (kCoefRing - 1)][1][lane] = (R)c.depth_db;
                                sm.cs[blk & (kCoefRing - 1)][2][lane] = (R)c.ratio_base;
                                ++blk; next_block_frame += a.block_len;
                            }
                            R x_det = mul((R)0.5, add((R)ib[f], (R)ib[kPitchF + f]));
                            if (f < nv) {
                                if (fs.hpf_on) { fs.hpf_lp = add(fs.hpf_lp, mul(fs.hpf_c, sub(x_det, fs.hpf_lp))); x_det = sub(x_det, fs.hpf_lp); }
                                if (fs.lpf_on) { fs.lpf_y = add(fs.lpf_y, mul(fs.lpf_c, sub(x_det, fs.lpf_y))); x_det = fs.lpf_y; }
                            }
                            xd[f] = x_det;
                        }
                    }
                }
            }
            __syncthreads();
        }
        if (live) { st[24 * ns] = (double)fs.hpf_lp; st[25 * ns] = (double)fs.lpf_y; }
    } else {
        // ============================ gain warp: lane = stream ============================================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        const GainK<R> gk{(R)a.k.open_c, (R)a.k.close_fast_c, (R)a.k.close_slow_c, (R)a.k.slow_c};
        GainS<R> gs{(R)st[26 * ns], (R)st[28 * ns], (R)st[29 * ns], (R)st[30 * ns], (R)0, (R)0, st[27 * ns] != 0.0};
        int64_t next_block_frame = 0;
        int blk = 0;
        for (int64_t i = 0; i < n_batches + 4; ++i) {
            if (live && i >= 3 && i - 3 < n_batches) {
                const int64_t b = i - 3;
                const int o = lane * kPitchF;
                const R* lin = sm.lin[b & 1] + o; const R* db = sm.db[b & 1] + o; const R* tg = sm.tg[b & 1] + o;
                R* gout = sm.gain[b & 1] + o;
                const int nv = (int)min((int64_t)kF, n - b * kF);
#pragma unroll 1
                for (int f0 = 0; f0 < kF; f0 += 8) {
                    R lv[8], dv[8], tv[8], go[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) { lv[k] = lin[f0 + k]; dv[k] = db[f0 + k]; tv[k] = tg[f0 + k]; }
                    const int64_t g0 = b * kF + f0;
                    if (next_block_frame >= g0 + 8 && f0 + 8 <= nv) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) { gain_frame<R>(lv[k], dv[k], tv[k], gk, gs); go[k] = gs.gain_db_s; }
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const int f = f0 + k;
                            if (f < nv) {
                                if (g0 + k >= next_block_frame) {
                                    if (smem_coef) gs.thresh_db = sm.cs[blk & (kCoefRing - 1)][0][lane];
                                    else gs.thresh_db = (R)a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * ns + sg].thresh_db;
                                    gs.thresh_hi = add(gs.thresh_db, hys_db);
                                    ++blk; next_block_frame += a.block_len;
                                }
                                gain_frame<R>(lv[k], dv[k], tv[k], gk, gs);
                            }
                            go[k] = gs.gain_db_s;
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) gout[f0 + k] = go[k];
                }
            }
            __syncthreads();
        }
        if (live) {
            st[26 * ns] = (double)gs.det_slow; st[27 * ns] = gs.expanding ? 1.0 : 0.0; st[28 * ns] = (double)gs.gain_db_s;
            st[29 * ns] = (double)gs.det_db_m; st[30 * ns] = (double)gs.gr_db_m;
        }
    }
}

// ======================================================================================================================
// fp32 variant with the frame-parallel stages on their own warps.  In the kernel above the band warps are the critical
// path (ncu, profiles/r01_easyexpander_ncu.md: they never wait at the barrier): besides the band recurrence they also
// load the batch, run the frame-parallel detector / gain-computer stage and write the output.  Here those lane = frame
// stages move to 8 helper warps, so a band warp only runs the recurrence; one more pipeline step separates them:
//   S0 helpers (batch i) | S1 filter (i-1) | P1 band warps (i-2) | P2 helpers (i-3) | S3 gain warp (i-4) | S4 helpers (i-5)
// The band recurrence itself is issue-bound (ncu, profiles/r01_easyexpander_split_ncu.md: `not_selected` dominates),
// so it runs on Blackwell's packed fp32 pipe: a lane owns TWO bands of a stream in the halves of 64-bit registers
// (`fma.rn.f32x2`, SASS FFMA2), 4 lanes per stream, 8 streams per band warp, 4 band warps; the SVF is advanced in the
// state-space form  u = 2 a2 ic2 + 2 bx x,  ic1' = (2 a1 - 1) ic1 + u,  s = 2 a1 ic1 + u (= 2 v1),  ic2' = ic2 + g s,
// env' = (1 - c) env + (c / 4) s^2  -- 10 packed instructions per frame for two bands.
// Warp roles are placed by scheduler (warp id & 3): every scheduler gets one band warp; the gain warp shares its
// scheduler with nothing else, the filter warp with two helpers, the other six helpers sit three to a scheduler.
// Needs block_len >= 32 (coefficient ring) -- other shapes use the kernel above.
namespace {
constexpr int kHelpWarps = 8;
constexpr int kBandWarpsS = 4;
constexpr int kSplitWarps = 16;                       // launched; two of them have no role and leave at once
constexpr int kSplitLive = kBandWarpsS + 2 + kHelpWarps;
constexpr int kSplitThreads = kSplitWarps * 32;
constexpr int kInBufsS = 8;       // a batch's input lives from S0 (i) to S4 (i-5)
constexpr int kCoefRingS = 8;     // host blocks alive between the filter warp (i-1) and the gain warp (i-4)
constexpr int kBpPitch = 12;      // floats per (stream, frame): 8 weighted band powers + pad, so the two LDS.128 of a
                                  // lane = frame reader and the STS.64 of the band lanes are bank-conflict free
constexpr int kBpStreamS = kF * kBpPitch + 8;
enum { ROLE_BAND = 0, ROLE_FILTER = 4, ROLE_GAIN = 5, ROLE_NONE = 6, ROLE_HELP = 8 };
//                                   sched:  0              1              2          3
__constant__ int kRole[kSplitWarps] = {0,             1,             2,         3,
                                       ROLE_HELP + 0, ROLE_HELP + 1, ROLE_GAIN, ROLE_FILTER,
                                       ROLE_HELP + 2, ROLE_HELP + 3, ROLE_NONE, ROLE_HELP + 4,
                                       ROLE_HELP + 5, ROLE_HELP + 6, ROLE_NONE, ROLE_HELP + 7};
struct EESmemS {
    float in[kInBufsS][kMaxS * kInPitch];
    float xd[2][kMaxS * kPitchF];
    alignas(16) float bp[2][kMaxS * kBpStreamS];       // [stream][frame][band]
    float lin[2][kMaxS * kPitchF], db[2][kMaxS * kPitchF], tg[2][kMaxS * kPitchF], gain[2][kMaxS * kPitchF];
    float cs[kCoefRingS][3][kMaxS];
};
__device__ __forceinline__ void split_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kSplitLive * 32) : "memory"); }

// two floats in one 64-bit register
typedef unsigned long long p2;
__device__ __forceinline__ p2 pk(float lo, float hi) { p2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpk(p2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a)); }
__device__ __forceinline__ p2 fma2(p2 a, p2 b, p2 c) { p2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ p2 mul2(p2 a, p2 b) { p2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ p2 add2(p2 a, p2 b) { p2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ p2 mul2s(float s, p2 b) { p2 r; asm("{.reg .b64 t; mov.b64 t, {%1, %1}; mul.rn.f32x2 %0, t, %2;}" : "=l"(r) : "f"(s), "l"(b)); return r; }

struct BandK2 { p2 c1, a12, a22, bx2, g, omc, cq, wj; };
struct BandS2 { p2 ic1, ic2, env; };

template <bool FULL>
__device__ __forceinline__ void band_group2(const float* xd, float* bpw, int f0, int nv, const BandK2& c, BandS2& s)
{
    float xv[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) xv[k] = xd[f0 + k];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (FULL || f0 + k < nv) {
            const p2 u2 = fma2(c.a22, s.ic2, mul2s(xv[k], c.bx2));
            const p2 sv = fma2(c.a12, s.ic1, u2);                  // 2 v1
            s.ic1 = fma2(c.c1, s.ic1, u2);
            s.ic2 = fma2(c.g, sv, s.ic2);
            s.env = fma2(c.cq, mul2(sv, sv), mul2(c.omc, s.env));
            *reinterpret_cast<p2*>(bpw + (f0 + k) * kBpPitch) = mul2(c.wj, s.env);
        }
    }
}
}  // namespace

__global__ void __launch_bounds__(kSplitThreads, 1)
easyexpander_split_kernel(const EEArgs a, const int spc)
{
    typedef float R;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    EESmemS& sm = *reinterpret_cast<EESmemS*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int role = kRole[warp];
    if (role == ROLE_NONE) return;
    const int s_base = blockIdx.x * spc;
    const int ns_cta = min(spc, a.n_streams - s_base);
    const size_t ns = (size_t)a.n_streams;
    const int64_t n = a.n_frames;
    const int64_t n_batches = (n + kF - 1) / kF;
    const int64_t n_iter = n_batches + 5;
    const R eps = (R)1e-12, db_per_ln = (R)8.685889638065037, ln_per_db = (R)0.115129254649702;
    const R knee_db = (R)6.0, hys_db = (R)2.0;

    if (role < kBandWarpsS) {
        // ============================ band warps: the SVF + power-follower recurrence only, two bands per lane ========
        const int bw = role, sl = lane >> 2, jj = lane & 3;
        const int ls = 8 * bw + sl;
        const bool act = 8 * bw < ns_cta, live = ls < ns_cta;
        const int sg = s_base + (live ? ls : (ns_cta - 1));
        const int jA = 2 * jj, jB = jA + 1;
        const double gA = a.k.g[jA], gB = a.k.g[jB], dA = a.k.den[jA], dB = a.k.den[jB], bc = a.k.band_c;
        const BandK2 bk{pk((float)(2.0 / dA - 1.0), (float)(2.0 / dB - 1.0)), pk((float)(2.0 / dA), (float)(2.0 / dB)),
                        pk((float)(-2.0 * gA / dA), (float)(-2.0 * gB / dB)),
                        pk((float)(2.0 * gA / dA), (float)(2.0 * gB / dB)), pk((float)gA, (float)gB),
                        pk((float)(1.0 - bc), (float)(1.0 - bc)), pk((float)(0.25 * bc), (float)(0.25 * bc)),
                        pk((float)a.k.w[jA], (float)a.k.w[jB])};
        double* st = a.state + sg;
        BandS2 bs{pk((float)st[(0 + jA) * ns], (float)st[(0 + jB) * ns]), pk((float)st[(8 + jA) * ns], (float)st[(8 + jB) * ns]),
                  pk((float)st[(16 + jA) * ns], (float)st[(16 + jB) * ns])};
        for (int64_t i = 0; i < n_iter; ++i) {
            if (act && i >= 2 && i - 2 < n_batches) {
                const int64_t b = i - 2;
                const float* xd = sm.xd[b & 1] + ls * kPitchF;
                const int nv = (int)min((int64_t)kF, n - b * kF);
                float* bpw = sm.bp[b & 1] + ls * kBpStreamS + jA;
                if (nv == kF) {
#pragma unroll 1
                    for (int f0 = 0; f0 < kF; f0 += 8) band_group2<true>(xd, bpw, f0, nv, bk, bs);
                } else {
#pragma unroll 1
                    for (int f0 = 0; f0 < kF; f0 += 8) band_group2<false>(xd, bpw, f0, nv, bk, bs);
                }
            }
            split_sync();
        }
        if (live) {
            st = a.state + s_base + ls;
            float lo_, hi_;
            unpk(bs.ic1, lo_, hi_); st[(0 + jA) * ns] = (double)lo_; st[(0 + jB) * ns] = (double)hi_;
            unpk(bs.ic2, lo_, hi_); st[(8 + jA) * ns] = (double)lo_; st[(8 + jB) * ns] = (double)hi_;
            unpk(bs.env, lo_, hi_); st[(16 + jA) * ns] = (double)lo_; st[(16 + jB) * ns] = (double)hi_;
        }
    } else if (role == ROLE_FILTER) {
        // ============================ filter warp: lane = stream (batch i-1) =======================================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        FiltS<R> fs{(R)st[24 * ns], (R)st[25 * ns], (R)0, (R)0, 0, 0};
        int64_t next_block_frame = 0;
        int blk = 0;
        for (int64_t i = 0; i < n_iter; ++i) {
            if (live && i >= 1 && i - 1 < n_batches) {
                const int64_t b = i - 1;
                const float* ib = sm.in[b & (kInBufsS - 1)] + lane * kInPitch;
                R* xd = sm.xd[b & 1] + lane * kPitchF;
                const int nv = (int)min((int64_t)kF, n - b * kF);
#pragma unroll 1
                for (int f0 = 0; f0 < kF; f0 += 8) {
                    const int64_t g0 = b * kF + f0;
                    if (next_block_frame >= g0 + 8 && f0 + 8 <= nv) {
                        filter_group<R, true>(ib, xd, f0, nv, fs);
                    } else {
                        for (int k = 0; k < 8; ++k) {
                            const int f = f0 + k;
                            if (f < nv && b * kF + f >= next_block_frame) {      // host-block boundary: sliders sampled once per block
                                const EECoef& c = a.coef[(size_t)(blk < a.n_coef_blocks ? blk : a.n_coef_blocks - 1) * ns + sg];
                                fs.hpf_c = (R)c.hpf_c; fs.lpf_c = (R)c.lpf_c; fs.hpf_on = c.hpf_on; fs.lpf_on = c.lpf_on;
                                sm.cs[blk & (kCoefRingS - 1)][0][lane] = (R)c.thresh_db;
                                sm.cs[blk & (kCoefRingS - 1)][1][lane] = (R)c.depth_db;
                                sm.cs[blk & (kCoefRingS - 1)][2][lane] = (R)c.ratio_base;
                                ++blk; next_block_frame += a.block_len;
                            }
                            R x_det = mul((R)0.5, add((R)ib[f], (R)ib[kPitchF + f]));
                            if (f < nv) {
                                if (fs.hpf_on) { fs.hpf_lp = add(fs.hpf_lp, mul(fs.hpf_c, sub(x_det, fs.hpf_lp))); x_det = sub(x_det, fs.hpf_lp); }
                                if (fs.lpf_on) { fs.lpf_y = add(fs.lpf_y, mul(fs.lpf_c, sub(x_det, fs.lpf_y))); x_det = fs.lpf_y; }
                            }
                            xd[f] = x_det;
                        }
                    }
                }
            }
            split_sync();
        }
        if (live) { st[24 * ns] = (double)fs.hpf_lp; st[25 * ns] = (double)fs.lpf_y; }
    } else if (role == ROLE_GAIN) {
        // ============================ gain warp: lane = stream (batch i-4) =========================================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        const GainK<R> gk{(R)a.k.open_c, (R)a.k.close_fast_c, (R)a.k.close_slow_c, (R)a.k.slow_c};
        GainS<R> gs{(R)st[26 * ns], (R)st[28 * ns], (R)st[29 * ns], (R)st[30 * ns], (R)0, (R)0, st[27 * ns] != 0.0};
        int64_t next_block_frame = 0;
        int blk = 0;
        for (int64_t i = 0; i < n_iter; ++i) {
            if (live && i >= 4 && i - 4 < n_batches) {
                const int64_t b = i - 4;
                const int o = lane * kPitchF;
                const R* lin = sm.lin[b & 1] + o; const R* db = sm.db[b & 1] + o; const R* tg = sm.tg[b & 1] + o;
                R* gout = sm.gain[b & 1] + o;
                const int nv = (int)min((int64_t)kF, n - b * kF);
#pragma unroll 1
                for (int f0 = 0; f0 < kF; f0 += 8) {
                    R lv[8], dv[8], tv[8], go[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) { lv[k] = lin[f0 + k]; dv[k] = db[f0 + k]; tv[k] = tg[f0 + k]; }
                    const int64_t g0 = b * kF + f0;
                    if (next_block_frame >= g0 + 8 && f0 + 8 <= nv) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) { gain_frame<R>(lv[k], dv[k], tv[k], gk, gs); go[k] = gs.gain_db_s; }
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            if (f0 + k < nv) {
                                if (g0 + k >= next_block_frame) {
                                    gs.thresh_db = sm.cs[blk & (kCoefRingS - 1)][0][lane];
                                    gs.thresh_hi = add(gs.thresh_db, hys_db);
                                    ++blk; next_block_frame += a.block_len;
                                }
                                gain_frame<R>(lv[k], dv[k], tv[k], gk, gs);
                            }
                            go[k] = gs.gain_db_s;
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) gout[f0 + k] = go[k];
                }
            }
            split_sync();
        }
        if (live) {
            st[26 * ns] = (double)gs.det_slow; st[27 * ns] = gs.expanding ? 1.0 : 0.0; st[28 * ns] = (double)gs.gain_db_s;
            st[29 * ns] = (double)gs.det_db_m; st[30 * ns] = (double)gs.gr_db_m;
        }
    } else {
        // ============================ helper warps: lane = frame; streams h, h + 8, h + 16, h + 24 of the CTA ========
        const int h = role - ROLE_HELP;
        const float* irow[4]; float* orow[4];
        bool has[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int lq = h + kHelpWarps * q;
            has[q] = lq < ns_cta;
            const size_t s = (size_t)(s_base + (has[q] ? lq : 0));
            irow[q] = a.in + s * 2 * a.row_stride + lane; orow[q] = a.out + s * 2 * a.row_stride + lane;
        }
        float hold[4][2];
        int64_t next2 = a.block_len;                          // first frame of host block blk2 + 1 (P2's block tracking)
        int blk2 = 0;
        for (int64_t i = 0; i < n_iter; ++i) {
            // ---- S0 (issue): batch i ---------------------------------------------------------------------------
            {
                const int64_t fo = i * kF;
                const bool ok = i < n_batches && fo + lane < n;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    hold[q][0] = (ok && has[q]) ? irow[q][fo] : 0.f;
                    hold[q][1] = (ok && has[q]) ? irow[q][a.row_stride + fo] : 0.f;
                }
            }
            // ---- P2: batch i-3: band sum in reference order, sqrt, log, memoryless gain computer --------------------
            if (i >= 3 && i - 3 < n_batches) {
                const int64_t b = i - 3;
                const int64_t F0 = b * kF;
                while (F0 >= next2) { ++blk2; next2 += a.block_len; }
                const int slot = (blk2 + (F0 + lane >= next2 ? 1 : 0)) & (kCoefRingS - 1);
                R det_lin[4], det_db[4], thr[4], dep[4], rat[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int lq = h + kHelpWarps * q;
                    const float4* bp = reinterpret_cast<const float4*>(sm.bp[b & 1] + lq * kBpStreamS + lane * kBpPitch);
                    const float4 b03 = bp[0], b47 = bp[1];
                    R sum = add(add(add(add(add(add(add(add((R)0.0, b03.x), b03.y), b03.z), b03.w), b47.x), b47.y), b47.z), b47.w);
                    det_lin[q] = sqrt_(jmax(sum, eps));
                    thr[q] = sm.cs[slot][0][lq]; dep[q] = sm.cs[slot][1][lq]; rat[q] = sm.cs[slot][2][lq];
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) det_db[q] = mul(db_per_ln, log_(jmax(det_lin[q], eps)));
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    R d = sub(thr[q], det_db[q]);
                    if (d < (R)0.0) d = (R)0.0;
                    const R xk = div_<R>(d, knee_db);
                    const R smo = mul(mul(xk, xk), sub((R)3.0, mul((R)2.0, xk)));
                    const R ratio_eff = (d < knee_db) ? add((R)1.0, mul(sub(rat[q], (R)1.0), smo)) : rat[q];
                    R target_gr = mul(d, sub(ratio_eff, (R)1.0));
                    if (target_gr > dep[q]) target_gr = dep[q];
                    const int o = (h + kHelpWarps * q) * kPitchF + lane;
                    sm.lin[b & 1][o] = det_lin[q]; sm.db[b & 1][o] = det_db[q]; sm.tg[b & 1][o] = target_gr;
                }
            }
            // ---- S4: batch i-5: exp and the output multiplies --------------------------------------------------------
            if (i >= 5) {
                const int64_t b = i - 5;
                const int64_t fo = b * kF;
                const bool okf = fo + lane < n;
                const float* ib = sm.in[b & (kInBufsS - 1)];
                R o0[4], o1[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int lq = h + kHelpWarps * q;
                    const R gain_lin = exp_(mul(sm.gain[b & 1][lq * kPitchF + lane], ln_per_db));
                    o0[q] = mul((R)ib[lq * kInPitch + lane], gain_lin);
                    o1[q] = mul((R)ib[lq * kInPitch + kPitchF + lane], gain_lin);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (okf && has[q]) { orow[q][fo] = (float)o0[q]; orow[q][a.row_stride + fo] = (float)o1[q]; }
            }
            // ---- S0 (commit): batch i -> shared memory ------------------------------------------------------------------
            {
                float* ib = sm.in[i & (kInBufsS - 1)];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int lq = h + kHelpWarps * q;
                    ib[lq * kInPitch + lane] = hold[q][0];
                    ib[lq * kInPitch + kPitchF + lane] = hold[q][1];
                }
            }
            split_sync();
        }
    }
}

template <typename R, bool TAP>
static bool launch_ee(const EEArgs& a, int n_sm, cudaStream_t stream)
{
    auto kern = easyexpander_kernel<R, TAP>;
    const size_t smem = sizeof(EESmem<R>);
    static DeviceOnce attr_once;
    if (attr_once.pending()) {
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "EasyExpander smem attr")) return false;
        attr_once.mark();
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

static bool launch_ee_split(const EEArgs& a, int n_sm, cudaStream_t stream)
{
    const size_t smem = sizeof(EESmemS);
    static DeviceOnce attr_once;
    if (attr_once.pending()) {
        if (!cuda_ok(cudaFuncSetAttribute(easyexpander_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "EasyExpander smem attr")) return false;
        attr_once.mark();
    }
    int spc = (a.n_streams + n_sm - 1) / n_sm;
    spc = (spc + 3) / 4 * 4;
    if (spc > kMaxS) spc = kMaxS;
    if (spc < 4) spc = 4;
    const int blocks = (a.n_streams + spc - 1) / spc;
    easyexpander_split_kernel<<<blocks, kSplitThreads, smem, stream>>>(a, spc);
    return cuda_ok(cudaGetLastError(), "easyexpander_split_kernel launch");
}

bool easyexpander_launch(const EEArgs& a, int precision, int n_sm, cudaStream_t stream, int variant)
{
    if (a.n_frames <= 0 || a.n_streams <= 0) return true;
    if (precision) return a.tap ? launch_ee<double, true>(a, n_sm, stream) : launch_ee<double, false>(a, n_sm, stream);
    // fp32: variant 0 = the v3 role pipeline where it applies, 2 = the round-1 split kernel, 1 = the single-role kernel
    if (variant == 0 && easyexpander_v3_eligible(a)) return easyexpander_v3_launch(a, n_sm, stream);
    if (variant != 1 && a.block_len >= kF) return launch_ee_split(a, n_sm, stream);
    return launch_ee<float, false>(a, n_sm, stream);
}

}  // namespace za
