// k_easyexpander_v3.cu -- EasyExpander fp32, third cut of the role pipeline (BASELINE config 3, the headline workload).
//
// Same algorithm and the same stage idea as easyexpander_split_kernel (k_easyexpander.cu; EasyExpander.jsfx:177-281), rebuilt
// around what its ncu capture showed (profiles/r01_easyexpander_split_ncu.md, tools/ncu_roles.py): nobody waits at the batch
// barrier for long, every role is busy -- the lane = stream gain warp is the critical path (879 instructions per 32-frame
// batch at ~0.4 IPC), and the lane = frame helper warps spend half their instructions on addresses, denormal guards and
// per-frame block bookkeeping.  Changes:
//   * host blocks are a whole number of 32-frame batches here (block_len % 32 == 0; other shapes keep the older kernels),
//     so every coefficient is a per-batch constant and no stage tracks blocks per frame;
//   * the sequential gain computer is cut into two warps one batch apart: G1 = slow detector follower + adaptive closing
//     coefficient + detector meter, G2 = hysteresis flag + reduction meter + switching smoother;
//   * rows that a lane = stream warp walks are [stream][36] (16-byte aligned, bank-conflict free both ways), read and written
//     with 128-bit shared-memory accesses: 4x fewer LDS/STS in the filter / G1 / G2 warps and for the band warps' input;
//   * band recurrence in scaled coordinates: states divided by the input gain 2 g/den, the power follower carried as
//     w_j * env_j with w_j c (g/den)^2 folded into one constant -- 7 packed instructions per frame and band pair instead of
//     10; the HBM state keeps the script's own variables (exact round trip through double);
//   * helpers: MUFU forms without the denormal guards (every argument is >= 1e-12), 10 log10(sum) instead of
//     sqrt -> log, pointers bumped instead of rebuilt, 7 helper warps own the 28 streams of a cfg3 CTA (no idle quarter);
//   * the band-power tile is [stream][frame][8] with a 16-byte XOR swizzle instead of a 12-float pitch (32 KB less shared
//     memory, which pays for the deeper db / tg rings G2 needs).
// Pipeline per 32-frame batch b (iteration i):  S0 helpers load b = i | filter b = i-1 | bands i-2 | P2 helpers i-3 |
// G1 i-4 | G2 i-5 | S4 helpers i-6 (exp, output multiplies, stores).  One named barrier per iteration.
// fp32 criterion: residual below -100 dBFS against the float64 reference (tests/test_easyexpander_gpu.py).
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

namespace {

constexpr int kF = 32;             // frames per batch
constexpr int kMaxS = 32;          // streams per CTA
constexpr int kRowP = 36;          // floats per [stream] row of a 32-frame tile
constexpr int kIoP = 64;           // I/O ring: [stream][ch][32 frames], 16-byte chunk c of a row stored at c ^ (stream & 7)
constexpr int kBpP = kF * 4 + 4;   // [stream][frame][4 band pairs] + 4: pair sums w_j env_j + w_j+1 env_j+1 of lane jj
constexpr int kSB = 4;             // batches per super-batch: global memory is read and written in 512-byte runs per row
constexpr int kIoSlots = 16;       // I/O ring: a batch's slot is filled by cp.async six iterations early, read by the filter,
                                   // overwritten in place by S4 (output) and released by the super-batch's 128-bit stores
constexpr int kLead = 6;           // iterations between the cp.async issue of a super-batch and its first batch
constexpr int kDbSlots = 3;        // det_db / target rows live from P2 (i) to G2 (i + 2)

#ifndef ZA_EE_KQ
#define ZA_EE_KQ 2
#endif
constexpr int kQ = ZA_EE_KQ;       // streams per helper warp (2: 16 helper warps, 24 warps per CTA; 4: 8 helper warps, 16 warps per CTA)
enum { R_FILTER = 4, R_G1 = 5, R_G2 = 6, R_NONE = 7, R_HELP = 8 };
// Warp w runs on scheduler w & 3.  Every scheduler gets one band warp; the sequential warps (G2, G1, filter) take one
// scheduler each and the helpers fill up: with the 28 streams of a cfg3 CTA helpers 0..13 (kQ = 2) or 0..6 (kQ = 4) are active.
#ifndef ZA_EE_PPL
#define ZA_EE_PPL 1
#endif
constexpr int kPPL = ZA_EE_PPL;    // band pairs per lane of a band warp: 1 (default) = four band warps of 8 streams, one dependent
                                   // chain per lane; 2 = two band warps of 16 streams with two independent chains per lane.
                                   // Measured on B200 (build switch ZA_EE_PPL=2): 3x SLOWER (31.4 against 10.6 ms per segment) --
                                   // the chain interleaves as hoped, but the band work (2300 issue slots per batch) then sits on
                                   // two schedulers instead of four and, under the 80-register cap of 768 threads, the compiler
                                   // rebuilds the store address every frame; kept as a negative result.
constexpr int kBandWarpsN = 4 / kPPL, kLPS = 4 / kPPL;      // band warps; lanes per stream
#if ZA_EE_KQ == 2 && ZA_EE_PPL == 2
constexpr int kWarps = 24;
// two band warps, each with a scheduler it shares with two helpers only; the sequential warps and the other helpers on the other two
//                                  scheduler: 0            1            2            3
__constant__ int kRoleV3[kWarps] = {0,           1,           R_G2,        R_G1,
                                    R_HELP + 0,  R_HELP + 2,  R_FILTER,    R_HELP + 8,
                                    R_HELP + 1,  R_HELP + 3,  R_HELP + 4,  R_HELP + 9,
                                    R_HELP + 14, R_HELP + 13, R_HELP + 5,  R_HELP + 10,
                                    R_NONE,      R_HELP + 15, R_HELP + 6,  R_HELP + 11,
                                    R_NONE,      R_NONE,      R_HELP + 7,  R_HELP + 12};
constexpr int kLiveWarps = 21;
#elif ZA_EE_KQ == 2
constexpr int kWarps = 24;
//                                  scheduler: 0            1            2            3
__constant__ int kRoleV3[kWarps] = {0,           1,           2,           3,
                                    R_G2,        R_G1,        R_FILTER,    R_HELP + 0,
                                    R_HELP + 1,  R_HELP + 2,  R_HELP + 3,  R_HELP + 4,
                                    R_HELP + 5,  R_HELP + 6,  R_HELP + 7,  R_HELP + 8,
                                    R_HELP + 14, R_HELP + 9,  R_HELP + 10, R_HELP + 11,
                                    R_NONE,      R_HELP + 15, R_HELP + 12, R_HELP + 13};
constexpr int kLiveWarps = 23;
#else
static_assert(ZA_EE_PPL == 1, "four streams per helper: only with four band warps");
constexpr int kWarps = 16;
__constant__ int kRoleV3[kWarps] = {0,           1,           2,           3,
                                    R_G2,        R_G1,        R_FILTER,    R_HELP + 3,
                                    R_HELP + 0,  R_HELP + 1,  R_HELP + 2,  R_HELP + 6,
                                    R_NONE,      R_HELP + 4,  R_HELP + 5,  R_HELP + 7};
constexpr int kLiveWarps = 15;
#endif
constexpr int kThreads = kWarps * 32;

struct SmemV3 {
    alignas(16) float io[kIoSlots][kMaxS * kIoP];
    alignas(16) float xd[2][kMaxS * kRowP];
    alignas(16) float bp[2][kMaxS * kBpP];
    alignas(16) float lin[2][kMaxS * kRowP];
    alignas(16) float db[kDbSlots][kMaxS * kRowP];
    alignas(16) float tg[kDbSlots][kMaxS * kRowP];
    alignas(16) float cc[2][kMaxS * kRowP];
    alignas(16) float gain[2][kMaxS * kRowP];
};
static_assert(sizeof(SmemV3) <= 227 * 1024, "EasyExpander v3 shared memory");

__device__ __forceinline__ void v3_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kLiveWarps * 32) : "memory"); }

typedef unsigned long long p2;     // two floats in one 64-bit register
__device__ __forceinline__ p2 pk(float lo, float hi) { p2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpk(p2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a)); }
__device__ __forceinline__ p2 fma2(p2 a, p2 b, p2 c) { p2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ p2 mul2(p2 a, p2 b) { p2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

__device__ __forceinline__ float rsqrt_fast(float a) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float lg2_fast(float a) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float ex2_fast(float a) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }

// first frame of the host block a batch starts in -> coefficient row (clamped to the rows there are)
struct BlockWalk {
    int64_t next;      // first frame of the next host block
    int row;           // row of the current one (-1 before the first batch)
    int block_len, n_rows;
    __device__ __forceinline__ bool enter(int64_t F0)       // true when batch F0 starts a new host block
    {
        if (F0 < next) return false;
        while (F0 >= next) { ++row; next += block_len; }
        return true;
    }
};

struct BandK3 { p2 a22, a12, c1, g, omc, kq; };
struct BandS3 { p2 i1, i2, e; };

// eight frames of one band pair: u = 2 a2 i2 + x,  s = 2 a1 i1 + u (= 2 v1 / k),  i1' = (2 a1 - 1) i1 + u,  i2' = i2 + g s,
// E' = (1 - c) E + (w c k^2 / 4) s^2, with k = 2 g / den the input gain the states are divided by (svf_bp, :74-93; :200-219)
template <bool FULL>
__device__ __forceinline__ void band_group3(const float* xd, float* pa, int f0, int nv, bool wr, const BandK3 (&c)[kPPL], BandS3 (&s)[kPPL])
{
    float xv[8];
    if (FULL) {
        const float4 x0 = *reinterpret_cast<const float4*>(xd + f0), x1 = *reinterpret_cast<const float4*>(xd + f0 + 4);
        xv[0] = x0.x; xv[1] = x0.y; xv[2] = x0.z; xv[3] = x0.w; xv[4] = x1.x; xv[5] = x1.y; xv[6] = x1.z; xv[7] = x1.w;
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) xv[k] = xd[f0 + k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (FULL || f0 + k < nv) {
            float ps[kPPL];
#pragma unroll
            for (int p = 0; p < kPPL; ++p) {                               // kPPL independent chains
                const p2 u = fma2(c[p].a22, s[p].i2, pk(xv[k], xv[k]));
                const p2 sv = fma2(c[p].a12, s[p].i1, u);
                s[p].i1 = fma2(c[p].c1, s[p].i1, u);
                s[p].i2 = fma2(c[p].g, sv, s[p].i2);
                s[p].e = fma2(c[p].kq, mul2(sv, sv), mul2(c[p].omc, s[p].e));
                float e0, e1;
                unpk(s[p].e, e0, e1);
                ps[p] = e0 + e1;
            }
            if (wr) {
                if (kPPL == 2) *reinterpret_cast<float2*>(pa + (f0 + k) * 4) = make_float2(ps[0], ps[kPPL - 1]);
                else pa[(f0 + k) * 4] = ps[0];
            }
        }
    }
}

// detector input + the detector-only one-poles (:180-194); MODE 0 = no stream of the warp has either filter on
template <int MODE>
__device__ __forceinline__ float filt_frame(float l, float r, float& hpf_lp, float& lpf_y, float hpf_c, float lpf_c, bool hpf_on, bool lpf_on)
{
    float x = 0.5f * (l + r);
    if (MODE) {
        const float lp = fmaf(hpf_c, x - hpf_lp, hpf_lp);
        hpf_lp = hpf_on ? lp : hpf_lp;
        x = hpf_on ? x - hpf_lp : x;
        const float ly = fmaf(lpf_c, x - lpf_y, lpf_y);
        lpf_y = lpf_on ? ly : lpf_y;
        x = lpf_on ? lpf_y : x;
    }
    return x;
}

template <int MODE>
__device__ __forceinline__ void filt_batch(const float* ib, int sw, float* xo, int nv, bool wr, float& hpf_lp, float& lpf_y, float hpf_c, float lpf_c,
                                           bool hpf_on, bool lpf_on)
{
    if (nv == kF) {
#pragma unroll
        for (int c0 = 0; c0 < 8; c0 += 2) {
            const int pa = ((c0 ^ sw) << 2), pb = (((c0 + 1) ^ sw) << 2);
            const float4 l0 = *reinterpret_cast<const float4*>(ib + pa), l1 = *reinterpret_cast<const float4*>(ib + pb);
            const float4 r0 = *reinterpret_cast<const float4*>(ib + 32 + pa), r1 = *reinterpret_cast<const float4*>(ib + 32 + pb);
            float4 o0, o1;
            o0.x = filt_frame<MODE>(l0.x, r0.x, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o0.y = filt_frame<MODE>(l0.y, r0.y, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o0.z = filt_frame<MODE>(l0.z, r0.z, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o0.w = filt_frame<MODE>(l0.w, r0.w, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o1.x = filt_frame<MODE>(l1.x, r1.x, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o1.y = filt_frame<MODE>(l1.y, r1.y, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o1.z = filt_frame<MODE>(l1.z, r1.z, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            o1.w = filt_frame<MODE>(l1.w, r1.w, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            if (wr) { *reinterpret_cast<float4*>(xo + 4 * c0) = o0; *reinterpret_cast<float4*>(xo + 4 * c0 + 4) = o1; }
        }
    } else {
        for (int f = 0; f < nv; ++f) {
            const int o = (((f >> 2) ^ sw) << 2) + (f & 3);
            const float v = filt_frame<MODE>(ib[o], ib[32 + o], hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            if (wr) xo[f] = v;
        }
    }
}

struct G1S { float det_slow, det_db_m; };
// :221-227 and :262: detector meter, slow follower, adaptive closing coefficient
__device__ __forceinline__ float g1_frame(float lin, float db, G1S& s, float slow_c, float close_fast_c, float close_slow_c)
{
    s.det_db_m = fmaf(0.02f, db - s.det_db_m, s.det_db_m);
    s.det_slow = fmaf(slow_c, lin - s.det_slow, s.det_slow);
    return lin < s.det_slow * 0.92f ? close_fast_c : close_slow_c;
}

struct G2S { float gain_db_s, gr_db_m, thresh_db, thresh_hi; bool expanding; };
// :229-236 hysteresis flag, :258-259 reduction meter, :264-268 switching smoother
__device__ __forceinline__ float g2_frame(float db, float tgv, float close_c, G2S& s, float open_c)
{
    const bool keep = s.expanding && !(db > s.thresh_hi);
    s.expanding = (db < s.thresh_db) || keep;
    const float target_gr = s.expanding ? tgv : 0.0f;
    s.gr_db_m = fmaf(0.05f, target_gr - s.gr_db_m, s.gr_db_m);
    const float target_gain_db = -target_gr;
    const float c = target_gain_db > s.gain_db_s ? open_c : close_c;
    s.gain_db_s = fmaf(c, target_gain_db - s.gain_db_s, s.gain_db_s);
    return s.gain_db_s;
}

}  // namespace

__global__ void __launch_bounds__(kThreads, 1)
easyexpander_v3_kernel(const EEArgs a, const int spc)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    SmemV3& sm = *reinterpret_cast<SmemV3*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int role = kRoleV3[warp];
    if (role == R_NONE) return;
    const int s_base = blockIdx.x * spc;
    const int ns_cta = min(spc, a.n_streams - s_base);
    const size_t ns = (size_t)a.n_streams;
    const int64_t n = a.n_frames;
    const int n_batches = (int)((n + kF - 1) / kF);
    const int n_iter = n_batches + 7;

    if (role < kBandWarpsN) {
        // ============================ band warps: SVF + power follower; kPPL band pairs per lane, 8 kPPL streams per warp ====
        const int sl = lane / kLPS, jj = lane % kLPS;
        const int ls = 8 * kPPL * role + sl;
        const bool act = 8 * kPPL * role < ns_cta, live = ls < ns_cta;
        const int sg = s_base + (live ? ls : (ns_cta - 1));
        const double bc = a.k.band_c;
        double* st = a.state + sg;
        BandK3 bk[kPPL];
        BandS3 bs[kPPL];
        float kA[kPPL], kB[kPPL], wA[kPPL], wB[kPPL];
#pragma unroll
        for (int p = 0; p < kPPL; ++p) {
            const int jA = 2 * (jj * kPPL + p), jB = jA + 1;
            const double gA = a.k.g[jA], gB = a.k.g[jB], dA = a.k.den[jA], dB = a.k.den[jB];
            kA[p] = (float)(2.0 * gA / dA); kB[p] = (float)(2.0 * gB / dB);      // input gain of the unscaled recurrence
            wA[p] = (float)a.k.w[jA]; wB[p] = (float)a.k.w[jB];
            bk[p] = BandK3{pk((float)(-2.0 * gA / dA), (float)(-2.0 * gB / dB)), pk((float)(2.0 / dA), (float)(2.0 / dB)),
                           pk((float)(2.0 / dA - 1.0), (float)(2.0 / dB - 1.0)), pk((float)gA, (float)gB),
                           pk((float)(1.0 - bc), (float)(1.0 - bc)),
                           pk((float)((double)wA[p] * bc * 0.25 * (double)kA[p] * (double)kA[p]), (float)((double)wB[p] * bc * 0.25 * (double)kB[p] * (double)kB[p]))};
            bs[p] = BandS3{pk((float)(st[(0 + jA) * ns] / (double)kA[p]), (float)(st[(0 + jB) * ns] / (double)kB[p])),
                           pk((float)(st[(8 + jA) * ns] / (double)kA[p]), (float)(st[(8 + jB) * ns] / (double)kB[p])),
                           pk((float)(st[(16 + jA) * ns] * (double)wA[p]), (float)(st[(16 + jB) * ns] * (double)wB[p]))};
        }
        for (int i = 0; i < n_iter; ++i) {
            if (act && i >= 2 && i - 2 < n_batches && !(a.role_off & 1)) {
                const int b = i - 2;
                const int lr = live ? ls : 0;                        // rows exist for kMaxS streams only; the others never store
                const float* xd = sm.xd[b & 1] + lr * kRowP;
                const int nv = (int)min((int64_t)kF, n - (int64_t)b * kF);
                float* pa = sm.bp[b & 1] + lr * kBpP + jj * kPPL;
                if (nv == kF) {
#pragma unroll 1
                    for (int f0 = 0; f0 < kF; f0 += 8) band_group3<true>(xd, pa, f0, nv, live, bk, bs);
                } else {
#pragma unroll 1
                    for (int f0 = 0; f0 < kF; f0 += 8) band_group3<false>(xd, pa, f0, nv, live, bk, bs);
                }
            }
            v3_sync();
        }
        if (live) {
#pragma unroll
            for (int p = 0; p < kPPL; ++p) {
                const int jA = 2 * (jj * kPPL + p), jB = jA + 1;
                float lo_, hi_;
                unpk(bs[p].i1, lo_, hi_); st[(0 + jA) * ns] = (double)lo_ * (double)kA[p]; st[(0 + jB) * ns] = (double)hi_ * (double)kB[p];
                unpk(bs[p].i2, lo_, hi_); st[(8 + jA) * ns] = (double)lo_ * (double)kA[p]; st[(8 + jB) * ns] = (double)hi_ * (double)kB[p];
                unpk(bs[p].e, lo_, hi_); st[(16 + jA) * ns] = (double)lo_ / (double)wA[p]; st[(16 + jB) * ns] = (double)hi_ / (double)wB[p];
            }
        }
    } else if (role == R_FILTER) {
        // ============================ filter warp: lane = stream (batch i - 1) ========================================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        float hpf_lp = (float)st[24 * ns], lpf_y = (float)st[25 * ns], hpf_c = 0.f, lpf_c = 0.f;
        bool hpf_on = false, lpf_on = false, any_on = false;
        BlockWalk bw{0, -1, a.block_len, a.n_coef_blocks};
        const EECoef* crow = a.coef + sg;
        double n_hc = crow[0].hpf_c, n_lc = crow[0].lpf_c;                    // the next host block's row, loaded one block ahead
        int n_hon = crow[0].hpf_on, n_lon = crow[0].lpf_on;
        for (int i = 0; i < n_iter; ++i) {
            if (i >= 1 && i - 1 < n_batches && !(a.role_off & 2)) {
                const int b = i - 1;
                if (bw.enter((int64_t)b * kF)) {
                    hpf_c = (float)n_hc; lpf_c = (float)n_lc; hpf_on = n_hon != 0; lpf_on = n_lon != 0;
                    any_on = __any_sync(0xffffffffu, hpf_on || lpf_on);
                    const EECoef& c = crow[(size_t)(bw.row + 1 < bw.n_rows ? bw.row + 1 : bw.n_rows - 1) * ns];
                    n_hc = c.hpf_c; n_lc = c.lpf_c; n_hon = c.hpf_on; n_lon = c.lpf_on;
                }
                const int lr = live ? lane : 0;                      // rows exist for kMaxS streams only; the others never store
                const float* ib = sm.io[b & (kIoSlots - 1)] + lr * kIoP;
                float* xo = sm.xd[b & 1] + lr * kRowP;
                const int nv = (int)min((int64_t)kF, n - (int64_t)b * kF);
                if (any_on) filt_batch<1>(ib, lr & 7, xo, nv, live, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
                else filt_batch<0>(ib, lr & 7, xo, nv, live, hpf_lp, lpf_y, hpf_c, lpf_c, hpf_on, lpf_on);
            }
            v3_sync();
        }
        if (live) { st[24 * ns] = (double)hpf_lp; st[25 * ns] = (double)lpf_y; }
    } else if (role == R_G1) {
        // ============================ G1: lane = stream (batch i - 4): followers + closing coefficient ===================
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        G1S gs{(float)st[26 * ns], (float)st[29 * ns]};
        const float slow_c = (float)a.k.slow_c, cfast = (float)a.k.close_fast_c, cslow = (float)a.k.close_slow_c;
        for (int i = 0; i < n_iter; ++i) {
            if (i >= 4 && i - 4 < n_batches && !(a.role_off & 4)) {
                const int b = i - 4;
                const int lr = live ? lane : 0;
                const float* lin = sm.lin[b & 1] + lr * kRowP;
                const float* db = sm.db[b % kDbSlots] + lr * kRowP;
                float* cc = sm.cc[b & 1] + lr * kRowP;
                const int nv = (int)min((int64_t)kF, n - (int64_t)b * kF);
                if (nv == kF) {
#pragma unroll 1
                    for (int f0 = 0; f0 < kF; f0 += 8) {
                        const float4 l0 = *reinterpret_cast<const float4*>(lin + f0), l1 = *reinterpret_cast<const float4*>(lin + f0 + 4);
                        const float4 d0 = *reinterpret_cast<const float4*>(db + f0), d1 = *reinterpret_cast<const float4*>(db + f0 + 4);
                        float4 o0, o1;
                        o0.x = g1_frame(l0.x, d0.x, gs, slow_c, cfast, cslow); o0.y = g1_frame(l0.y, d0.y, gs, slow_c, cfast, cslow);
                        o0.z = g1_frame(l0.z, d0.z, gs, slow_c, cfast, cslow); o0.w = g1_frame(l0.w, d0.w, gs, slow_c, cfast, cslow);
                        o1.x = g1_frame(l1.x, d1.x, gs, slow_c, cfast, cslow); o1.y = g1_frame(l1.y, d1.y, gs, slow_c, cfast, cslow);
                        o1.z = g1_frame(l1.z, d1.z, gs, slow_c, cfast, cslow); o1.w = g1_frame(l1.w, d1.w, gs, slow_c, cfast, cslow);
                        if (live) { *reinterpret_cast<float4*>(cc + f0) = o0; *reinterpret_cast<float4*>(cc + f0 + 4) = o1; }
                    }
                } else {
                    for (int f = 0; f < nv; ++f) { const float v = g1_frame(lin[f], db[f], gs, slow_c, cfast, cslow); if (live) cc[f] = v; }
                }
            }
            v3_sync();
        }
        if (live) { st[26 * ns] = (double)gs.det_slow; st[29 * ns] = (double)gs.det_db_m; }
    } else if (role == R_G2) {
        // ============================ G2: lane = stream (batch i - 5): hysteresis flag + switching smoother ==============
        const bool live = lane < ns_cta;
        const int sg = s_base + (live ? lane : 0);
        double* st = a.state + sg;
        G2S gs{(float)st[28 * ns], (float)st[30 * ns], 0.f, 0.f, st[27 * ns] != 0.0};
        const float open_c = (float)a.k.open_c;
        BlockWalk bw{0, -1, a.block_len, a.n_coef_blocks};
        const EECoef* crow = a.coef + sg;
        double n_thr = crow[0].thresh_db;                                     // the next host block's threshold, loaded one block ahead
        for (int i = 0; i < n_iter; ++i) {
            if (i >= 5 && i - 5 < n_batches && !(a.role_off & 8)) {
                const int b = i - 5;
                if (bw.enter((int64_t)b * kF)) {
                    gs.thresh_db = (float)n_thr;
                    n_thr = crow[(size_t)(bw.row + 1 < bw.n_rows ? bw.row + 1 : bw.n_rows - 1) * ns].thresh_db;
                    gs.thresh_hi = gs.thresh_db + 2.0f;                                   // hys_db (:33)
                }
                const int lr = live ? lane : 0;
                const float* db = sm.db[b % kDbSlots] + lr * kRowP;
                const float* tg = sm.tg[b % kDbSlots] + lr * kRowP;
                const float* cc = sm.cc[b & 1] + lr * kRowP;
                float* go = sm.gain[b & 1] + lr * kRowP;
                const int nv = (int)min((int64_t)kF, n - (int64_t)b * kF);
                if (nv == kF) {
#pragma unroll 1
                    for (int f0 = 0; f0 < kF; f0 += 8) {
                        const float4 d0 = *reinterpret_cast<const float4*>(db + f0), d1 = *reinterpret_cast<const float4*>(db + f0 + 4);
                        const float4 t0 = *reinterpret_cast<const float4*>(tg + f0), t1 = *reinterpret_cast<const float4*>(tg + f0 + 4);
                        const float4 c0 = *reinterpret_cast<const float4*>(cc + f0), c1 = *reinterpret_cast<const float4*>(cc + f0 + 4);
                        float4 o0, o1;
                        o0.x = g2_frame(d0.x, t0.x, c0.x, gs, open_c); o0.y = g2_frame(d0.y, t0.y, c0.y, gs, open_c);
                        o0.z = g2_frame(d0.z, t0.z, c0.z, gs, open_c); o0.w = g2_frame(d0.w, t0.w, c0.w, gs, open_c);
                        o1.x = g2_frame(d1.x, t1.x, c1.x, gs, open_c); o1.y = g2_frame(d1.y, t1.y, c1.y, gs, open_c);
                        o1.z = g2_frame(d1.z, t1.z, c1.z, gs, open_c); o1.w = g2_frame(d1.w, t1.w, c1.w, gs, open_c);
                        if (live) { *reinterpret_cast<float4*>(go + f0) = o0; *reinterpret_cast<float4*>(go + f0 + 4) = o1; }
                    }
                } else {
                    for (int f = 0; f < nv; ++f) { const float v = g2_frame(db[f], tg[f], cc[f], gs, open_c); if (live) go[f] = v; }
                }
            }
            v3_sync();
        }
        if (live) {
            st[27 * ns] = gs.expanding ? 1.0 : 0.0; st[28 * ns] = (double)gs.gain_db_s; st[30 * ns] = (double)gs.gr_db_m;
        }
    } else {
        // ============================ helper warps: lane = frame; streams kQ h .. kQ h + kQ - 1 of the CTA ===============
        // Global memory is touched in 512-byte runs per row: a super-batch (4 batches = 128 frames) of a helper's rows is
        // fetched with one 16-byte cp.async per lane and row into the I/O ring kLead iterations before its first batch is
        // due; S4 overwrites the ring slot in place with the output, and once the super-batch's last batch is through S4
        // the same warp stores it with one 128-bit store per lane and row.
        const int h = role - R_HELP;
        const bool act = kQ * h < ns_cta;
        const float* irow[kQ]; float* orow[kQ];
        const EECoef* crow[kQ];
        bool has[kQ];
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int lq = kQ * h + q;
            has[q] = lq < ns_cta;
            const size_t s = (size_t)(s_base + (has[q] ? lq : 0));
            irow[q] = a.in + s * 2 * a.row_stride;
            orow[q] = a.out + s * 2 * a.row_stride;
            crow[q] = a.coef + s;
        }
        const int64_t rs = a.row_stride;
        const int bpA = lane * 4;                                            // the four pair sums of this lane's frame
        // ring offsets of this lane: as frame `lane` of a batch (S4), and as 16-byte chunk `lane` of a super-batch (I/O)
        int fo[kQ], co[kQ];
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int lq = act ? (has[q] ? kQ * h + q : kQ * h) : 0;
            fo[q] = lq * kIoP + ((((lane >> 2) ^ (lq & 7)) << 2) | (lane & 3));
            co[q] = (lane >> 3) * (kMaxS * kIoP) + lq * kIoP + (((lane & 7) ^ (lq & 7)) << 2);
        }
        float thr[kQ], dep[kQ], rm1[kQ];
        double nthr[kQ], ndep[kQ], nrat[kQ];                                 // the next host block's row, loaded one block ahead
        BlockWalk bw{0, -1, a.block_len, a.n_coef_blocks};
        const float eps = 1e-12f;
        const float db_per_lg2_pow = 3.0102999566398120f;                    // 10 log10(2): det_db = 10 log10(sum) = 20 log10(sqrt(sum))
        const float lg2_per_db = 0.16609640474436813f;                       // ln(10)/20 / ln(2)
        const int n_sb = (n_batches + kSB - 1) / kSB;
        // cp.async of super-batch k (frames 128 k + 4 lane .. + 3 of every row); bytes beyond the call are zero-filled
        auto fetch = [&](const int k) {
            if (act && k < n_sb && !(a.role_off & 32)) {
                const int64_t f0 = (int64_t)k * (kSB * kF) + 4 * lane;
                const int64_t left = n - f0;
                const int bytes = left >= 4 ? 16 : (left > 0 ? 4 * (int)left : 0);
                const int64_t fsrc = bytes ? f0 : 0;
                float* ring = sm.io[(k * kSB) & (kIoSlots - 1)];
#pragma unroll
                for (int q = 0; q < kQ; ++q) {
                    if (has[q]) {
#pragma unroll
                        for (int ch = 0; ch < 2; ++ch)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(ring + co[q] + 32 * ch)),
                                         "l"(irow[q] + ch * rs + fsrc), "r"(bytes) : "memory");
                    }
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            thr[q] = dep[q] = rm1[q] = 0.f;
            nthr[q] = crow[q][0].thresh_db; ndep[q] = crow[q][0].depth_db; nrat[q] = crow[q][0].ratio_base;
        }
        fetch(0); fetch(1);
        // P2 of one batch: band sum in the script's order, level, memoryless half of the gain computer
        auto p2_math = [&](const float4 pr, const int q, float& det_lin, float& det_db, float& target_gr) {
            const float sum = fmaxf(((pr.x + pr.y) + pr.z) + pr.w, eps);
            det_lin = sum * rsqrt_fast(sum);
            det_db = db_per_lg2_pow * lg2_fast(sum);
            float d = thr[q] - det_db;
            d = d < 0.f ? 0.f : d;
            const float xk = d * (1.0f / 6.0f);                              // knee_db = 6 (:34)
            const float smo = (xk * xk) * fmaf(-2.0f, xk, 3.0f);
            const float re1 = d < 6.0f ? rm1[q] * smo : rm1[q];               // ratio_eff - 1
            target_gr = d * re1;
            target_gr = target_gr > dep[q] ? dep[q] : target_gr;
        };
        auto next_block_rows = [&]() {                                        // entering a host block: its row was fetched a block ago
            const int nr = bw.row + 1 < bw.n_rows ? bw.row + 1 : bw.n_rows - 1;
#pragma unroll
            for (int q = 0; q < kQ; ++q) {
                thr[q] = (float)nthr[q]; dep[q] = (float)ndep[q]; rm1[q] = (float)(nrat[q] - 1.0);
                const EECoef& c = crow[q][(size_t)nr * ns];
                nthr[q] = c.thresh_db; ndep[q] = c.depth_db; nrat[q] = c.ratio_base;
            }
        };
        // one iteration, any position in the render (pipeline fill and drain, ragged last batch, partial helpers)
        auto body_any = [&](const int i) {
            if ((i & (kSB - 1)) == 0) asm volatile("cp.async.wait_group 1;" ::: "memory");
            if (((i + kLead) & (kSB - 1)) == 0) fetch((i + kLead) / kSB);
            if (act) {
                if (i >= 3 && i - 3 < n_batches && !(a.role_off & 16)) {
                    const int b = i - 3;
                    if (bw.enter((int64_t)b * kF)) next_block_rows();
                    const float* bp = sm.bp[b & 1] + (kQ * h) * kBpP;
                    float* lin = sm.lin[b & 1] + (kQ * h) * kRowP + lane;
                    float* db = sm.db[b % kDbSlots] + (kQ * h) * kRowP + lane;
                    float* tg = sm.tg[b % kDbSlots] + (kQ * h) * kRowP + lane;
#pragma unroll
                    for (int q = 0; q < kQ; ++q) {
                        float det_lin, det_db, target_gr;
                        p2_math(*reinterpret_cast<const float4*>(bp + (has[q] ? q : 0) * kBpP + bpA), q, det_lin, det_db, target_gr);
                        if (has[q]) { lin[q * kRowP] = det_lin; db[q * kRowP] = det_db; tg[q * kRowP] = target_gr; }
                    }
                }
                if (i >= 6 && i - 6 < n_batches && !(a.role_off & 32)) {
                    const int b = i - 6;
                    const float* io = sm.io[b & (kIoSlots - 1)];
                    const float* gp = sm.gain[b & 1] + (kQ * h) * kRowP + lane;
                    const int64_t f = (int64_t)b * kF + lane;
#pragma unroll
                    for (int q = 0; q < kQ; ++q) {
                        const float gain_lin = ex2_fast(gp[(has[q] ? q : 0) * kRowP] * lg2_per_db);
                        const float o0 = io[fo[q]] * gain_lin, o1 = io[fo[q] + 32] * gain_lin;
                        if (has[q] && f < n) { orow[q][f] = o0; orow[q][rs + f] = o1; }
                    }
                }
            }
            v3_sync();
        };
        // Steady state (6 <= i < n_batches + 3, a helper with all its streams, no measurement switches): every stage is due,
        // every S4 batch is whole, so the iteration carries no validity checks; host blocks are counted in batches.
        bool all = act && a.role_off == 0;
#pragma unroll
        for (int q = 0; q < kQ; ++q) all = all && has[q];
        int i = 0;
        const int i_steady = all ? 6 : n_iter, i_drain = all ? n_batches + 3 : n_iter;
        for (; i < i_steady && i < n_iter; ++i) body_any(i);
        if (i < i_drain) {
            const int bpb = a.block_len / kF;                                 // batches per host block
            // batch i - 3 of P2: batches left in its host block (bw is positioned by the last body_any call)
            int left = (int)((bw.next - (int64_t)(i - 3) * kF) / kF);
            const float* bpq = sm.bp[0] + (kQ * h) * kBpP + bpA;
            float* linq = sm.lin[0] + (kQ * h) * kRowP + lane;
            float* dbq = sm.db[0] + (kQ * h) * kRowP + lane;
            float* tgq = sm.tg[0] + (kQ * h) * kRowP + lane;
            const float* gq = sm.gain[0] + (kQ * h) * kRowP + lane;
            int slot3 = (i - 3) % kDbSlots;
            float* op[kQ];
#pragma unroll
            for (int q = 0; q < kQ; ++q) op[q] = orow[q] + (int64_t)(i - 6) * kF + lane;
            for (; i < i_drain; ++i) {
                if ((i & (kSB - 1)) == 0) asm volatile("cp.async.wait_group 1;" ::: "memory");
                if (((i + kLead) & (kSB - 1)) == 0) fetch((i + kLead) / kSB);
                if (left == 0) { bw.row += 1; bw.next += a.block_len; next_block_rows(); left = bpb; }
                --left;
                const int par3 = (i - 3) & 1, par6 = i & 1;                      // (i - 6) & 1 == i & 1
                const float* io = sm.io[(i - 6) & (kIoSlots - 1)];
#pragma unroll
                for (int q = 0; q < kQ; ++q) {
                    float det_lin, det_db, target_gr;
                    p2_math(*reinterpret_cast<const float4*>(bpq + par3 * (kMaxS * kBpP) + q * kBpP), q, det_lin, det_db, target_gr);
                    linq[par3 * (kMaxS * kRowP) + q * kRowP] = det_lin;
                    dbq[slot3 * (kMaxS * kRowP) + q * kRowP] = det_db;
                    tgq[slot3 * (kMaxS * kRowP) + q * kRowP] = target_gr;
                }
#pragma unroll
                for (int q = 0; q < kQ; ++q) {
                    const float gain_lin = ex2_fast(gq[par6 * (kMaxS * kRowP) + q * kRowP] * lg2_per_db);
                    op[q][0] = io[fo[q]] * gain_lin;
                    op[q][rs] = io[fo[q] + 32] * gain_lin;
                    op[q] += kF;
                }
                slot3 = slot3 == kDbSlots - 1 ? 0 : slot3 + 1;
                v3_sync();
            }
            // hand the block position back to the general body
            // (bw.row / bw.next were advanced block by block above; batches of the current block already consumed do not matter
            //  to BlockWalk::enter, which only compares a batch's first frame with bw.next)
        }
        for (; i < n_iter; ++i) body_any(i);
    }
}

bool easyexpander_v3_eligible(const EEArgs& a)
{
    if (a.block_len < kF || a.block_len % kF != 0) return false;
    // 16-byte cp.async / 128-bit stores per row: rows have to start on 16-byte boundaries
    if (a.row_stride % 4 != 0 || ((uintptr_t)a.in & 15) != 0 || ((uintptr_t)a.out & 15) != 0) return false;
    for (int j = 0; j < 8; ++j) {
        const float w = (float)a.k.w[j], k = (float)(2.0 * a.k.g[j] / a.k.den[j]);
        if (!(w > 1e-30f) || !(k > 1e-30f)) return false;                  // the scaled band states divide by these
    }
    return true;
}

bool easyexpander_v3_launch(const EEArgs& a, int n_sm, cudaStream_t stream)
{
    const size_t smem = sizeof(SmemV3);
    static DeviceOnce attr_once;
    if (attr_once.pending()) {
        if (!cuda_ok(cudaFuncSetAttribute(easyexpander_v3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "EasyExpander v3 smem attr")) return false;
        attr_once.mark();
    }
    int spc = (a.n_streams + n_sm - 1) / n_sm;
    spc = (spc + 3) / 4 * 4;
    if (spc > kMaxS) spc = kMaxS;
    if (spc < 4) spc = 4;
    const int blocks = (a.n_streams + spc - 1) / spc;
    easyexpander_v3_kernel<<<blocks, kThreads, smem, stream>>>(a, spc);
    return cuda_ok(cudaGetLastError(), "easyexpander_v3_kernel launch");
}

}  // namespace za
