// k_cascade8_x2.cu -- linear IIR class, fp32 arithmetic, packed-FP32 (FFMA2) edition for sm_100a.
//
// Same job as k_cascade8.cu (plugin "cascade8": the generated per-sample loop of
// dsp_jsfx_aot.py:5713-5905 for n_streams instances), restructured around what ncu showed for v1
// (profiles/r01_cascade8_v1_ncu.md): DRAM traffic == algorithmic bytes, the kernel is bound by
// instruction count.  Blackwell's packed `fma.rn.f32x2` (SASS FFMA2) retires two FMAs per issue
// slot with the same latency as FFMA, so:
//
//   * every thread owns 64 consecutive frames = rows 2t and 2t+1 of the tile = FOUR 16-frame chunks
//     (A, B | C, D) and keeps them as 32 register PAIRS (A[k], B[k]), (C[k], D[k]); pass A (zero-
//     state FIR) and pass B (the recurrence) run on the pairs, one FFMA2 per formerly two FFMAs,
//     as two independent chains (AB, CD) so a single warp can keep the FMA pipe fed.  The pairs
//     are carried as 64-bit values (inline-PTX fma.rn.f32x2 on .b64 registers) so ptxas keeps them
//     in aligned register pairs across the run-time stage loops; CTA-uniform coefficients use
//     FFMA2's scalar-broadcast operand form (`R.F32`), negated addends fold into its operand
//     modifier, and pass B is arranged so that every read of x[k] precedes y (z2' = (b2-b0) x - z1
//     + d2 y), which lets y overwrite x[k] in place: the stage loops hold no register copies;
//   * the warp-shuffle scan runs once per 64 frames (the thread first folds its four chunks);
//   * pass A of stage s+1 is fused into the pass-B loop of stage s, and the six biquads run as
//     run-time loops over two loop bodies (plain / band-pass+mix) so the hot code stays inside the
//     instruction cache (v2a, fully unrolled, lost 12 % of its issue slots to `no_instruction`);
//   * one CTA = one stream; each WARP moves its own 64-row share of the tile: ONE TMA tensor load
//     into the warp's smem rows, compute in registers, results written back to the SAME rows
//     (a thread only touches its own two rows), one TMA tensor store.  The next tile is pulled
//     into L2 by a TMA prefetch while this one computes, so the exposed load is an L2 hit that
//     the other resident warps cover.  smem per CTA = 16 KB (W = 2) + 6 KB tables, so 8 CTAs
//     (16 warps at 128 registers) share an SM.
//
// What bounds it (profiles/r01_cascade8_x2_ncu.md): 56 FMAs per frame are irreducible for this
// scan formulation (2 pass A + 6 pass B per biquad frame); an FFMA2 holds the sub-partition's
// dispatch port for 2 cycles, as does every 16-lane ALU-pipe instruction (tools/ubench_issue.cu),
// so the kernel is dispatch-bound, not HBM-bound: DRAM traffic equals the algorithmic bytes.
//
// Arithmetic notes (fp32 mode only has to null below -100 dBFS against the fp64 oracle):
//   * feedback coefficients are delta-coded, a1 = -2 + d1, a2 = 1 - d2, as in v1;
//   * both one-poles run as s += c (x - s) (exact DC gain in float);
//   * the scan runs in the real normal form of each biquad (|N^n| <= 1), tables built in double.
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

constexpr int kH = 16;   // frames per thread-chunk of the packed kernel: 4 chunks (A, B | C, D) = 2 rows of a tile

struct alignas(16) C8TablesX2 {
    float HA[7][kH][2];     // stages 0-5, biquads: (h1k, h2k), h_k = T A^(15-k) B (zero-state end state of a
                            // 16-frame chunk, scan coords); stage 6 = one-pole S7 in the same row format: (h_k, 0)
    float HB[kH];           // one-pole S8: h_k
    float P2[6][36][4];     // (N^64)^j for j = 0..32 (row-major 2x2); [33] = N^16, [34] = N^32
    float T2[6][8];         // T (4), T^-1 (4)
    float P1[2][36];        // (alpha^64)^j, j = 0..32; [33] = alpha^16, [34] = alpha^32
    float coef[40];         // per biquad k: b0 b1 (b2-b0) -d1 d2 at [5k..5k+4]; [30] = c7, [31] = c8, [32] = mix
};

size_t cascade8_x2_tables_bytes() { return sizeof(C8TablesX2); }

// one thread per (stream, section); double arithmetic, stored as float
__global__ void cascade8_x2_tables_kernel(const double* __restrict__ coef, C8TablesX2* __restrict__ tabs, int n_streams)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int s = g >> 3, sec = g & 7;
    if (s >= n_streams) return;
    const double* c = coef + (size_t)s * kC8Coefs;
    C8TablesX2& t = tabs[s];
    if (sec < 6) {
        const double b0 = c[5 * sec], b1 = c[5 * sec + 1], b2 = c[5 * sec + 2], a1 = c[5 * sec + 3], a2 = c[5 * sec + 4];
        // s' = A s + B x with A = [[-a1, 1], [-a2, 0]]; scan coordinates w = T s, T A T^-1 = N (scaled rotation)
        const double sg = -0.5 * a1, w2 = a2 - sg * sg;
        double t00 = 1, t01 = 0, t10 = 0, t11 = 1, i00 = 1, i01 = 0, i10 = 0, i11 = 1;
        double n00 = -a1, n01 = 1, n10 = -a2, n11 = 0;
        if (w2 > 1e-14 && a2 > 1e-14) {
            const double om = sqrt(w2);
            t00 = 0; t01 = -1.0 / a2; t10 = 1.0 / om; t11 = sg / (om * a2);
            i00 = sg; i01 = om; i10 = -a2; i11 = 0;
            n00 = sg; n01 = om; n10 = -om; n11 = sg;
        }
        const double bb1 = b1 - a1 * b0, bb2 = b2 - a2 * b0;
        double v1 = t00 * bb1 + t01 * bb2, v2 = t10 * bb1 + t11 * bb2;
        for (int k = kH - 1; k >= 0; --k) {
            t.HA[sec][k][0] = (float)v1; t.HA[sec][k][1] = (float)v2;
            const double x1 = n00 * v1 + n01 * v2, x2 = n10 * v1 + n11 * v2;
            v1 = x1; v2 = x2;
        }
        double m00 = 1, m01 = 0, m10 = 0, m11 = 1;       // N^16, then squared to N^32 and N^64
        for (int k = 0; k < kH; ++k) {
            const double x00 = n00 * m00 + n01 * m10, x01 = n00 * m01 + n01 * m11;
            const double x10 = n10 * m00 + n11 * m10, x11 = n10 * m01 + n11 * m11;
            m00 = x00; m01 = x01; m10 = x10; m11 = x11;
        }
        for (int q = 33; q <= 35; ++q) {
            if (q < 35) { t.P2[sec][q][0] = (float)m00; t.P2[sec][q][1] = (float)m01; t.P2[sec][q][2] = (float)m10; t.P2[sec][q][3] = (float)m11; }
            else        { t.P2[sec][q][0] = 0; t.P2[sec][q][1] = 0; t.P2[sec][q][2] = 0; t.P2[sec][q][3] = 0; break; }
            const double x00 = m00 * m00 + m01 * m10, x01 = m00 * m01 + m01 * m11;
            const double x10 = m10 * m00 + m11 * m10, x11 = m10 * m01 + m11 * m11;
            m00 = x00; m01 = x01; m10 = x10; m11 = x11;
        }
        double p00 = 1, p01 = 0, p10 = 0, p11 = 1;
        for (int j = 0; j <= 32; ++j) {
            t.P2[sec][j][0] = (float)p00; t.P2[sec][j][1] = (float)p01; t.P2[sec][j][2] = (float)p10; t.P2[sec][j][3] = (float)p11;
            const double x00 = m00 * p00 + m01 * p10, x01 = m00 * p01 + m01 * p11;
            const double x10 = m10 * p00 + m11 * p10, x11 = m10 * p01 + m11 * p11;
            p00 = x00; p01 = x01; p10 = x10; p11 = x11;
        }
        t.T2[sec][0] = (float)t00; t.T2[sec][1] = (float)t01; t.T2[sec][2] = (float)t10; t.T2[sec][3] = (float)t11;
        t.T2[sec][4] = (float)i00; t.T2[sec][5] = (float)i01; t.T2[sec][6] = (float)i10; t.T2[sec][7] = (float)i11;
        t.coef[5 * sec + 0] = (float)b0; t.coef[5 * sec + 1] = (float)b1; t.coef[5 * sec + 2] = (float)(b2 - b0);
        t.coef[5 * sec + 3] = (float)(-(a1 + 2.0));      // -d1
        t.coef[5 * sec + 4] = (float)(1.0 - a2);         //  d2
    } else {
        const int o = sec - 6;
        // S7: lp = (1-a7) x + a7 lp ; S8: lp += c8 (x - lp); both are s' = (1-c) s + c x
        const double cc = o == 0 ? c[31] : c[32];
        const double alpha = o == 0 ? c[30] : 1.0 - c[32];
        double v = cc;
        for (int k = kH - 1; k >= 0; --k) {
            if (o == 0) { t.HA[6][k][0] = (float)v; t.HA[6][k][1] = 0; }
            else        { t.HB[k] = (float)v; }
            v *= alpha;
        }
        double m = 1;
        for (int k = 0; k < kH; ++k) m *= alpha;
        t.P1[o][33] = (float)m;
        m *= m;
        t.P1[o][34] = (float)m;
        m *= m;
        double p = 1;
        for (int j = 0; j <= 32; ++j) { t.P1[o][j] = (float)p; p *= m; }
        t.P1[o][35] = 0;
        t.coef[30 + o] = (float)cc;
        if (o == 0) { t.coef[32] = (float)c[33]; for (int i = 33; i < 40; ++i) t.coef[i] = 0; }
    }
}

bool cascade8_x2_build_tables(const double* d_coef, void* d_tables, int n_streams, cudaStream_t s)
{
    const int threads = 128, blocks = (n_streams * 8 + threads - 1) / threads;
    cascade8_x2_tables_kernel<<<blocks, threads, 0, s>>>(d_coef, (C8TablesX2*)d_tables, n_streams);
    return cuda_ok(cudaGetLastError(), "cascade8_x2_tables_kernel");
}


// ---------------------------------------------------------------------------------------------
// device side
// ---------------------------------------------------------------------------------------------
// A pair of floats in one 64-bit register (lo = chunk A, hi = chunk B).
typedef unsigned long long p2;
__device__ __forceinline__ p2 pk(float lo, float hi) { p2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo(p2 a) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a)); (void)h; return l; }
__device__ __forceinline__ float hi(p2 a) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a)); (void)l; return h; }
// a*b + c on both halves
__device__ __forceinline__ p2 fma2(p2 a, p2 b, p2 c) { p2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
// s*b + c, scalar s broadcast to both halves (SASS: FFMA2 Rd, Rs.F32, Rb.F32x2, Rc.F32x2)
__device__ __forceinline__ p2 fma2s(float s, p2 b, p2 c)
{
    p2 r;
    asm("{.reg .b64 t; mov.b64 t, {%1, %1}; fma.rn.f32x2 %0, t, %2, %3;}" : "=l"(r) : "f"(s), "l"(b), "l"(c));
    return r;
}
// s*b - c (the negation folds into FFMA2's operand modifier)
__device__ __forceinline__ p2 fma2s_sub(float s, p2 b, p2 c)
{
    p2 r;
    asm("{.reg .b64 t, u; .reg .f32 l, h; mov.b64 t, {%1, %1}; mov.b64 {l, h}, %3; neg.f32 l, l; neg.f32 h, h;"
        " mov.b64 u, {l, h}; fma.rn.f32x2 %0, t, %2, u;}" : "=l"(r) : "f"(s), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ p2 add2(p2 a, p2 b) { p2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ p2 sub2(p2 a, p2 b) { p2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float4 lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }

template <int W, int NBUF> struct X2Smem {
    alignas(1024) float tile[NBUF][64 * W * kL]; // 2T rows x 32 frames per buffer, SWIZZLE_128B; warp w owns rows [64w, 64w+64)
    alignas(16) C8TablesX2 tab;
    alignas(16) float carry[2][kC8State];        // tile-parity double buffer (W > 1 only)
    alignas(16) float wtot[8][W][2];             // per-stage warp totals (W > 1 only)
    alignas(8) uint64_t full[NBUF][W];           // one mbarrier per warp and buffer
};

template <int W, int NBUF> struct X2Ctx {
    X2Smem<W, NBUF>* sm;
    int tid, lane, warp;
    int t_last, half_last;  // owner thread of the tile's last valid chunk and which of its two chunks that is
    int par;
    float c[kC8State];      // W == 1: carry lives in registers
};

// Register layout of a thread's 64 frames (rows 2t and 2t+1 of the tile, 4 chunks of 16 in time order
// A, B | C, D):  x[k] = (A[k], B[k]),  x[16 + k] = (C[k], D[k]),  k = 0..15.  Every pass runs two
// independent FFMA2 chains (AB and CD) so one warp alone can keep the FMA pipe busy.
struct ZS { p2 ab1, ab2, cd1, cd2; };     // zero-state end states: ab1 = (z1 of A, z1 of B), ...

// pass A of stage 0 (no previous stage to fuse it into)
__device__ __forceinline__ void pass_a0(const C8TablesX2& tb, const p2 (&x)[kL], ZS& z)
{
    const p2 zero = pk(0.f, 0.f);
    z.ab1 = zero; z.ab2 = zero; z.cd1 = zero; z.cd2 = zero;
#pragma unroll
    for (int k = 0; k < kH; k += 2) {
        const float4 h = lds4(&tb.HA[0][k][0]);        // (h1[k], h2[k], h1[k+1], h2[k+1])
        z.ab1 = fma2s(h.x, x[k], z.ab1); z.ab2 = fma2s(h.y, x[k], z.ab2);
        z.cd1 = fma2s(h.x, x[kH + k], z.cd1); z.cd2 = fma2s(h.y, x[kH + k], z.cd2);
        z.ab1 = fma2s(h.z, x[k + 1], z.ab1); z.ab2 = fma2s(h.w, x[k + 1], z.ab2);
        z.cd1 = fma2s(h.z, x[kH + k + 1], z.cd1); z.cd2 = fma2s(h.w, x[kH + k + 1], z.cd2);
    }
}

// ---- one biquad stage over the four chunks of the thread -------------------------------------------
// In: z = zero-state end states of THIS stage; out: the same for stage sec+1, accumulated while pass B runs.
// BP: band-pass stages S3/S4 (b1 == 0 exactly, output mixed onto the through signal).
template <int W, int NBUF, bool BP>
__device__ __forceinline__ void biquad_stage(X2Ctx<W, NBUF>& cx, p2 (&x)[kL], const int sec, ZS& z)
{
    const C8TablesX2& tb = cx.sm->tab;
    const float zA1 = lo(z.ab1), zB1 = hi(z.ab1), zA2 = lo(z.ab2), zB2 = hi(z.ab2);
    const float zC1 = lo(z.cd1), zD1 = hi(z.cd1), zC2 = lo(z.cd2), zD2 = hi(z.cd2);
    const float4 n16 = lds4(&tb.P2[sec][33][0]);
    const float4 n32 = lds4(&tb.P2[sec][34][0]);
    const p2 zero = pk(0.f, 0.f);
    // fold the chunks: (cAB, cCD) = N^16 (zA, zC) + (zB, zD), then c = N^32 cAB + cCD (64-frame zero-state end state)
    const p2 zAC1 = pk(zA1, zC1), zAC2 = pk(zA2, zC2);
    const p2 f1 = fma2s(n16.x, zAC1, fma2s(n16.y, zAC2, pk(zB1, zD1)));
    const p2 f2 = fma2s(n16.z, zAC1, fma2s(n16.w, zAC2, pk(zB2, zD2)));
    const float cAB1 = lo(f1), cCD1 = hi(f1), cAB2 = lo(f2), cCD2 = hi(f2);
    float c1 = fmaf(n32.x, cAB1, fmaf(n32.y, cAB2, cCD1));
    float c2 = fmaf(n32.z, cAB1, fmaf(n32.w, cAB2, cCD2));
    // inclusive Kogge-Stone scan of w -> M w + c over the warp's 32 threads, M = N^64
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float u1 = shfl_up(c1, d), u2 = shfl_up(c2, d);
        const float4 m = lds4(&tb.P2[sec][d][0]);
        if (cx.lane >= d) {
            c1 = fmaf(m.x, u1, fmaf(m.y, u2, c1));
            c2 = fmaf(m.z, u1, fmaf(m.w, u2, c2));
        }
    }
    // state entering my warp: exact sequential state (z1, z2) -> scan coordinates
    const float4 tf = lds4(&tb.T2[sec][0]);
    const float4 ti = lds4(&tb.T2[sec][4]);
    float s1, s2;
    if (W == 1) {
        float a = 0, b = 0;
#pragma unroll
        for (int q = 0; q < 6; ++q)                   // sec is a run-time value: select from the register file
            if (q == sec) { a = cx.c[2 * q]; b = cx.c[2 * q + 1]; }
        s1 = fmaf(tf.x, a, tf.y * b); s2 = fmaf(tf.z, a, tf.w * b);
    } else {
        if (cx.lane == 31) { cx.sm->wtot[sec][cx.warp][0] = c1; cx.sm->wtot[sec][cx.warp][1] = c2; }
        __syncthreads();
        const float a = cx.sm->carry[cx.par][2 * sec], b = cx.sm->carry[cx.par][2 * sec + 1];
        s1 = fmaf(tf.x, a, tf.y * b); s2 = fmaf(tf.z, a, tf.w * b);
        const float4 m = lds4(&tb.P2[sec][32][0]);
        for (int w = 0; w < cx.warp; ++w) {
            const float t1 = fmaf(m.x, s1, fmaf(m.y, s2, cx.sm->wtot[sec][w][0]));
            const float t2 = fmaf(m.z, s1, fmaf(m.w, s2, cx.sm->wtot[sec][w][1]));
            s1 = t1; s2 = t2;
        }
    }
    // chunk start states in scan coordinates: wA = exclusive prefix + M^lane * warp-incoming state,
    // wC = N^32 wA + cAB, (wB, wD) = N^16 (wA, wC) + (zA, zC)
    float e1 = shfl_up(c1, 1), e2 = shfl_up(c2, 1);
    if (cx.lane == 0) { e1 = 0; e2 = 0; }
    const float4 pl = lds4(&tb.P2[sec][cx.lane][0]);
    const float wA1 = fmaf(pl.x, s1, fmaf(pl.y, s2, e1));
    const float wA2 = fmaf(pl.z, s1, fmaf(pl.w, s2, e2));
    const float wC1 = fmaf(n32.x, wA1, fmaf(n32.y, wA2, cAB1));
    const float wC2 = fmaf(n32.z, wA1, fmaf(n32.w, wA2, cAB2));
    const p2 wAC1 = pk(wA1, wC1), wAC2 = pk(wA2, wC2);
    const p2 wBD1 = fma2s(n16.x, wAC1, fma2s(n16.y, wAC2, zAC1));
    const p2 wBD2 = fma2s(n16.z, wAC1, fma2s(n16.w, wAC2, zAC2));
    // back to the recurrence's own coordinates (z1, z2) = T^-1 w, as (A, B) and (C, D) pairs
    const p2 wab1 = pk(wA1, lo(wBD1)), wab2 = pk(wA2, lo(wBD2)), wcd1 = pk(wC1, hi(wBD1)), wcd2 = pk(wC2, hi(wBD2));
    p2 sa1 = fma2s(ti.x, wab1, fma2s(ti.y, wab2, zero)), sa2 = fma2s(ti.z, wab1, fma2s(ti.w, wab2, zero));
    p2 sc1 = fma2s(ti.x, wcd1, fma2s(ti.y, wcd2, zero)), sc2 = fma2s(ti.z, wcd1, fma2s(ti.w, wcd2, zero));
    // pass B: the TDF-II recurrence (Spectral Stabilizer.jsfx:98-111), chains AB and CD in lockstep,
    // with pass A of the next stage riding along on the fresh outputs
    const float* cf = &tb.coef[5 * sec];
    const float b0 = cf[0], b1 = cf[1], b20 = cf[2], nd1 = cf[3], d2 = cf[4], mix = tb.coef[32];
    const float* hn = &tb.HA[sec + 1][0][0];
    z.ab1 = zero; z.ab2 = zero; z.cd1 = zero; z.cd2 = zero;
#pragma unroll
    for (int k = 0; k < kH; k += 2) {
        const float4 h = lds4(hn + 2 * k);             // next stage's (h1[k], h2[k], h1[k+1], h2[k+1])
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const float h1 = j ? h.z : h.x, h2 = j ? h.w : h.y;
            // z1' = b1 x - a1 y + z2 = (b1 x + z2) + 2 y - d1 y
            // z2' = b2 x - a2 y = (b2 x - y) + d2 y = ((b2 - b0) x - z1) + d2 y     [y = b0 x + z1]
            // -> every read of x[k] precedes y, so y is written in place (no register copies in the loop)
            const p2 xa = x[k + j], xc = x[kH + k + j];
            const p2 ta = BP ? sa2 : fma2s(b1, xa, sa2), tc = BP ? sc2 : fma2s(b1, xc, sc2);
            const p2 ua = fma2s_sub(b20, xa, sa1), uc = fma2s_sub(b20, xc, sc1);
            const p2 ya = fma2s(b0, xa, sa1), yc = fma2s(b0, xc, sc1);
            sa1 = fma2s(2.0f, ya, fma2s(nd1, ya, ta)); sc1 = fma2s(2.0f, yc, fma2s(nd1, yc, tc));
            sa2 = fma2s(d2, ya, ua); sc2 = fma2s(d2, yc, uc);
            const p2 oa = BP ? fma2s(mix, ya, xa) : ya, oc = BP ? fma2s(mix, yc, xc) : yc;
            x[k + j] = oa; x[kH + k + j] = oc;
            z.ab1 = fma2s(h1, oa, z.ab1); z.ab2 = fma2s(h2, oa, z.ab2);
            z.cd1 = fma2s(h1, oc, z.cd1); z.cd2 = fma2s(h2, oc, z.cd2);
        }
    }
    // publish the exact sequential state after the tile's last valid 32-frame chunk (B or D of thread t_last)
    const float sn1 = cx.half_last ? hi(sc1) : hi(sa1), sn2 = cx.half_last ? hi(sc2) : hi(sa2);
    if (W == 1) {
        const float g1 = shfl_idx(sn1, cx.t_last), g2 = shfl_idx(sn2, cx.t_last);
#pragma unroll
        for (int q = 0; q < 6; ++q)
            if (q == sec) { cx.c[2 * q] = g1; cx.c[2 * q + 1] = g2; }
    } else if (cx.tid == cx.t_last) {
        cx.sm->carry[cx.par ^ 1][2 * sec] = sn1; cx.sm->carry[cx.par ^ 1][2 * sec + 1] = sn2;
    }
}

// ---- one one-pole stage: s += c (x - s)  (ERB Tilt.jsfx:397 and EasyExpander.jsfx:186 forms) --------
// In: (z.ab1, z.cd1) = zero-state end states of this stage; O == 0 also accumulates pass A of S8 into them.
template <int W, int NBUF, int O>
__device__ __forceinline__ void onepole_stage(X2Ctx<W, NBUF>& cx, p2 (&x)[kL], ZS& z)
{
    const C8TablesX2& tb = cx.sm->tab;
    constexpr int sec = 6 + O, ci = 12 + O;
    const float zA = lo(z.ab1), zB = hi(z.ab1), zC = lo(z.cd1), zD = hi(z.cd1);
    const float a16 = tb.P1[O][33], a32 = tb.P1[O][34];
    const float cAB = fmaf(a16, zA, zB), cCD = fmaf(a16, zC, zD);
    float c = fmaf(a32, cAB, cCD);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float u = shfl_up(c, d);
        const float m = tb.P1[O][d];
        if (cx.lane >= d) c = fmaf(m, u, c);
    }
    float s;
    if (W == 1) {
        s = cx.c[ci];
    } else {
        if (cx.lane == 31) cx.sm->wtot[sec][cx.warp][0] = c;
        __syncthreads();
        s = cx.sm->carry[cx.par][ci];
        const float m = tb.P1[O][32];
        for (int w = 0; w < cx.warp; ++w) s = fmaf(m, s, cx.sm->wtot[sec][w][0]);
    }
    float e = shfl_up(c, 1);
    if (cx.lane == 0) e = 0;
    const float sA = fmaf(tb.P1[O][cx.lane], s, e);
    const float sB = fmaf(a16, sA, zA);
    const float sC = fmaf(a32, sA, cAB);
    const float sD = fmaf(a16, sC, zC);
    p2 sa = pk(sA, sB), sc = pk(sC, sD);
    const float cc = tb.coef[30 + O];
    const p2 zero = pk(0.f, 0.f);
    p2 pa = zero, pc = zero, qa = zero, qc = zero;
#pragma unroll
    for (int k = 0; k < kH; k += 4) {
        float4 h = make_float4(0, 0, 0, 0);
        if (O == 0) h = lds4(&tb.HB[k]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sa = fma2s(cc, sub2(x[k + j], sa), sa); sc = fma2s(cc, sub2(x[kH + k + j], sc), sc);
            x[k + j] = sa; x[kH + k + j] = sc;
            if (O == 0) {
                const float hh = j == 0 ? h.x : j == 1 ? h.y : j == 2 ? h.z : h.w;
                if (j & 1) { qa = fma2s(hh, sa, qa); qc = fma2s(hh, sc, qc); }
                else       { pa = fma2s(hh, sa, pa); pc = fma2s(hh, sc, pc); }
            }
        }
    }
    if (O == 0) { z.ab1 = add2(pa, qa); z.cd1 = add2(pc, qc); }
    const float sn = cx.half_last ? hi(sc) : hi(sa);
    if (W == 1) cx.c[ci] = shfl_idx(sn, cx.t_last);
    else if (cx.tid == cx.t_last) cx.sm->carry[cx.par ^ 1][ci] = sn;
}

// ---- the whole cascade over one tile (cascade8.jsfx @sample) ----------------------------------
template <int W, int NBUF>
__device__ __forceinline__ void cascade8_x2_tile(X2Ctx<W, NBUF>& cx, p2 (&x)[kL])
{
    ZS z;
    pass_a0(cx.sm->tab, x, z);
    // S1 HP, S2 LP | S3 BP(peak) + mix, S4 BP(skirt) + mix | S5 LP, S6 HP: two loop bodies, run-time stage index
#pragma unroll 1
    for (int g = 0; g < 3; ++g) {
        if (g == 1) {
#pragma unroll 1
            for (int j = 0; j < 2; ++j) biquad_stage<W, NBUF, true>(cx, x, 2 + j, z);
        } else {
#pragma unroll 1
            for (int j = 0; j < 2; ++j) biquad_stage<W, NBUF, false>(cx, x, 2 * g + j, z);
        }
    }
    onepole_stage<W, NBUF, 0>(cx, x, z);    // S7 (its pass A came out of S6's loop)
    onepole_stage<W, NBUF, 1>(cx, x, z);    // S8
}

__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap* map, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];"
                 ::"l"(map), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

struct X2Args {
    int64_t n_chunks;          // 32-frame chunks per stream in this launch
    const void* tables; double* state;
};

// One CTA per stream, persistent over time.  n_frames % 32 == 0, rows 16-byte aligned (TMA).
// The tensor maps have box {32 frames, 64 chunks, 1 stream}: one box = one warp's share of a tile.
template <int W, int NBUF>
__global__ void __launch_bounds__(32 * W, 16 / W)     // <= 128 registers: 4 warps per SM sub-partition (16 K registers each)
cascade8_x2_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map, const X2Args a)
{
    constexpr int T = 32 * W;
    constexpr int kWarpBytes = 64 * kL * 4;            // 8 KB: 64 rows of 128 B
    extern __shared__ uint8_t smem_raw[];
    using Smem = X2Smem<W, NBUF>;
    // pointer arithmetic on the __shared__ array itself keeps the address space known to ptxas (LDS/STS, not generic LD/ST)
    Smem* sm = reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    const int stream = blockIdx.x;
    X2Ctx<W, NBUF> cx;
    cx.sm = sm;
    cx.tid = threadIdx.x; cx.lane = threadIdx.x & 31; cx.warp = threadIdx.x >> 5; cx.par = 0;

    const int64_t n_chunks = a.n_chunks;
    const int n_tiles = (int)((n_chunks + 2 * T - 1) / (2 * T));
    const int wrow = 64 * cx.warp;                     // first row (chunk) of this warp inside a tile

    if (cx.tid == 0) {
        prefetch_tensormap(&in_map); prefetch_tensormap(&out_map);
        for (int s = 0; s < NBUF; ++s)
            for (int w = 0; w < W; ++w) mbar_init(&sm->full[s][w], 1);
        fence_barrier_init();
    }
    {   // per-stream tables -> shared memory (once per launch)
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const C8TablesX2*>(a.tables) + stream);
        uint4* dst = reinterpret_cast<uint4*>(&sm->tab);
        for (int i = cx.tid; i < (int)(sizeof(C8TablesX2) / 16); i += T) dst[i] = src[i];
    }
    const double* gstate = a.state + (size_t)stream * kC8State;
    if (W == 1) {
#pragma unroll
        for (int i = 0; i < kC8State; ++i) cx.c[i] = (float)gstate[i];
    } else if (cx.tid < kC8State) {
        sm->carry[0][cx.tid] = (float)gstate[cx.tid];
    }
    __syncthreads();
    // a warp whose 64 chunks lie entirely past the end takes no part in the copies of that tile
    auto warp_live = [&](int tile_idx) { return (int64_t)tile_idx * 2 * T + wrow < n_chunks; };
    if (cx.lane == 0) {
        for (int s = 0; s < NBUF && s < n_tiles; ++s)
            if (warp_live(s)) {
                mbar_expect_tx(&sm->full[s][cx.warp], kWarpBytes);
                tma_load_3d(sm->tile[s] + wrow * kL, &in_map, &sm->full[s][cx.warp], 0, s * 2 * T + wrow, stream);
            }
        if (NBUF == 1 && n_tiles > 1 && warp_live(1)) tma_prefetch_3d(&in_map, 0, 2 * T + wrow, stream);
    }

    p2 x[kL];
    for (int i = 0; i < n_tiles; ++i) {
        const int b = NBUF == 1 ? 0 : (i & 1);
        float* tile = sm->tile[b];
        const int64_t c0 = (int64_t)i * 2 * T;
        const int nvc = (int)((n_chunks - c0) < 2 * T ? (n_chunks - c0) : 2 * T);    // valid chunks in this tile
        cx.t_last = (nvc - 1) >> 1;
        cx.half_last = (nvc - 1) & 1;
        const bool live = warp_live(i);
        if (live) {
            mbar_wait(&sm->full[b][cx.warp], (uint32_t)((NBUF == 1 ? i : (i >> 1)) & 1));
            // row 2t = chunks A | B, row 2t+1 = chunks C | D  ->  register pairs (A[k], B[k]), (C[k], D[k])
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 va = lds4(tile + swz_piece(2 * cx.tid, j)), vb = lds4(tile + swz_piece(2 * cx.tid, j + 4));
                const float4 vc = lds4(tile + swz_piece(2 * cx.tid + 1, j)), vd = lds4(tile + swz_piece(2 * cx.tid + 1, j + 4));
                x[4 * j + 0] = pk(va.x, vb.x); x[4 * j + 1] = pk(va.y, vb.y);
                x[4 * j + 2] = pk(va.z, vb.z); x[4 * j + 3] = pk(va.w, vb.w);
                x[kH + 4 * j + 0] = pk(vc.x, vd.x); x[kH + 4 * j + 1] = pk(vc.y, vd.y);
                x[kH + 4 * j + 2] = pk(vc.z, vd.z); x[kH + 4 * j + 3] = pk(vc.w, vd.w);
            }
        } else {
#pragma unroll
            for (int k = 0; k < kL; ++k) x[k] = pk(0.f, 0.f);
        }

        cascade8_x2_tile<W, NBUF>(cx, x);

        if (live) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                *reinterpret_cast<float4*>(tile + swz_piece(2 * cx.tid, j)) =
                    make_float4(lo(x[4 * j]), lo(x[4 * j + 1]), lo(x[4 * j + 2]), lo(x[4 * j + 3]));
                *reinterpret_cast<float4*>(tile + swz_piece(2 * cx.tid, j + 4)) =
                    make_float4(hi(x[4 * j]), hi(x[4 * j + 1]), hi(x[4 * j + 2]), hi(x[4 * j + 3]));
                *reinterpret_cast<float4*>(tile + swz_piece(2 * cx.tid + 1, j)) =
                    make_float4(lo(x[kH + 4 * j]), lo(x[kH + 4 * j + 1]), lo(x[kH + 4 * j + 2]), lo(x[kH + 4 * j + 3]));
                *reinterpret_cast<float4*>(tile + swz_piece(2 * cx.tid + 1, j + 4)) =
                    make_float4(hi(x[kH + 4 * j]), hi(x[kH + 4 * j + 1]), hi(x[kH + 4 * j + 2]), hi(x[kH + 4 * j + 3]));
            }
            fence_proxy_async();
            __syncwarp();
            if (cx.lane == 0) {
                tma_store_3d(&out_map, tile + wrow * kL, 0, (int)c0 + wrow, stream);
                tma_commit();
                const int nx = i + NBUF;                      // the tile that reuses this buffer
                if (nx < n_tiles && warp_live(nx)) {
                    tma_wait_read<0>();                       // the store has read the rows: refill them
                    mbar_expect_tx(&sm->full[b][cx.warp], kWarpBytes);
                    tma_load_3d(tile + wrow * kL, &in_map, &sm->full[b][cx.warp], 0, nx * 2 * T + wrow, stream);
                    if (NBUF == 1 && nx + 1 < n_tiles && warp_live(nx + 1))
                        tma_prefetch_3d(&in_map, 0, (nx + 1) * 2 * T + wrow, stream);
                }
            }
        }
        cx.par ^= 1;
    }

    // state after the last frame -> HBM (double), carried into the next call
    double* gs = a.state + (size_t)stream * kC8State;
    if (W == 1) {
        if (cx.tid == 0) {
#pragma unroll
            for (int i = 0; i < 14; ++i) gs[i] = (double)cx.c[i];
        }
    } else {
        __syncthreads();
        if (cx.tid < 14) gs[cx.tid] = (double)sm->carry[cx.par][cx.tid];
    }
    if (cx.lane == 0) tma_wait_all<0>();
}

// ---------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------
template <int W, int NBUF>
static bool launch_x2(const C8Launch& L)
{
    X2Args a{L.n_frames / kL, L.tables_x2, L.state};
    CUtensorMap in_map, out_map;
    if (!make_chunk_tensor_map(&in_map, L.in, a.n_chunks, L.n_streams, L.row_stride, 64)) return false;
    if (!make_chunk_tensor_map(&out_map, L.out, a.n_chunks, L.n_streams, L.row_stride, 64)) return false;
    auto kern = cascade8_x2_kernel<W, NBUF>;
    const size_t smem = sizeof(X2Smem<W, NBUF>) + 1024;
    static DeviceOnce attr_once;    // per template instantiation
    if (attr_once.pending()) {
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "smem attr")) return false;
        // all of L1 as shared memory: the kernel's only L1 traffic is TMA, and 8 CTAs x 26 KB must fit
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared), "carveout")) return false;
        attr_once.mark();
    }
    kern<<<L.n_streams, 32 * W, smem, L.stream>>>(in_map, out_map, a);
    return cuda_ok(cudaGetLastError(), "cascade8_x2_kernel launch");
}

bool cascade8_x2_launch(const C8Launch& L)
{
    if (L.n_frames <= 0 || L.n_streams <= 0) return true;
    if ((L.n_frames % kL) != 0) { set_error("cascade8_x2: needs n_frames % 32 == 0"); return false; }
    const int key = L.warps * 10 + L.nbuf;
    switch (key) {
    case 11: return launch_x2<1, 1>(L);
    case 12: return launch_x2<1, 2>(L);
    case 21: return launch_x2<2, 1>(L);
    case 22: return launch_x2<2, 2>(L);
    case 41: return launch_x2<4, 1>(L);
    case 42: return launch_x2<4, 2>(L);
    default: set_error("cascade8_x2: warps_per_stream must be 1, 2 or 4 and tile_buffers 1 or 2"); return false;
    }
}

// Warps per stream and tile buffers: all streams resident in one wave when they fit (one CTA = one
// stream runs the whole time axis), up to 16 warps per SM (128 registers per thread), double-buffered
// tiles only if shared memory still fits that many CTAs.
void cascade8_x2_pick(int n_streams, int device, int* warps, int* nbuf)
{
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int per_sm = (n_streams + sms - 1) / sms;      // CTAs each SM must host for one wave
    int w = 4;
    while (w > 1 && per_sm * w > 16) w >>= 1;
    const int ctas = per_sm < 16 / w ? per_sm : 16 / w;  // resident CTAs per SM we plan for
    const size_t smem2 = (w == 1 ? sizeof(X2Smem<1, 2>) : w == 2 ? sizeof(X2Smem<2, 2>) : sizeof(X2Smem<4, 2>)) + 2048;
    *warps = w;
    *nbuf = (smem2 * (size_t)ctas <= 227u * 1024u) ? 2 : 1;
}

}  // namespace za
