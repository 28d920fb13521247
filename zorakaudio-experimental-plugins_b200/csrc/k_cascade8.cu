// k_cascade8.cu -- linear IIR class: the 8-section cascade of BASELINE config 2 on sm_100a.
//
// Replaces, for plugin "cascade8", the reference's generated per-sample loop
// (jsfx_process_block -> jsfx_sample, dsp_jsfx_aot.py:5713-5905) for n_streams instances at once.
//
// Mapping.  One CTA per stream, persistent over time; T = 32*W threads.  Time is cut into tiles of
// T chunks x 32 frames.  A tile arrives by ONE TMA tensor load (SWIZZLE_128B, 3-stage mbarrier
// ring) and leaves by one TMA tensor store; thread t owns chunk t (a 128-byte row) in registers.
// Each section is then evaluated time-parallel as a blocked linear-recurrence scan:
//   pass A  zero-state end state of every chunk  z_t = sum_k A^(31-k) B x_k  (FIR against a
//           per-stream table, independent FMAs),
//   scan    warp-shuffle Kogge-Stone over the affine maps s -> M s + z_t with M = A^32 (2x2 / 1x1
//           transfer matrices, powers M^(2^k) from a per-stream table), cross-warp carry through
//           shared memory, tile-to-tile carry = the exact sequential state,
//   pass B  every thread re-runs the section's own recurrence over its chunk from its exact start
//           state s_t, in the reference's operation order (fp64: separate mul/add, bit-for-bit the
//           oracle's sequence), overwriting the chunk with the section output.
// Only the chunk start states are re-associated; their error decays with the filter, and the carry
// between tiles is the true sequential state, so fp64 output stays ~1e-15 of the oracle's.
#include "za_device.cuh"
#include "za_internal.h"

namespace za {

template <typename R> struct C8Tables {
    R H2[6][kL][2];    // biquads: H2[s][k] = T A^(31-k) B  (zero-state end-state response to x_k, scan coords)
    R P2[6][33][4];    // biquads: (N^32)^j, j = 0..32, row-major 2x2
    R T2[6][8];        // biquads: T (4) and T^-1 (4): scan coordinates w = T s with T A T^-1 = N normal
    R H1[2][kL];       // one-poles
    R P1[2][36];       // one-poles: (alpha^32)^j, j = 0..32 (padded)
    R coef[kC8Coefs];
};

size_t cascade8_tables_bytes(int precision) { return precision ? sizeof(C8Tables<double>) : sizeof(C8Tables<float>); }

// ---------------------------------------------------------------------------------------------
// tables: one thread per (stream, section); double arithmetic, stored as R
// ---------------------------------------------------------------------------------------------
template <typename R>
__global__ void cascade8_tables_kernel(const double* __restrict__ coef, C8Tables<R>* __restrict__ tabs, int n_streams)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int s = g >> 3, sec = g & 7;
    if (s >= n_streams) return;
    const double* c = coef + (size_t)s * kC8Coefs;
    C8Tables<R>& t = tabs[s];
    if (sec < 6) {
        const double b0 = c[5 * sec], b1 = c[5 * sec + 1], b2 = c[5 * sec + 2], a1 = c[5 * sec + 3], a2 = c[5 * sec + 4];
        // state update of the TDF-II recurrence: s' = A s + B x, A = [[-a1, 1], [-a2, 0]].
        // For poles near z = 1 the powers of A are large (|A^n| ~ 1/theta) and a scan in these
        // coordinates loses ~cond^2 * eps.  The scan therefore runs in the real normal form of A:
        // w = T s, T A T^-1 = N = [[sg, om], [-om, sg]] (a scaled rotation, |N^n| = r^n <= 1).
        const double sg = -0.5 * a1, w2 = a2 - sg * sg;
        double t00 = 1, t01 = 0, t10 = 0, t11 = 1, i00 = 1, i01 = 0, i10 = 0, i11 = 1;   // T, T^-1
        double n00 = -a1, n01 = 1, n10 = -a2, n11 = 0;                                   // N
        if (w2 > 1e-14 && a2 > 1e-14) {              // complex pole pair
            const double om = sqrt(w2);
            t00 = 0; t01 = -1.0 / a2; t10 = 1.0 / om; t11 = sg / (om * a2);
            i00 = sg; i01 = om; i10 = -a2; i11 = 0;
            n00 = sg; n01 = om; n10 = -om; n11 = sg;
        }
        const double bb1 = b1 - a1 * b0, bb2 = b2 - a2 * b0;
        double v1 = t00 * bb1 + t01 * bb2, v2 = t10 * bb1 + t11 * bb2;
        for (int k = kL - 1; k >= 0; --k) {
            t.H2[sec][k][0] = (R)v1; t.H2[sec][k][1] = (R)v2;
            const double x1 = n00 * v1 + n01 * v2, x2 = n10 * v1 + n11 * v2;
            v1 = x1; v2 = x2;
        }
        double m00 = 1, m01 = 0, m10 = 0, m11 = 1;       // N^32 by 32 left-multiplications
        for (int k = 0; k < kL; ++k) {
            const double x00 = n00 * m00 + n01 * m10, x01 = n00 * m01 + n01 * m11;
            const double x10 = n10 * m00 + n11 * m10, x11 = n10 * m01 + n11 * m11;
            m00 = x00; m01 = x01; m10 = x10; m11 = x11;
        }
        double p00 = 1, p01 = 0, p10 = 0, p11 = 1;
        for (int j = 0; j <= 32; ++j) {
            t.P2[sec][j][0] = (R)p00; t.P2[sec][j][1] = (R)p01; t.P2[sec][j][2] = (R)p10; t.P2[sec][j][3] = (R)p11;
            const double x00 = m00 * p00 + m01 * p10, x01 = m00 * p01 + m01 * p11;
            const double x10 = m10 * p00 + m11 * p10, x11 = m10 * p01 + m11 * p11;
            p00 = x00; p01 = x01; p10 = x10; p11 = x11;
        }
        t.T2[sec][0] = (R)t00; t.T2[sec][1] = (R)t01; t.T2[sec][2] = (R)t10; t.T2[sec][3] = (R)t11;
        t.T2[sec][4] = (R)i00; t.T2[sec][5] = (R)i01; t.T2[sec][6] = (R)i10; t.T2[sec][7] = (R)i11;
    } else {
        const int o = sec - 6;
        // S7: lp = (1-a7) x + a7 lp ; S8: lp += c8 (x - lp)  ->  s' = alpha s + beta x
        const double alpha = o == 0 ? c[30] : 1.0 - c[32];
        const double beta = o == 0 ? c[31] : c[32];
        double v = beta;
        for (int k = kL - 1; k >= 0; --k) { t.H1[o][k] = (R)v; v *= alpha; }
        double m = 1;
        for (int k = 0; k < kL; ++k) m *= alpha;
        double p = 1;
        for (int j = 0; j <= 32; ++j) { t.P1[o][j] = (R)p; p *= m; }
        for (int j = 33; j < 36; ++j) t.P1[o][j] = (R)0;
    }
    if (sec == 0) {
        for (int i = 0; i < kC8Coefs; ++i) t.coef[i] = (R)c[i];
        if (sizeof(R) == 4)     // fp32 mode: delta-coded feedback, a1 = -2 + d1, a2 = 1 - d2 (see biquad pass B)
            for (int k = 0; k < 6; ++k) { t.coef[5 * k + 3] = (R)(c[5 * k + 3] + 2.0); t.coef[5 * k + 4] = (R)(1.0 - c[5 * k + 4]); }
    }
}

bool cascade8_build_tables(const double* d_coef, void* d_tables, int n_streams, int precision, cudaStream_t s)
{
    const int threads = 128, blocks = (n_streams * 8 + threads - 1) / threads;
    if (precision)
        cascade8_tables_kernel<double><<<blocks, threads, 0, s>>>(d_coef, (C8Tables<double>*)d_tables, n_streams);
    else
        cascade8_tables_kernel<float><<<blocks, threads, 0, s>>>(d_coef, (C8Tables<float>*)d_tables, n_streams);
    return cuda_ok(cudaGetLastError(), "cascade8_tables_kernel");
}

// ---------------------------------------------------------------------------------------------
// per-CTA context
// ---------------------------------------------------------------------------------------------
template <typename R, int W> struct C8Smem {
    alignas(1024) float tile[3][32 * W * kL];    // 3-stage ring, each stage a swizzled T x 32 tile
    alignas(16) C8Tables<R> tab;
    alignas(16) R carry[2][kC8State];            // tile-parity double buffer (W > 1 only)
    alignas(16) R wtot[8][W][2];                 // per-section warp totals (W > 1 only)
    alignas(8) uint64_t full[3];
};

template <typename R, int W> struct Ctx {
    C8Smem<R, W>* sm;
    int tid, lane, warp;
    int t_last, v_last;       // owner thread of the tile's last valid frame, and its valid count
    int par;                  // tile parity for the carry double buffer
    R c[kC8State];            // W == 1: carry lives in registers
};

// ---- one biquad section over the thread's chunk ------------------------------------------------
template <typename R, int W, bool PARTIAL, bool MIX>
__device__ __forceinline__ void biquad_section(Ctx<R, W>& cx, R (&x)[kL], const int sec)
{
    const C8Tables<R>& tb = cx.sm->tab;
    // pass A: zero-state end state of my chunk (independent FMAs, 2 accumulator pairs)
    R z1 = 0, z2 = 0, y1 = 0, y2 = 0;
#pragma unroll
    for (int k = 0; k < kL; k += 2) {
        const Quad<R> h = ld4(&tb.H2[sec][k][0]);
        z1 = fma_(h.x, x[k], z1); z2 = fma_(h.y, x[k], z2);
        y1 = fma_(h.z, x[k + 1], y1); y2 = fma_(h.w, x[k + 1], y2);
    }
    z1 += y1; z2 += y2;
    // inclusive Kogge-Stone scan of s -> M s + z over the warp's 32 chunks
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const R n1 = shfl_up(z1, d), n2 = shfl_up(z2, d);
        const Quad<R> m = ld4(&tb.P2[sec][d][0]);
        if (cx.lane >= d) {
            z1 = fma_(m.x, n1, fma_(m.y, n2, z1));
            z2 = fma_(m.z, n1, fma_(m.w, n2, z2));
        }
    }
    // state entering my warp: exact sequential state (z1, z2) -> scan coordinates w = T s
    const Quad<R> tf = ld4(&tb.T2[sec][0]);
    const Quad<R> ti = ld4(&tb.T2[sec][4]);
    R s1, s2;
    if (W == 1) {
        const R c1 = cx.c[2 * sec], c2 = cx.c[2 * sec + 1];
        s1 = fma_(tf.x, c1, tf.y * c2); s2 = fma_(tf.z, c1, tf.w * c2);
    } else {
        if (cx.lane == 31) { cx.sm->wtot[sec][cx.warp][0] = z1; cx.sm->wtot[sec][cx.warp][1] = z2; }
        __syncthreads();
        const R c1 = cx.sm->carry[cx.par][2 * sec], c2 = cx.sm->carry[cx.par][2 * sec + 1];
        s1 = fma_(tf.x, c1, tf.y * c2); s2 = fma_(tf.z, c1, tf.w * c2);
        const Quad<R> m = ld4(&tb.P2[sec][32][0]);
        for (int w = 0; w < cx.warp; ++w) {
            const R t1 = fma_(m.x, s1, fma_(m.y, s2, cx.sm->wtot[sec][w][0]));
            const R t2 = fma_(m.z, s1, fma_(m.w, s2, cx.sm->wtot[sec][w][1]));
            s1 = t1; s2 = t2;
        }
    }
    // my start state: exclusive prefix + N^(32 lane) * warp-incoming state, back to (z1, z2) = T^-1 w
    R e1 = shfl_up(z1, 1), e2 = shfl_up(z2, 1);
    if (cx.lane == 0) { e1 = 0; e2 = 0; }
    const Quad<R> pl = ld4(&tb.P2[sec][cx.lane][0]);
    const R w1 = fma_(pl.x, s1, fma_(pl.y, s2, e1));
    const R w2 = fma_(pl.z, s1, fma_(pl.w, s2, e2));
    R st1 = fma_(ti.x, w1, ti.y * w2);
    R st2 = fma_(ti.z, w1, ti.w * w2);
    // pass B: the reference recurrence, reference operation order (Spectral Stabilizer.jsfx:98-111)
    const R b0 = tb.coef[5 * sec], b1 = tb.coef[5 * sec + 1], b2 = tb.coef[5 * sec + 2];
    const R a1 = tb.coef[5 * sec + 3], a2 = tb.coef[5 * sec + 4];
    const R mix = tb.coef[33];
    R sn1 = st1, sn2 = st2;
#pragma unroll
    for (int k = 0; k < kL; ++k) {
        const R xx = x[k];
        R y, n1, n2;
        if (sizeof(R) == 8) {
            y = add(mul(b0, xx), st1);
            n1 = add(sub(mul(b1, xx), mul(a1, y)), st2);
            n2 = sub(mul(b2, xx), mul(a2, y));
        } else {
            // fp32 mode: a1, a2 hold d1 = a1 + 2 and d2 = 1 - a2.  Poles near z = 1 have a1 ~ -2,
            // a2 ~ 1; rounding those to float moves a 40 Hz pole by ~1e-3 of its frequency, while
            // d1, d2 keep full relative precision.  Same recurrence, one extra FMA.
            y = fma_(b0, xx, st1);
            n1 = fma_((R)2, y, fma_(-a1, y, fma_(b1, xx, st2)));
            n2 = fma_(a2, y, fma_(b2, xx, -y));
        }
        st1 = n1; st2 = n2;
        x[k] = MIX ? add(xx, mul(mix, y)) : y;
        if (PARTIAL && k + 1 == cx.v_last) { sn1 = st1; sn2 = st2; }
    }
    if (!PARTIAL) { sn1 = st1; sn2 = st2; }
    // publish the exact sequential state after the tile's last valid frame
    if (W == 1) {
        cx.c[2 * sec] = shfl_idx(sn1, cx.t_last); cx.c[2 * sec + 1] = shfl_idx(sn2, cx.t_last);
    } else if (cx.tid == cx.t_last) {
        cx.sm->carry[cx.par ^ 1][2 * sec] = sn1; cx.sm->carry[cx.par ^ 1][2 * sec + 1] = sn2;
    }
}

// ---- one one-pole section ----------------------------------------------------------------------
// KIND 0: lp = (1-a)*x + a*lp   (ERB Tilt.jsfx:397);   KIND 1: lp += c*(x - lp)   (EasyExpander.jsfx:186)
template <typename R, int W, bool PARTIAL, int KIND>
__device__ __forceinline__ void onepole_section(Ctx<R, W>& cx, R (&x)[kL])
{
    const C8Tables<R>& tb = cx.sm->tab;
    constexpr int o = KIND, sec = 6 + KIND, ci = 12 + KIND;
    R z = 0, za = 0, zb = 0, zc = 0;
#pragma unroll
    for (int k = 0; k < kL; k += 4) {
        const Quad<R> h = ld4(&tb.H1[o][k]);
        z = fma_(h.x, x[k], z); za = fma_(h.y, x[k + 1], za);
        zb = fma_(h.z, x[k + 2], zb); zc = fma_(h.w, x[k + 3], zc);
    }
    z = (z + za) + (zb + zc);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const R n = shfl_up(z, d);
        const R m = tb.P1[o][d];
        if (cx.lane >= d) z = fma_(m, n, z);
    }
    R s;
    if (W == 1) {
        s = cx.c[ci];
    } else {
        if (cx.lane == 31) cx.sm->wtot[sec][cx.warp][0] = z;
        __syncthreads();
        s = cx.sm->carry[cx.par][ci];
        const R m = tb.P1[o][32];
        for (int w = 0; w < cx.warp; ++w) s = fma_(m, s, cx.sm->wtot[sec][w][0]);
    }
    R e = shfl_up(z, 1);
    if (cx.lane == 0) e = 0;
    R st = fma_(tb.P1[o][cx.lane], s, e);
    const R a7 = tb.coef[30], k7 = tb.coef[31], c8 = tb.coef[32];
    R sn = st;
#pragma unroll
    for (int k = 0; k < kL; ++k) {
        const R xx = x[k];
        if (KIND == 0) st = add(mul(k7, xx), mul(a7, st));
        else           st = add(st, mul(c8, sub(xx, st)));
        x[k] = st;
        if (PARTIAL && k + 1 == cx.v_last) sn = st;
    }
    if (!PARTIAL) sn = st;
    if (W == 1) cx.c[ci] = shfl_idx(sn, cx.t_last);
    else if (cx.tid == cx.t_last) cx.sm->carry[cx.par ^ 1][ci] = sn;
}

// ---- the whole cascade over one tile (cascade8.jsfx @sample) ----------------------------------
template <typename R, int W, bool PARTIAL>
__device__ __forceinline__ void cascade8_tile(Ctx<R, W>& cx, R (&x)[kL])
{
    biquad_section<R, W, PARTIAL, false>(cx, x, 0);   // S1 HP
    biquad_section<R, W, PARTIAL, false>(cx, x, 1);   // S2 LP
    biquad_section<R, W, PARTIAL, true>(cx, x, 2);    // S3 BP (peak), u3 = y2 + mix*y3
    biquad_section<R, W, PARTIAL, true>(cx, x, 3);    // S4 BP (skirt), u4 = u3 + mix*y4
    biquad_section<R, W, PARTIAL, false>(cx, x, 4);   // S5 LP
    biquad_section<R, W, PARTIAL, false>(cx, x, 5);   // S6 HP
    onepole_section<R, W, PARTIAL, 0>(cx, x);         // S7
    onepole_section<R, W, PARTIAL, 1>(cx, x);         // S8
}

template <int W> __device__ __forceinline__ void cta_sync()
{
    if (W == 1) __syncwarp(); else __syncthreads();
}

struct C8Args {
    const float* in; float* out; double* tap;
    int64_t row_stride, n_frames;
    const void* tables; double* state;
    int n_streams;
};

// ---------------------------------------------------------------------------------------------
// main kernel.  TMA = true: n_frames % 32 == 0, tiles move by cp.async.bulk.tensor (3-stage ring).
//               TMA = false: any n_frames / alignment, guarded plain loads and stores (fallback + tails).
// ---------------------------------------------------------------------------------------------
template <typename R, int W, bool TMA, bool TAP>
__global__ void __launch_bounds__(32 * W)
cascade8_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map, const C8Args a)
{
    constexpr int T = 32 * W;
    extern __shared__ uint8_t smem_raw[];
    C8Smem<R, W>* sm = reinterpret_cast<C8Smem<R, W>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    const int stream = blockIdx.x;
    Ctx<R, W> cx;
    cx.sm = sm; cx.tid = threadIdx.x; cx.lane = threadIdx.x & 31; cx.warp = threadIdx.x >> 5; cx.par = 0;

    // per-stream tables -> shared memory (once per launch)
    {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const C8Tables<R>*>(a.tables) + stream);
        uint4* dst = reinterpret_cast<uint4*>(&sm->tab);
        for (int i = cx.tid; i < (int)(sizeof(C8Tables<R>) / 16); i += T) dst[i] = src[i];
    }
    const double* gstate = a.state + (size_t)stream * kC8State;
    if (W == 1) {
#pragma unroll
        for (int i = 0; i < kC8State; ++i) cx.c[i] = (R)gstate[i];
    } else if (cx.tid < kC8State) {
        sm->carry[0][cx.tid] = (R)gstate[cx.tid];
    }

    const int64_t n = a.n_frames;
    const int64_t tile_frames = (int64_t)T * kL;
    const int n_tiles = (int)((n + tile_frames - 1) / tile_frames);
    const float* in_row = a.in + (size_t)stream * a.row_stride;
    float* out_row = a.out + (size_t)stream * a.row_stride;

    if (TMA) {
        if (cx.tid == 0) {
            prefetch_tensormap(&in_map); prefetch_tensormap(&out_map);
            for (int s = 0; s < 3; ++s) mbar_init(&sm->full[s], 1);
            fence_barrier_init();
        }
        __syncthreads();
        if (cx.tid == 0) {
            for (int s = 0; s < 2 && s < n_tiles; ++s) {
                mbar_expect_tx(&sm->full[s], T * kL * 4);
                tma_load_3d(sm->tile[s], &in_map, &sm->full[s], 0, s * T, stream);
            }
        }
    } else {
        __syncthreads();
    }

    R x[kL];
    for (int i = 0; i < n_tiles; ++i) {
        const int b = TMA ? i % 3 : 0;
        float* tile = sm->tile[b];
        const int64_t f0 = (int64_t)i * tile_frames;
        const int nv = (int)((n - f0) < tile_frames ? (n - f0) : tile_frames);   // valid frames in this tile
        cx.t_last = (nv - 1) / kL;
        cx.v_last = nv - cx.t_last * kL;
        if (TMA) {
            mbar_wait(&sm->full[b], (uint32_t)((i / 3) & 1));
        } else {
            for (int idx = cx.tid; idx < T * kL; idx += T)
                tile[swz_elem(idx >> 5, idx & 31)] = idx < nv ? in_row[f0 + idx] : 0.0f;
            cta_sync<W>();
        }
        tile_row_load<R>(tile, cx.tid, x);

        if (!TMA && nv < tile_frames) cascade8_tile<R, W, true>(cx, x);
        else                          cascade8_tile<R, W, false>(cx, x);

        if (TAP) {
            double* tp = a.tap + (size_t)stream * a.row_stride + f0 + (int64_t)cx.tid * kL;
#pragma unroll
            for (int k = 0; k < kL; ++k)
                if (cx.tid * kL + k < nv) tp[k] = (double)x[k];
        }
        tile_row_store<R>(tile, cx.tid, x);
        if (TMA) {
            fence_proxy_async();
            cta_sync<W>();
            if (cx.tid == 0) {
                tma_store_3d(&out_map, tile, 0, i * T, stream);
                tma_commit();
                tma_wait_read<1>();                       // store i-1 has left its buffer
                const int nx = i + 2;
                if (nx < n_tiles) {
                    const int nb = nx % 3;
                    mbar_expect_tx(&sm->full[nb], T * kL * 4);
                    tma_load_3d(sm->tile[nb], &in_map, &sm->full[nb], 0, nx * T, stream);
                }
            }
        } else {
            cta_sync<W>();
            for (int idx = cx.tid; idx < T * kL; idx += T)
                if (idx < nv) out_row[f0 + idx] = tile[swz_elem(idx >> 5, idx & 31)];
            cta_sync<W>();
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
    if (TMA && cx.tid == 0) tma_wait_all<0>();
}

// ---------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------
template <typename R, int W, bool TMA, bool TAP>
static bool launch_variant(const C8Launch& L)
{
    C8Args a{L.in, L.out, L.tap, L.row_stride, L.n_frames, L.tables, L.state, L.n_streams};
    CUtensorMap in_map, out_map;
    memset(&in_map, 0, sizeof in_map); memset(&out_map, 0, sizeof out_map);
    if (TMA) {
        const int64_t n_chunks = L.n_frames / kL;
        if (!make_chunk_tensor_map(&in_map, L.in, n_chunks, L.n_streams, L.row_stride, 32 * W)) return false;
        if (!make_chunk_tensor_map(&out_map, L.out, n_chunks, L.n_streams, L.row_stride, 32 * W)) return false;
    }
    auto kern = cascade8_kernel<R, W, TMA, TAP>;
    const size_t smem = sizeof(C8Smem<R, W>) + 1024;
    static DeviceOnce attr_once;    // per template instantiation
    if (attr_once.pending()) {
        if (!cuda_ok(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "smem attr")) return false;
        attr_once.mark();
    }
    kern<<<L.n_streams, 32 * W, smem, L.stream>>>(in_map, out_map, a);
    return cuda_ok(cudaGetLastError(), "cascade8_kernel launch");
}

template <typename R, bool TMA, bool TAP> static bool launch_w(const C8Launch& L)
{
    switch (L.warps) {
    case 1: return launch_variant<R, 1, TMA, TAP>(L);
    case 2: return launch_variant<R, 2, TMA, TAP>(L);
    case 4: return launch_variant<R, 4, TMA, TAP>(L);
    default: set_error("cascade8: warps_per_stream must be 1, 2 or 4"); return false;
    }
}

bool cascade8_launch(const C8Launch& L)
{
    if (L.n_frames <= 0 || L.n_streams <= 0) return true;
    if (L.tap && !L.precision) { set_error("cascade8: the double tap needs precision f64"); return false; }
    if (L.use_tma && (L.n_frames % kL) != 0) { set_error("cascade8: TMA launch needs n_frames % 32 == 0"); return false; }
    if (L.precision) {
        if (L.tap) return L.use_tma ? launch_w<double, true, true>(L) : launch_w<double, false, true>(L);
        return L.use_tma ? launch_w<double, true, false>(L) : launch_w<double, false, false>(L);
    }
    return L.use_tma ? launch_w<float, true, false>(L) : launch_w<float, false, false>(L);
}

// Warps per stream: as many resident warps per SM as registers allow without spilling into a
// second wave of CTAs (one CTA = one stream, all streams resident at once when they fit).
int cascade8_pick_warps(int n_streams, int precision, int device)
{
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int per_sm = (n_streams + sms - 1) / sms;          // CTAs each SM must host for one wave
    const int max_warps_sm = precision ? 16 : 24;            // register budget: ~128 (f64) / ~80 (f32) regs per thread
    int w = 4;
    while (w > 1 && per_sm * w > max_warps_sm) w >>= 1;
    return w;
}

}  // namespace za
