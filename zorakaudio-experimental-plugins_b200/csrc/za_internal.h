// za_internal.h -- host-side declarations shared between the C-ABI layer and the kernel files.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <string>

namespace za {

void set_error(const std::string& msg);          // thread-local text behind za_gpu_last_error()
bool cuda_ok(cudaError_t e, const char* what);   // false + set_error on failure
// cudaFuncSetAttribute is a per-device setting: one of these per kernel remembers which devices already have it
// (a process may drive several GPUs -- za_batch_* runs one host thread per device).
struct DeviceOnce {
    std::atomic<unsigned long long> mask{0};
    static unsigned long long bit() { int d = 0; cudaGetDevice(&d); return 1ull << (d & 63); }
    bool pending() const { return !(mask.load(std::memory_order_acquire) & bit()); }
    void mark() { mask.fetch_or(bit(), std::memory_order_release); }
};

// Encodes a 3-D SWIZZLE_128B tensor map over planar float rows:
//   dim0 = 32 frames (one chunk), dim1 = n_chunks, dim2 = n_rows; box = {32, box_rows, 1}.
// base must be 16-byte aligned and row_stride_frames a multiple of 4.
bool make_chunk_tensor_map(CUtensorMap* map, const float* base, int64_t n_chunks, int64_t n_rows,
                           int64_t row_stride_frames, int box_rows);

// ---- cascade8 (LTI cascade: 6 TDF-II biquads + 2 one-poles) -----------------------------------
constexpr int kC8Coefs = 36;     // per stream: 6 x (b0,b1,b2,a1,a2), a7, 1-a7, c8, mix, pad
constexpr int kC8State = 16;     // per stream: z[6][2], lp7, lp8, pad (double, persists across calls)
// k_cascade8_auto.cu: renders with per-block coefficient rows [n_rows][n_streams][kC8Coefs] in one launch (tables rebuilt on the device)
bool cascade8_auto_launch(const float* in, float* out, double* tap, int64_t row_stride, int64_t n_frames, int n_streams, int block_len,
                          int n_rows, const double* coef, double* state, int precision, cudaStream_t stream);

struct C8Launch {
    const float* in;         // row 0, first frame of this launch
    float* out;
    double* tap;             // optional double tap (same indexing), fp64 only
    int64_t row_stride;      // frames between consecutive streams
    int64_t n_frames;        // frames in this launch
    int n_streams;
    const void* tables;      // C8Tables<R>[n_streams] (device)
    double* state;           // [n_streams][kC8State] (device)
    int precision;           // 0 = f32, 1 = f64
    int warps;               // warps per stream (1, 2 or 4)
    bool use_tma;            // needs n_frames % 32 == 0, 16-byte aligned rows
    const void* tables_x2 = nullptr;   // C8TablesX2[n_streams] (fp32 packed kernel, k_cascade8_x2.cu)
    int nbuf = 1;            // tile buffers per CTA of the packed kernel (1 or 2)
    cudaStream_t stream;
};

size_t cascade8_tables_bytes(int precision);
bool cascade8_build_tables(const double* d_coef, void* d_tables, int n_streams, int precision, cudaStream_t s);
bool cascade8_launch(const C8Launch& a);
int cascade8_pick_warps(int n_streams, int precision, int device);

// fp32 packed-FP32 (FFMA2) kernel, two chunks per thread (k_cascade8_x2.cu); TMA-aligned spans only
size_t cascade8_x2_tables_bytes();
bool cascade8_x2_build_tables(const double* d_coef, void* d_tables, int n_streams, cudaStream_t s);
bool cascade8_x2_launch(const C8Launch& a);
void cascade8_x2_pick(int n_streams, int device, int* warps, int* nbuf);

// ---- EasyExpander (nonlinear dynamics class) -----------------------------------------------------
constexpr int kEEState = 32;     // per stream: ic1[8], ic2[8], env[8], hpf_lp, lpf_y, det_slow, expanding,
                                 // gain_db_s, det_db_m, gr_db_m, pad; stored [var][stream] (double)
struct EEConsts {                // @init results (EasyExpander.jsfx:27-171), identical for all streams
    double g[8], k[8], den[8], w[8];
    double band_c, open_c, close_fast_c, close_slow_c, slow_c;
};
struct EECoef {                  // @slider results (:154-170) of one stream for one host block
    double thresh_db, depth_db, ratio_base, hpf_c, lpf_c;
    int hpf_on, lpf_on;
};
struct EEArgs {
    const float* in; float* out; double* tap;      // [stream][2][row_stride]
    int64_t row_stride, n_frames;
    int n_streams, block_len, n_coef_blocks;
    const EECoef* coef;                            // [n_coef_blocks][n_streams]
    double* state;                                 // [kEEState][n_streams]
    EEConsts k;
    int role_off;                                  // measurement only (option "role_off"): 1 bands, 2 filter, 4 G1, 8 G2, 16 helpers' P2, 32 helpers' I/O skip their work
};
bool easyexpander_launch(const EEArgs& a, int precision, int n_sm, cudaStream_t stream, int variant = 0);
bool easyexpander_v3_eligible(const EEArgs& a);          // k_easyexpander_v3.cu: fp32, block_len a multiple of 32
bool easyexpander_v3_launch(const EEArgs& a, int n_sm, cudaStream_t stream);

// ---- stft_mask2048 (STFT class) ---------------------------------------------------------------------
constexpr int kStftN = 2048, kStftHop = 512, kStftBins = 1025;
struct StftState {               // per stream, persists across calls (the reference keeps these in mem[])
    float xin[2][4 * kStftHop];  // last four hops of input (slot = absolute hop index mod 4)
    double acc[2][kStftN];       // overlap-add accumulator (slot = absolute hop index mod 4)
    double base[kStftBins + 3];  // per-bin baseline EMA
    double prev[kStftBins + 3];  // per-bin previous magnitude
};
struct StftCoef {                // @slider results of one stream for one host block
    double amt, thresh, ab, bb, floor_g, mixv;
};
struct StftArgs {
    const float* in; float* out; double* tap;      // [stream][2][row_stride]
    int64_t row_stride, n_frames, t0;              // t0 = frames rendered since prepare (hop phase)
    int n_streams, block_len, n_coef_blocks;
    const StftCoef* coef;                          // [n_coef_blocks][n_streams]
    StftState* state;                              // [n_streams]
    const double* win; const double* syn;          // [2048] analysis window, [2048] synthesis window x 1/N
    const double* tw1; const double* tw2;          // [15][128][2] exp(-2 pi i t q / 2048), [15][8][2] exp(-2 pi i j q / 128)
};
size_t stft_state_bytes();
bool stft_launch(const StftArgs& a, int precision, int n_sm, cudaStream_t stream);

// ---- ModTilt (Faust leaf, float32; plugins/Dynamics/ModTilt/src/ModTilt.dsp) ---------------------------------------
constexpr int kMTState = 5;      // per stream: the five onepole() states env2, base, lp_piv, ratio, mean_rdb; [var][stream] float
struct alignas(16) MTCoef {      // slider-rate values of one stream for one compute() call (ModTilt.dsp:37,44-45,17-19)
    float a_piv, na_piv, g_lo, g_hi, mix, nmix, piv_index, pad1;     // piv_index: grid index of the pivot slider (int bits) -> power table row
};
struct MTArgs {
    const float* in; float* out;                   // [stream][2][row_stride]
    int64_t row_stride, n_frames;
    int n_streams, block_len, n_coef_blocks;
    const MTCoef* coef;                            // [n_coef_blocks][n_streams]
    float* state;                                  // [kMTState][n_streams]
    float a_env, a_base, a_trim;                   // sample-rate constants (:35-36,40)
};
// time-parallel form (k_modtilt_scan.cu, modtilt_scan_core.h): one warp per stream, five blocked scans per 256-frame tile
namespace mtscan { struct Consts; }
struct MTScanArgs;
bool modtilt_scan_ok(const MTArgs& a);             // shape the scan kernel takes: host blocks of whole tiles, 16-byte aligned rows
bool modtilt_scan_launch(const MTScanArgs& A, cudaStream_t stream, bool exact_maps = false);   // exact_maps: correctly rounded sqrt / div / log10 / 10^x
int modtilt_pick_lanes(int n_streams, int n_sm);   // streams per warp: 8, 16 or 32
bool modtilt_launch(const MTArgs& a, int lanes, cudaStream_t stream, int run = 8);   // run = frames per straight-line run

// ---- ERB Tilt (IIR class, hand-written scan kernel: k_erbtilt.cu / erbtilt_core.h) -------------------------------------
constexpr int kErbChunkF32 = 8, kErbChunkF64 = 4;       // frames per lane of a main tile (tile = 32 x chunk frames)
// state slots, [slot][stream] doubles: the script's mem[] arrays lpL, lpR, envF, envS, preE, postE, then block-rate variables
constexpr int kErbLpL = 0, kErbLpR = 15, kErbEnvF = 30, kErbEnvS = 46, kErbPreE = 62, kErbPostE = 78, kErbGlobalGain = 94,
              kErbGBand = 106, kErbSlots = 122;

// ---- Spectral Stabilizer (IIR class, hand-written scan kernel: k_specstab.cu / specstab_core.h) ---------------------------
constexpr int kSsChunkF32 = 8, kSsChunkF64 = 4;
// state slots, [slot][stream] doubles: ten per-band carries (12 bands each: z1L z2L z1R z2R envlin envlog d1 d2 mot gs), then the
// block-rate values gn[11] gw[11] res[12] gt[12] depth mb low_min_gain action_amt
constexpr int kSsCarry = 0, kSsGn = 120, kSsGw = 131, kSsRes = 142, kSsGt = 154, kSsScalars = 166, kSsSlots = 170;

// ---- translated DSP-JSFX leaves (csrc/gen/g_*.cu from za_gpu/jsfxc.py, runtime in jsfx_rt.cuh) -------------------
struct GenLaunch {
    const float* in; float* out; double* tap;      // [stream][ch][row_stride]
    int64_t row_stride, n_frames, mem_slots;
    int n_streams, block_len, n_rows, n_decl;
    const double* q;                               // device [n_rows][n_streams][n_decl] quantised slider values
    const int* decl_index0;                        // device [n_decl]: slider number - 1 of each column
    void* state;                                   // [n_vars + 64 + 1][n_streams] of R: vars, st.sliders, lastSlidersValid
    void* mem;                                     // [mem_slots][n_streams] of R: the script's mem[] heap
    unsigned* mt;                                  // [625][n_streams]: MT19937 words + index
    int* err;                                      // device flag: 1 = mem[] access beyond mem_slots
    double srate;
    int64_t fft_scratch;                           // slots of FFT scratch per stream behind the heap (0: the script calls no FFT hook)
    const double* tw;                              // device: e^{-2 pi i k / 32768}, k < 16384, (cos, -sin) pairs
    int lanes;                                     // streams per warp of jsfx_process_kernel: 8 / 16 / 32, 0 = picked at launch
    int unstaged;                                  // leaves with a stage pipeline: 0 = picked by batch size, 1 = one thread per stream, 2 = pipeline (option "unstaged")
};
struct GenDesc {
    const char* key; const char* slider_text;
    int n_in, n_out, n_vars; int64_t mem_slots;
    const char* const* var_names;                  // n_vars names (DSPJSFX_VARS order of this translation), null-terminated
    const int* alias;                              // [64]: vars slot mirrored by slider i0, or -1
    bool (*prepare)(const GenLaunch&, int precision, cudaStream_t);
    bool (*process)(const GenLaunch&, int precision, cudaStream_t);
    int n_decl; const int* decl_index0;            // declared sliders the generated code was compiled for (declaration order)
    int64_t fft_scratch;                           // FFT scratch slots per stream this script needs (2 x its largest transform)
};

}  // namespace za
