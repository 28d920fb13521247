// za_capi.cu -- the C ABI of include/za_gpu.h: catalogue, instance life cycle, and the batch
// entry points that mirror prepareToPlay / processBlock (src/JSFXJuceProcessor.cpp:3239-3342,
// 3435-3772) over n_streams instances.  No CPU fallback: every compute call needs a CUDA device.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "za_host.h"

namespace za {

// ---- errors -------------------------------------------------------------------------------------
static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
bool cuda_ok(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return true;
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return false;
}

// ---- tensor maps (driver entry point resolved at run time; libcuda is not linked) ---------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

bool make_chunk_tensor_map(CUtensorMap* map, const float* base, int64_t n_chunks, int64_t n_rows,
                           int64_t row_stride_frames, int box_rows)
{
    EncodeTiledFn fn = encode_tiled();
    if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return false; }
    const cuuint64_t dims[3] = {32, (cuuint64_t)n_chunks, (cuuint64_t)n_rows};
    const cuuint64_t strides[2] = {128, (cuuint64_t)row_stride_frames * 4};   // bytes, dims 1 and 2
    const cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[160];
        std::snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled failed (%d): chunks=%lld rows=%lld stride=%lld", (int)r,
                      (long long)n_chunks, (long long)n_rows, (long long)row_stride_frames);
        set_error(buf);
        return false;
    }
    return true;
}

// ---- catalogue ----------------------------------------------------------------------------------
const char* cascade8_slider_text();
const char* easyexpander_slider_text();
const char* stft_mask2048_slider_text();
const char* modtilt_slider_text();
const char* erbtilt_slider_text();
const char* specstab_slider_text();
PluginImpl* make_specstab();
const char* red_slider_text();
const char* clickbegonesg_slider_text();
const char* var_slider_text();
const char* gts_slider_text();
PluginImpl* make_red();
PluginImpl* make_clickbegonesg();
PluginImpl* make_var();
PluginImpl* make_gts();

void append_generated_plugins(std::vector<PluginDesc>& t);    // p_generic.cu: the translated DSP-JSFX leaves

const std::vector<PluginDesc>& plugin_table()
{
    static const std::vector<PluginDesc> t = [] {
        std::vector<PluginDesc> v = {
            {"cascade8", 1, 1, 2, 0, cascade8_slider_text(), &make_cascade8},
            {"EasyExpander", 2, 2, 0, 0, easyexpander_slider_text(), &make_easyexpander},
            {"stft_mask2048", 2, 2, 2, 2048, stft_mask2048_slider_text(), &make_stft_mask2048},
            {"ModTilt", 2, 2, 1, 0, modtilt_slider_text(), &make_modtilt},
            {"ERBTilt", 2, 2, 0, 0, erbtilt_slider_text(), &make_erbtilt},
            {"SpectralStabilizer", 2, 2, 0, 0, specstab_slider_text(), &make_specstab},
            {"RED", 6, 6, 1, 0, red_slider_text(), &make_red},
            {"ClickBeGoneSG", 2, 2, 1, 15, clickbegonesg_slider_text(), &make_clickbegonesg},
            {"VAR", 2, 2, 1, 0, var_slider_text(), &make_var},
            {"GTS", 2, 2, 1, 128, gts_slider_text(), &make_gts},
        };
        append_generated_plugins(v);
        return v;
    }();
    return t;
}
const PluginDesc* find_plugin(const char* key)
{
    if (!key) return nullptr;
    for (const auto& p : plugin_table())
        if (std::strcmp(p.key, key) == 0) return &p;
    return nullptr;
}

// ---- checkpoint visitor ---------------------------------------------------------------------------
void StateIO::dev(void* d_ptr, size_t bytes)
{
    if (mode == Save && ok) ok = cuda_ok(cudaMemcpyAsync(blob + pos, d_ptr, bytes, cudaMemcpyDeviceToHost, stream), "state save");
    if (mode == Load && ok) ok = cuda_ok(cudaMemcpyAsync(d_ptr, blob + pos, bytes, cudaMemcpyHostToDevice, stream), "state load");
    pos += (bytes + 15) & ~(size_t)15;
}
void StateIO::host(void* h_ptr, size_t bytes)
{
    if (mode == Save) std::memcpy(blob + pos, h_ptr, bytes);
    if (mode == Load) std::memcpy(h_ptr, blob + pos, bytes);
    pos += (bytes + 15) & ~(size_t)15;
}

struct StateHeader {              // 64 bytes; everything a blob must agree on with the handle it is loaded into
    char magic[8];                // "ZAGPUST1"
    char key[32];
    int32_t n_streams, block_len, precision, n_decl;
    double srate;
};

static size_t state_walk(Instance& I, StateIO& io)
{
    io.pos = sizeof(StateHeader);
    io.host(I.sliders.data(), I.sliders.size() * sizeof(double));
    io.host(I.sliders_valid.data(), I.sliders_valid.size());
    io.host(&I.prepared, sizeof(I.prepared));
    I.impl->state_io(I, io);
    return io.pos;
}

// ---- kernel timing ------------------------------------------------------------------------------
bool Instance::time_begin()
{
    if (ev_used + 2 > 2048) return false;
    while ((int)ev.size() < ev_used + 2) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return false;
        ev.push_back(e);
    }
    cudaEventRecord(ev[ev_used], stream);
    return true;
}
void Instance::time_end()
{
    cudaEventRecord(ev[ev_used + 1], stream);
    ev_used += 2;
}

// ---- processBlock over a stream range with device buffers ----------------------------------------
// params: [n_param_blocks][n_streams][n_decl] host floats (may be null with n_param_blocks == 0).
// in/out/tap point at stream s0's first row.
static bool process_range_engine(Instance& I, const float* params, int n_param_blocks, int s0, int ns, const float* in,
                                 float* out, double* tap, int64_t n_frames, int64_t stride)
{
    const int nd = (int)I.decls.size();
    const int64_t bl = I.block_len;
    const int64_t n_blocks = (n_frames + bl - 1) / bl;
    if (I.impl->takes_block_params()) {
        // quantise the whole automation trajectory once (pushParamsToStateSliders per host block, :9286-9357)
        const int nbq = (params && n_param_blocks > 0) ? (int)std::min<int64_t>(n_param_blocks, n_blocks) : 1;
        if (params && n_param_blocks > 0 && nbq > 1 && s0 == 0 && ns == I.n_streams) {
            const int r = I.impl->run_blocks_raw(I, params, nbq, in, out, tap, n_frames, stride);
            if (r < 0) return false;
            if (r > 0) {   // st.sliders after the call = the last row, quantised
                for (int s = 0; s < ns; ++s) {
                    for (int j = 0; j < nd; ++j)
                        I.sliders[(size_t)s * nd + j] = quantise_slider(I.decls[j], params[((size_t)(nbq - 1) * ns + s) * nd + j]);
                    I.sliders_valid[s] = 1;
                }
                return true;
            }
        }
        std::vector<double> q((size_t)nbq * ns * nd);
        parallel_for(nbq, 8, [&](int64_t b0, int64_t b1) {
            for (int64_t b = b0; b < b1; ++b)
                for (int s = 0; s < ns; ++s)
                    for (int j = 0; j < nd; ++j)
                        q[((size_t)b * ns + s) * nd + j] = (params && n_param_blocks > 0)
                            ? quantise_slider(I.decls[j], params[((size_t)b * I.n_streams + s0 + s) * nd + j])
                            : I.sliders[(size_t)(s0 + s) * nd + j];
        });
        for (int s = 0; s < ns; ++s) {
            for (int j = 0; j < nd; ++j) I.sliders[(size_t)(s0 + s) * nd + j] = q[((size_t)(nbq - 1) * ns + s) * nd + j];
            I.sliders_valid[s0 + s] = 1;
        }
        return I.impl->run_blocks(I, s0, ns, q.data(), nbq, in, out, tap, n_frames, stride);
    }
    // plugins without per-block coefficient rows in their throughput kernels may still render an automated call in one go
    if (params && n_param_blocks > 1 && n_blocks > 1 && s0 == 0 && ns == I.n_streams) {
        const int nbq = (int)std::min<int64_t>(n_param_blocks, n_blocks);
        const int r = I.impl->run_blocks_raw(I, params, nbq, in, out, tap, n_frames, stride);
        if (r < 0) return false;
        if (r > 0) {   // st.sliders after the call = the last row, quantised
            for (int s = 0; s < ns; ++s) {
                for (int j = 0; j < nd; ++j)
                    I.sliders[(size_t)s * nd + j] = quantise_slider(I.decls[j], params[((size_t)(nbq - 1) * ns + s) * nd + j]);
                I.sliders_valid[s] = 1;
            }
            return true;
        }
    }
    std::vector<int> changed;
    int64_t run_start = 0;     // first frame of the pending run
    for (int64_t b = 0; b < n_blocks; ++b) {
        changed.clear();
        // pushParamsToStateSliders (:9286-9357): only rows that can differ need a look
        if (params && n_param_blocks > 0 && (b < n_param_blocks)) {
            const float* row = params + ((size_t)b * I.n_streams + s0) * nd;
            for (int s = 0; s < ns; ++s) {
                double* cur = &I.sliders[(size_t)(s0 + s) * nd];
                bool ch = false;
                for (int j = 0; j < nd; ++j) {
                    const double v = quantise_slider(I.decls[j], row[(size_t)s * nd + j]);
                    if (!I.sliders_valid[s0 + s] || v != cur[j]) { ch = true; cur[j] = v; }
                }
                I.sliders_valid[s0 + s] = 1;
                if (ch) changed.push_back(s0 + s);
            }
        }
        if (!changed.empty()) {
            const int64_t f = b * bl;
            if (f > run_start) {   // flush the run rendered with the old parameters
                if (!I.impl->run(I, s0, ns, in + run_start, out + run_start, tap ? tap + run_start : nullptr, f - run_start, stride))
                    return false;
                run_start = f;
            }
            if (!I.impl->sliders_changed(I, s0, ns, changed)) return false;   // @slider (:3545-3547)
        }
        if (!(params && n_param_blocks > 0 && b + 1 < n_param_blocks)) break;  // nothing can change any more
    }
    if (n_frames > run_start)
        if (!I.impl->run(I, s0, ns, in + run_start, out + run_start, tap ? tap + run_start : nullptr, n_frames - run_start, stride))
            return false;
    return true;
}

// ---- oversampling x2 / x4 / x8 around any plugin (src/JSFXJuceProcessor.cpp:8704-8789) ---------------------------------
// Per HOST block: linear up-sampling inside the block (the block's last frame is held: `b = (i + 1 < hostSamples) ? src[i + 1]
// : a`, float arithmetic a + (b - a) * ((float) sub / (float) factor)), the engine renders block_len x factor frames at srate x
// factor, then a box average back (sum in double over `factor` frames / factor -> float).
__global__ void __launch_bounds__(256) os_up_kernel(const float* __restrict__ in, int64_t in_stride, float* __restrict__ out, int64_t out_stride,
                                                   int64_t n_host, int host_block, int factor)
{
    const float* src = in + (size_t)blockIdx.y * in_stride;
    float* dst = out + (size_t)blockIdx.y * out_stride;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_host * factor; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = e / factor;
        const int sub = (int)(e - i * factor);
        const bool last = (i + 1 >= n_host) || ((i + 1) % host_block == 0);
        const float a = src[i], b = last ? a : src[i + 1];
        const float t = __fdiv_rn((float)sub, (float)factor);
        dst[e] = __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), t));
    }
}
__global__ void __launch_bounds__(256) os_down_kernel(const float* __restrict__ eng, int64_t eng_stride, float* __restrict__ out, int64_t out_stride,
                                                     int64_t n_host, int factor)
{
    const float* src = eng + (size_t)blockIdx.y * eng_stride;
    float* dst = out + (size_t)blockIdx.y * out_stride;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_host; i += (int64_t)gridDim.x * blockDim.x) {
        double sum = 0.0;
        for (int sub = 0; sub < factor; ++sub) sum = __dadd_rn(sum, (double)src[i * factor + sub]);
        dst[i] = (float)__ddiv_rn(sum, (double)factor);
    }
}

static bool process_range(Instance& I, const float* params, int n_param_blocks, int s0, int ns, const float* in,
                          float* out, double* tap, int64_t n_frames, int64_t stride)
{
    if (I.oversample <= 1) return process_range_engine(I, params, n_param_blocks, s0, ns, in, out, tap, n_frames, stride);
    if (tap) { set_error("za_gpu_process_tap64: no double tap on an oversampled handle (the host-rate output is a box average of floats)"); return false; }
    const int f = I.oversample, n_in = I.desc->n_in, n_out = I.desc->n_out, nd = (int)I.decls.size();
    const int64_t hb = I.host_block, chunk_blocks = 64, cf = hb * chunk_blocks;      // host frames per chunk
    const int64_t need = (cf * f + 31) / 32 * 32;
    if (need != I.os_stride) {
        cudaFree(I.os_in); cudaFree(I.os_out); I.os_in = I.os_out = nullptr; I.os_stride = 0;
        if (!cuda_ok(cudaMalloc(&I.os_in, (size_t)I.n_streams * std::max(n_in, 1) * need * 4), "oversampling input") ||
            !cuda_ok(cudaMalloc(&I.os_out, (size_t)I.n_streams * n_out * need * 4), "oversampling output")) return false;
        I.os_stride = need;
    }
    for (int64_t f0 = 0; f0 < n_frames; f0 += cf) {
        const int64_t nh = std::min<int64_t>(cf, n_frames - f0);
        const unsigned gx = (unsigned)std::min<int64_t>((nh * f + 255) / 256, 32);
        if (n_in > 0) {
            os_up_kernel<<<dim3(gx, (unsigned)(ns * n_in)), 256, 0, I.stream>>>(in + f0, stride, I.os_in, need, nh, (int)hb, f);
            if (!cuda_ok(cudaGetLastError(), "os_up_kernel")) return false;
            ++I.launches;
        }
        const int64_t b0 = f0 / hb;
        const float* p = nullptr; int np = 0;
        if (params && n_param_blocks > 0) {
            const int64_t r0 = std::min<int64_t>(b0, n_param_blocks - 1);
            p = params + (size_t)r0 * I.n_streams * nd; np = (int)(n_param_blocks - r0);
        }
        if (!process_range_engine(I, p, np, s0, ns, I.os_in, I.os_out, nullptr, nh * f, need)) return false;
        os_down_kernel<<<dim3(gx, (unsigned)(ns * n_out)), 256, 0, I.stream>>>(I.os_out, need, out + f0, stride, nh, f);
        if (!cuda_ok(cudaGetLastError(), "os_down_kernel")) return false;
        ++I.launches;
    }
    return true;
}

static void free_staging(Instance& I)
{
    for (int i = 0; i < 3; ++i) {
        cudaFree(I.stage_in[i]); cudaFree(I.stage_out[i]); cudaFree(I.stage_tap[i]);
        I.stage_in[i] = I.stage_out[i] = nullptr; I.stage_tap[i] = nullptr;
    }
    I.stage_in_bytes = I.stage_out_bytes = I.stage_tap_bytes = 0; I.stage_stride = 0;
}

// Host buffers: TIME chunks of all streams flow H2D -> kernels -> D2H through three staging slots on
// three CUDA streams, so PCIe traffic in both directions overlaps the kernels.  Chunks are whole host
// blocks (the block grid of the call is kept) and every chunk launches over all streams, so the
// kernels see the same grid shapes as a device-resident render; state carries from chunk to chunk.
static bool process_host(Instance& I, const float* params, int npb, const float* in, float* out, double* tap,
                         int64_t n_frames, int64_t stride)
{
    const int n_in = I.desc->n_in, n_out = I.desc->n_out, nd = (int)I.decls.size();
    const int64_t bl = I.host_block;
    int64_t cf;                                                        // frames per chunk
    if (I.opt_host_chunk > 0) {
        cf = (int64_t)I.opt_host_chunk;
    } else {
        const int64_t per_frame = (int64_t)I.n_streams * (4 * (n_in + n_out) + (tap ? 8 * n_out : 0));
        cf = (384ll << 20) / std::max<int64_t>(per_frame, 1);
    }
    cf = std::max<int64_t>(bl, cf / bl * bl);
    cf = std::min<int64_t>(cf, (n_frames + bl - 1) / bl * bl);
    const int64_t dstride = (cf + 31) / 32 * 32;                       // device rows stay TMA-aligned
    const size_t need_in = (size_t)I.n_streams * n_in * dstride * 4, need_out = (size_t)I.n_streams * n_out * dstride * 4;
    const size_t need_tap = tap ? (size_t)I.n_streams * n_out * dstride * 8 : 0;
    if (need_in > I.stage_in_bytes || need_out > I.stage_out_bytes || need_tap > I.stage_tap_bytes || dstride != I.stage_stride) {
        free_staging(I);
        for (int i = 0; i < 3; ++i) {
            if (!cuda_ok(cudaMalloc(&I.stage_in[i], need_in), "staging in")) return false;
            if (!cuda_ok(cudaMalloc(&I.stage_out[i], need_out), "staging out")) return false;
            if (need_tap && !cuda_ok(cudaMalloc(&I.stage_tap[i], need_tap), "staging tap")) return false;
        }
        I.stage_in_bytes = need_in; I.stage_out_bytes = need_out; I.stage_tap_bytes = need_tap; I.stage_stride = dstride;
    }
    if (!I.h2d) {
        if (!cuda_ok(cudaStreamCreateWithFlags(&I.h2d, cudaStreamNonBlocking), "h2d stream")) return false;
        if (!cuda_ok(cudaStreamCreateWithFlags(&I.d2h, cudaStreamNonBlocking), "d2h stream")) return false;
        for (int i = 0; i < 3; ++i) {
            cudaEventCreateWithFlags(&I.ev_h2d[i], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&I.ev_comp[i], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&I.ev_d2h[i], cudaEventDisableTiming);
        }
    }
    // On any failure the copies already enqueued still target the caller's buffers: drain the three streams before
    // returning, so `out` is never written after the call has reported its error.
    struct Drain {
        Instance& I; bool armed = true;
        ~Drain() { if (armed) { cudaStreamSynchronize(I.h2d); cudaStreamSynchronize(I.stream); cudaStreamSynchronize(I.d2h); } }
    } drain{I};
    int c = 0;
    for (int64_t f0 = 0; f0 < n_frames; f0 += cf, ++c) {
        const int64_t nf = std::min<int64_t>(cf, n_frames - f0);
        const int slot = c % 3;
        if (c >= 3) cudaStreamWaitEvent(I.h2d, I.ev_d2h[slot], 0);     // slot drained
        if (!cuda_ok(cudaMemcpy2DAsync(I.stage_in[slot], dstride * 4, in + f0, stride * 4, nf * 4,
                                       (size_t)I.n_streams * n_in, cudaMemcpyHostToDevice, I.h2d), "H2D"))
            return false;
        cudaEventRecord(I.ev_h2d[slot], I.h2d);
        cudaStreamWaitEvent(I.stream, I.ev_h2d[slot], 0);
        if (c >= 3) cudaStreamWaitEvent(I.stream, I.ev_d2h[slot], 0);  // its previous output has left
        // automation rows of this chunk: row b of the call applies to host block b (the last row persists)
        const int64_t b0 = f0 / bl;
        const float* p = nullptr;
        int np = 0;
        if (params && npb > 0 && b0 < npb) { p = params + (size_t)b0 * I.n_streams * nd; np = (int)(npb - b0); }
        // "copy_only" (bench: the box's H2D + D2H ceiling for this exact copy pattern) skips the kernels
        if (!I.opt_copy_only &&
            !process_range(I, p, np, 0, I.n_streams, I.stage_in[slot], I.stage_out[slot], tap ? I.stage_tap[slot] : nullptr,
                           nf, dstride))
            return false;
        cudaEventRecord(I.ev_comp[slot], I.stream);
        cudaStreamWaitEvent(I.d2h, I.ev_comp[slot], 0);
        if (!cuda_ok(cudaMemcpy2DAsync(out + f0, stride * 4, I.stage_out[slot], dstride * 4, nf * 4,
                                       (size_t)I.n_streams * n_out, cudaMemcpyDeviceToHost, I.d2h), "D2H"))
            return false;
        if (tap && !cuda_ok(cudaMemcpy2DAsync(tap + f0, stride * 8, I.stage_tap[slot], dstride * 8, nf * 8,
                                              (size_t)I.n_streams * n_out, cudaMemcpyDeviceToHost, I.d2h), "D2H tap"))
            return false;
        cudaEventRecord(I.ev_d2h[slot], I.d2h);
    }
    drain.armed = false;
    const bool ok_d2h = cuda_ok(cudaStreamSynchronize(I.d2h), "D2H sync");
    return cuda_ok(cudaStreamSynchronize(I.stream), "compute sync") && ok_d2h;
}

static int do_process(Instance* I, const float* params, int npb, const float* in, float* out, double* tap,
                      int64_t n_frames, int64_t stride, int mem)
{
    if (!I || (!in && I->desc->n_in > 0) || !out || n_frames < 0 || stride < n_frames || npb < 0 || (npb > 0 && !params)) {
        set_error("za_gpu_process: bad argument");
        return ZA_ERR_ARG;
    }
    if (!I->prepared) { set_error("za_gpu_process before za_gpu_prepare"); return ZA_ERR_STATE; }
    if (tap && I->precision != ZA_F64) { set_error("za_gpu_process_tap64 needs precision ZA_F64"); return ZA_ERR_ARG; }
    if (I->desc->n_in != I->desc->n_out && in && n_frames > 0) {
        // rows of stream s sit at (s * n_ch + c) * stride: with different channel counts the two layouts interleave, so
        // overlapping buffers would let one stream's output land in another stream's unread input
        const char* a0 = (const char*)in;  const char* a1 = a0 + ((size_t)I->n_streams * I->desc->n_in - 1) * stride * 4 + n_frames * 4;
        const char* b0 = (const char*)out; const char* b1 = b0 + ((size_t)I->n_streams * I->desc->n_out - 1) * stride * 4 + n_frames * 4;
        if (a0 < b1 && b0 < a1) { set_error("za_gpu_process: in and out overlap but the plugin's channel counts differ"); return ZA_ERR_ARG; }
    }
    if (!cuda_ok(cudaSetDevice(I->device), "cudaSetDevice")) return ZA_ERR_CUDA;
    I->ev_used = 0;
    if (I->impl->deferred_error(*I)) return ZA_ERR_STATE;           // flagged by a kernel of an earlier call
    if (n_frames == 0) return ZA_OK;
    bool ok;
    if (mem == ZA_MEM_DEVICE) ok = process_range(*I, params, npb, 0, I->n_streams, in, out, tap, n_frames, stride);
    else if (mem == ZA_MEM_HOST) ok = process_host(*I, params, npb, in, out, tap, n_frames, stride);
    else { set_error("za_gpu_process: unknown memory kind"); return ZA_ERR_ARG; }
    if (ok && mem == ZA_MEM_HOST && I->impl->deferred_error(*I)) return ZA_ERR_STATE;   // host calls have synchronised
    return ok ? ZA_OK : ZA_ERR_CUDA;
}

}  // namespace za

using namespace za;

struct za_gpu_instance { Instance I; };

extern "C" {

ZA_API const char* za_gpu_last_error(void) { return g_err.c_str(); }
ZA_API const char* za_gpu_version(void) { return "za_gpu 0.1 (sm_100a)"; }
ZA_API int za_gpu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

ZA_API int za_gpu_plugin_count(void) { return (int)plugin_table().size(); }
ZA_API const char* za_gpu_plugin_key(int i)
{
    return (i >= 0 && i < (int)plugin_table().size()) ? plugin_table()[i].key : nullptr;
}
ZA_API int za_gpu_plugin_info(const char* key, za_plugin_info* out)
{
    const PluginDesc* d = find_plugin(key);
    if (!d) { set_error(std::string("unknown plugin_key: ") + (key ? key : "(null)")); return ZA_ERR_PLUGIN; }
    if (!out) return ZA_ERR_ARG;
    out->n_in = d->n_in; out->n_out = d->n_out; out->kind = d->kind; out->latency = d->latency;
    out->n_sliders = (int)parse_slider_decls(d->slider_text).size();
    return ZA_OK;
}
static void fill_slider(const SliderDecl& s, za_slider_info* o)
{
    std::memset(o, 0, sizeof *o);
    o->index0 = s.index0; o->is_choice = s.is_choice ? 1 : 0;
    o->def = s.def; o->min = s.min; o->max = s.max; o->step = s.step;
    std::strncpy(o->alias, s.alias.c_str(), sizeof(o->alias) - 1);
}
ZA_API int za_gpu_slider_info(const char* key, int column, za_slider_info* out)
{
    const PluginDesc* d = find_plugin(key);
    if (!d) { set_error(std::string("unknown plugin_key: ") + (key ? key : "(null)")); return ZA_ERR_PLUGIN; }
    const auto decls = parse_slider_decls(d->slider_text);
    if (!out || column < 0 || column >= (int)decls.size()) return ZA_ERR_ARG;
    fill_slider(decls[column], out);
    return ZA_OK;
}
ZA_API int za_gpu_parse_sliders(const char* text, za_slider_info* out, int cap)
{
    if (!text || cap < 0 || (cap > 0 && !out)) return ZA_ERR_ARG;
    const auto decls = parse_slider_decls(text);
    for (int i = 0; i < (int)decls.size() && i < cap; ++i) fill_slider(decls[i], &out[i]);
    return (int)decls.size();
}
ZA_API int za_gpu_quantise(const za_slider_info* s, float host_value, double* out)
{
    if (!s || !out) return ZA_ERR_ARG;
    SliderDecl d;
    d.index0 = s->index0; d.is_choice = s->is_choice != 0; d.def = s->def; d.min = s->min; d.max = s->max; d.step = s->step;
    *out = quantise_slider(d, host_value);
    return ZA_OK;
}

ZA_API int za_gpu_create(const char* key, int n_streams, double srate, int block_len, int precision, int device,
                         za_gpu_handle* out)
{
    return za_gpu_create_os(key, n_streams, srate, block_len, precision, device, 1, out);
}

ZA_API int za_gpu_create_os(const char* key, int n_streams, double srate, int block_len, int precision, int device, int oversampling,
                            za_gpu_handle* out)
{
    if (!out) return ZA_ERR_ARG;
    *out = nullptr;
    if (oversampling != 1 && oversampling != 2 && oversampling != 4 && oversampling != 8) { set_error("za_gpu_create_os: oversampling must be 1, 2, 4 or 8"); return ZA_ERR_ARG; }
    if (block_len > 0 && block_len > (1 << 28) / oversampling) { set_error("za_gpu_create_os: block too long"); return ZA_ERR_ARG; }
    const PluginDesc* d = find_plugin(key);
    if (!d) { set_error(std::string("unknown plugin_key: ") + (key ? key : "(null)")); return ZA_ERR_PLUGIN; }
    if (n_streams <= 0 || block_len <= 0 || !(srate > 0) || (precision != ZA_F32 && precision != ZA_F64)) {
        set_error("za_gpu_create: bad argument");
        return ZA_ERR_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        set_error("no CUDA device: this library has no CPU fallback");
        return ZA_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { set_error("za_gpu_create: bad device index"); return ZA_ERR_ARG; }
    if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice")) return ZA_ERR_CUDA;
    auto* h = new za_gpu_instance();
    Instance& I = h->I;
    I.desc = d; I.decls = parse_slider_decls(d->slider_text);
    // the engine (every plugin back end) sees the oversampled rate and block length (reinitialiseJsfxForCurrentEngineRate, :8791-8800)
    I.oversample = oversampling; I.host_block = block_len;
    I.n_streams = n_streams; I.block_len = block_len * oversampling; I.precision = precision; I.device = device; I.srate = srate * oversampling;
    I.sliders.assign((size_t)n_streams * I.decls.size(), 0.0);
    for (int s = 0; s < n_streams; ++s)
        for (size_t j = 0; j < I.decls.size(); ++j) I.sliders[(size_t)s * I.decls.size() + j] = (double)I.decls[j].def;
    I.sliders_valid.assign(n_streams, 0);
    if (!cuda_ok(cudaStreamCreateWithFlags(&I.own_stream, cudaStreamNonBlocking), "stream create")) { delete h; return ZA_ERR_CUDA; }
    I.stream = I.own_stream;
    I.impl.reset(d->make());
    if (!I.impl->create(I)) { const int rc = I.create_status; za_gpu_destroy(h); return rc; }
    *out = h;
    return ZA_OK;
}

ZA_API int za_gpu_destroy(za_gpu_handle h)
{
    if (!h) return ZA_OK;
    Instance& I = h->I;
    cudaSetDevice(I.device);
    cudaDeviceSynchronize();
    I.impl.reset();
    free_staging(I);
    cudaFree(I.os_in); cudaFree(I.os_out);
    for (auto e : I.ev) cudaEventDestroy(e);
    for (int i = 0; i < 3; ++i) {
        if (I.ev_h2d[i]) cudaEventDestroy(I.ev_h2d[i]);
        if (I.ev_comp[i]) cudaEventDestroy(I.ev_comp[i]);
        if (I.ev_d2h[i]) cudaEventDestroy(I.ev_d2h[i]);
    }
    if (I.h2d) cudaStreamDestroy(I.h2d);
    if (I.d2h) cudaStreamDestroy(I.d2h);
    if (I.own_stream) cudaStreamDestroy(I.own_stream);
    delete h;
    return ZA_OK;
}

ZA_API int za_gpu_set_stream(za_gpu_handle h, void* s)
{
    if (!h) return ZA_ERR_ARG;
    h->I.stream = s ? (cudaStream_t)s : h->I.own_stream;
    return ZA_OK;
}

ZA_API int za_gpu_prepare(za_gpu_handle h, const float* p0)
{
    if (!h) return ZA_ERR_ARG;
    Instance& I = h->I;
    if (!cuda_ok(cudaSetDevice(I.device), "cudaSetDevice")) return ZA_ERR_CUDA;
    const int nd = (int)I.decls.size();
    // resetStateStructOnly + lastSlidersValid=false (:3250, 3298), push sliders BEFORE @init (:3297-3303)
    std::vector<int> all(I.n_streams);
    for (int s = 0; s < I.n_streams; ++s) {
        all[s] = s;
        for (int j = 0; j < nd; ++j) {
            const float hv = p0 ? p0[(size_t)s * nd + j] : I.decls[j].def;
            I.sliders[(size_t)s * nd + j] = quantise_slider(I.decls[j], hv);
        }
        I.sliders_valid[s] = 1;
    }
    if (!I.impl->reset(I)) return ZA_ERR_CUDA;                       // @init
    if (!I.impl->sliders_changed(I, 0, I.n_streams, all)) return ZA_ERR_CUDA;   // alias re-apply + @slider (:3308-3318)
    I.prepared = true;
    return ZA_OK;
}

ZA_API int za_gpu_process(za_gpu_handle h, const float* params, int npb, const float* in, float* out, int64_t n_frames,
                          int64_t stride, int mem)
{
    return do_process(h ? &h->I : nullptr, params, npb, in, out, nullptr, n_frames, stride, mem);
}
ZA_API int za_gpu_process_tap64(za_gpu_handle h, const float* params, int npb, const float* in, float* out, double* tap,
                                int64_t n_frames, int64_t stride, int mem)
{
    if (!tap) { set_error("za_gpu_process_tap64: tap64 is null"); return ZA_ERR_ARG; }
    return do_process(h ? &h->I : nullptr, params, npb, in, out, tap, n_frames, stride, mem);
}

ZA_API int za_gpu_read_state(za_gpu_handle h, const char* name, double* out)
{
    if (!h || !name || !out) return ZA_ERR_ARG;
    if (!cuda_ok(cudaSetDevice(h->I.device), "cudaSetDevice")) return ZA_ERR_CUDA;
    return h->I.impl->read_state(h->I, name, out) ? ZA_OK : ZA_ERR_ARG;
}
ZA_API int za_gpu_sync(za_gpu_handle h)
{
    if (!h) return ZA_ERR_ARG;
    cudaSetDevice(h->I.device);
    if (!cuda_ok(cudaStreamSynchronize(h->I.stream), "sync")) return ZA_ERR_CUDA;
    return h->I.impl && h->I.impl->deferred_error(h->I) ? ZA_ERR_STATE : ZA_OK;
}
ZA_API int za_gpu_state_size(za_gpu_handle h, size_t* bytes)
{
    if (!h || !bytes) return ZA_ERR_ARG;
    StateIO io; io.mode = StateIO::Size;
    *bytes = state_walk(h->I, io);
    return ZA_OK;
}
ZA_API int za_gpu_state_save(za_gpu_handle h, void* blob, size_t bytes)
{
    if (!h || !blob) return ZA_ERR_ARG;
    Instance& I = h->I;
    if (!cuda_ok(cudaSetDevice(I.device), "cudaSetDevice")) return ZA_ERR_CUDA;
    StateIO sz; sz.mode = StateIO::Size;
    if (bytes < state_walk(I, sz)) { set_error("za_gpu_state_save: blob too small (za_gpu_state_size)"); return ZA_ERR_ARG; }
    if (I.impl->deferred_error(I)) return ZA_ERR_STATE;
    StateHeader hd{};
    std::memcpy(hd.magic, "ZAGPUST1", 8);
    std::strncpy(hd.key, I.desc->key, sizeof(hd.key) - 1);
    hd.n_streams = I.n_streams; hd.block_len = I.block_len; hd.precision = I.precision; hd.n_decl = (int)I.decls.size(); hd.srate = I.srate;
    std::memcpy(blob, &hd, sizeof hd);
    StateIO io; io.mode = StateIO::Save; io.blob = (char*)blob; io.stream = I.stream;
    state_walk(I, io);
    if (!io.ok || !cuda_ok(cudaStreamSynchronize(I.stream), "state save sync")) return ZA_ERR_CUDA;
    return ZA_OK;
}
ZA_API int za_gpu_state_load(za_gpu_handle h, const void* blob, size_t bytes)
{
    if (!h || !blob || bytes < sizeof(StateHeader)) return ZA_ERR_ARG;
    Instance& I = h->I;
    if (!cuda_ok(cudaSetDevice(I.device), "cudaSetDevice")) return ZA_ERR_CUDA;
    StateHeader hd;
    std::memcpy(&hd, blob, sizeof hd);
    StateIO sz; sz.mode = StateIO::Size;
    if (std::memcmp(hd.magic, "ZAGPUST1", 8) != 0 || std::strncmp(hd.key, I.desc->key, sizeof hd.key) != 0 || hd.n_streams != I.n_streams ||
        hd.block_len != I.block_len || hd.precision != I.precision || hd.n_decl != (int)I.decls.size() || hd.srate != I.srate ||
        bytes < state_walk(I, sz)) {
        set_error("za_gpu_state_load: the blob was saved from a different plugin / batch shape / precision");
        return ZA_ERR_ARG;
    }
    StateIO io; io.mode = StateIO::Load; io.blob = (char*)const_cast<void*>(blob); io.stream = I.stream;
    state_walk(I, io);
    if (!io.ok || !cuda_ok(cudaStreamSynchronize(I.stream), "state load sync")) return ZA_ERR_CUDA;
    return I.impl->after_state_load(I) ? ZA_OK : ZA_ERR_CUDA;
}
ZA_API int za_gpu_host_alloc(size_t bytes, void** out)
{
    if (!out || bytes == 0) return ZA_ERR_ARG;
    *out = nullptr;
    return cuda_ok(cudaHostAlloc(out, bytes, cudaHostAllocPortable), "cudaHostAlloc") ? ZA_OK : ZA_ERR_CUDA;
}
ZA_API int za_gpu_host_free(void* p)
{
    if (!p) return ZA_OK;
    return cuda_ok(cudaFreeHost(p), "cudaFreeHost") ? ZA_OK : ZA_ERR_CUDA;
}
ZA_API int za_gpu_launch_count(za_gpu_handle h, int64_t* n)
{
    if (!h || !n) return ZA_ERR_ARG;
    *n = h->I.launches;
    return ZA_OK;
}
ZA_API int za_gpu_last_kernel_ms(za_gpu_handle h, double* ms, int* launches)
{
    if (!h || !ms) return ZA_ERR_ARG;
    Instance& I = h->I;
    cudaSetDevice(I.device);
    if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync")) return ZA_ERR_CUDA;
    double tot = 0;
    for (int i = 0; i + 1 < I.ev_used; i += 2) {
        float t = 0;
        if (cudaEventElapsedTime(&t, I.ev[i], I.ev[i + 1]) == cudaSuccess) tot += t;
    }
    *ms = tot;
    if (launches) *launches = I.ev_used / 2;
    return ZA_OK;
}
ZA_API int za_gpu_set_option(za_gpu_handle h, const char* name, int value)
{
    if (!h || !name) return ZA_ERR_ARG;
    Instance& I = h->I;
    if (!std::strcmp(name, "force_generic_io")) I.opt_force_generic = value;
    else if (!std::strcmp(name, "warps_per_stream")) I.opt_warps = value;
    else if (!std::strcmp(name, "host_chunk_frames")) I.opt_host_chunk = value;
    else if (!std::strcmp(name, "tile_buffers")) I.opt_nbuf = value;
    else if (!std::strcmp(name, "scalar_f32")) I.opt_scalar_f32 = value;
    else if (!std::strcmp(name, "heap_slots")) I.opt_heap_slots = value;
    else if (!std::strcmp(name, "copy_only")) I.opt_copy_only = value;
    else if (!std::strcmp(name, "unstaged")) I.opt_unstaged = value;        // translated leaves with a stage pipeline: 1 = the one-thread-per-stream kernel (cross-check)
    else if (!std::strcmp(name, "lanes_per_warp")) I.opt_lanes = value;     // stream-per-thread kernels: 8 / 16 streams per warp instead of 32 (measurement)
    else if (!std::strcmp(name, "role_off")) I.opt_role_off = value;      // measurement only: warp roles of a pipeline kernel that skip their work
    else { set_error(std::string("unknown option ") + name); return ZA_ERR_ARG; }
    return ZA_OK;
}

ZA_API int za_gpu_render(const char* key, const float* params, const float* in, float* out, int n_streams, int64_t n_frames)
{
    za_gpu_handle h = nullptr;
    const PluginDesc* d = find_plugin(key);
    // Faust leaves compute in float32 (cmake/plugin/CMakeLists.txt:144 passes no -double); DSP-JSFX computes in double
    int rc = za_gpu_create(key, n_streams, 48000.0, 512, (d && d->kind == 1) ? ZA_F32 : ZA_F64, 0, &h);
    if (rc != ZA_OK) return rc;
    rc = za_gpu_prepare(h, params);
    if (rc == ZA_OK) rc = za_gpu_process(h, nullptr, 0, in, out, n_frames, n_frames, ZA_MEM_HOST);
    za_gpu_destroy(h);
    return rc;
}

}  // extern "C"
