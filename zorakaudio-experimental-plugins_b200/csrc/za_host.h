// za_host.h -- host-side runtime behind the C ABI: plugin registry, slider layout, instances.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/za_gpu.h"
#include "za_internal.h"

namespace za {

struct SliderDecl {          // JsfxSliderDecl, src/JSFXJuceProcessor.cpp:520-545 (numeric part)
    int index0 = 0;
    bool is_choice = false;
    float def = 0, min = 0, max = 1, step = 0.001f;
    std::string alias;
};

std::vector<SliderDecl> parse_slider_decls(const char* jsfx_text);
double quantise_slider(const SliderDecl& d, float host_value);

// Host-side data parallelism for the parameter path (quantiser + @slider over [block][stream] rows):
// f(begin, end) over [0, n) on up to 32 host threads; runs inline below `grain` items.
template <class F>
inline void parallel_for(int64_t n, int64_t grain, F&& f)
{
    unsigned hw = std::thread::hardware_concurrency();
    int64_t nt = std::min<int64_t>(hw ? hw : 1, 32);
    nt = std::min<int64_t>(nt, (n + grain - 1) / std::max<int64_t>(grain, 1));
    if (nt <= 1) { f((int64_t)0, n); return; }
    std::vector<std::thread> th;
    const int64_t per = (n + nt - 1) / nt;
    for (int64_t t = 0; t < nt; ++t) {
        const int64_t b = t * per, e = std::min<int64_t>(n, b + per);
        if (b >= e) break;
        th.emplace_back([&f, b, e] { f(b, e); });
    }
    for (auto& x : th) x.join();
}

struct Instance;

// Checkpoint visitor (za_gpu_state_size / save / load): a plugin lists the buffers that ARE its per-stream DSP state --
// device arrays and the few host scalars beside them (e.g. a hop phase) -- in a fixed order; derived data (coefficient
// tables, scan tables) is rebuilt from the restored sliders by after_state_load().
struct StateIO {
    enum Mode { Size, Save, Load } mode = Size;
    char* blob = nullptr;            // host blob cursor base
    size_t pos = 0;
    cudaStream_t stream = nullptr;
    bool ok = true;
    void dev(void* d_ptr, size_t bytes);
    void host(void* h_ptr, size_t bytes);
};

// Per-plugin device back end.  All state lives in SoA device buffers owned by the implementation.
// Streams [s0, s0+ns) may be driven independently of other ranges (host-memory pipelining).
class PluginImpl {
public:
    virtual ~PluginImpl() {}
    virtual bool create(Instance& inst) = 0;
    virtual bool reset(Instance& inst) = 0;                                            // memset + @init
    virtual bool sliders_changed(Instance& inst, int s0, int ns, const std::vector<int>& streams) = 0;  // @slider
    // frames [0, n) of rows starting at in/out (device pointers to stream s0, first frame of the run),
    // parameters constant over the run, split into host blocks of inst.block_len (@block per block).
    virtual bool run(Instance& inst, int s0, int ns, const float* in, float* out, double* tap,
                     int64_t n, int64_t stride) = 0;
    virtual bool read_state(Instance& inst, const char* name, double* out) = 0;
    virtual void state_io(Instance& inst, StateIO& io) = 0;                     // every buffer of the per-stream DSP state
    virtual bool after_state_load(Instance&) { return true; }                   // rebuild what is derived from the sliders
    // Whole-call form for plugins that take per-block automation on the device: q = quantised slider
    // values [nbq][ns][n_decl] (host doubles, stream range s0..s0+ns), row b applies from host block b on,
    // the last row persists.  One launch renders all n frames; @slider semantics (recompute derived
    // coefficients when a value changed, src/JSFXJuceProcessor.cpp:3545-3547) are the plugin's job.
    virtual bool takes_block_params() const { return false; }
    // Device-side parameter path: the raw host float rows [nb][n_streams][n_decl] of the call are uploaded
    // and quantised + turned into @slider results by a kernel.  Returns 1 = done, 0 = not applicable for
    // this call (the caller falls back to the host path), -1 = error.
    // A failure a kernel could only flag on the device (e.g. a translated leaf ran past its mem[] heap): true + error
    // text if one has been seen.  Polled without synchronising at the start of every process call and after every
    // synchronisation the C ABI performs, so the failure surfaces on that call or, for device-memory calls, on the next.
    virtual bool deferred_error(Instance&) { return false; }
    virtual int run_blocks_raw(Instance&, const float*, int, const float*, float*, double*, int64_t, int64_t) { return 0; }
    virtual bool run_blocks(Instance&, int, int, const double*, int, const float*, float*, double*, int64_t, int64_t) { return false; }
};

struct PluginDesc {
    const char* key;
    int n_in, n_out, kind, latency;
    const char* slider_text;                 // the plugin's slider declaration lines (its parameter API)
    PluginImpl* (*make)();
};

const std::vector<PluginDesc>& plugin_table();
const PluginDesc* find_plugin(const char* key);

struct Instance {
    const PluginDesc* desc = nullptr;
    std::vector<SliderDecl> decls;
    int n_streams = 0, block_len = 512, precision = 1, device = 0;
    double srate = 48000.0;
    cudaStream_t stream = nullptr, own_stream = nullptr, h2d = nullptr, d2h = nullptr;
    // oversampling (src/JSFXJuceProcessor.cpp:3126-3146, 8704-8789): the engine runs at srate x factor on host blocks of
    // block_len x factor frames; `srate` / `block_len` above are the ENGINE's, host_block the caller's block length
    int oversample = 1, host_block = 512;
    float* os_in = nullptr; float* os_out = nullptr; int64_t os_stride = 0;
    bool prepared = false;
    std::vector<double> sliders;             // [n_streams][n_decl] current quantised values (st.sliders)
    std::vector<uint8_t> sliders_valid;      // lastSlidersValid per stream
    std::unique_ptr<PluginImpl> impl;
    int64_t launches = 0;
    // options
    int opt_force_generic = 0, opt_warps = 0, opt_host_chunk = 0, opt_nbuf = 0, opt_scalar_f32 = 0, opt_heap_slots = 0, opt_copy_only = 0, opt_role_off = 0, opt_lanes = 0, opt_unstaged = 0;
    int create_status = ZA_ERR_CUDA;         // what za_gpu_create returns when impl->create() fails (a refusal sets ZA_ERR_UNSUPPORTED)
    // timing of the dominant kernel
    std::vector<cudaEvent_t> ev;             // pairs
    int ev_used = 0;
    bool time_begin();                       // records a start event on `stream` (false if pool exhausted)
    void time_end();
    // host-memory staging
    float* stage_in[3] = {nullptr, nullptr, nullptr};
    float* stage_out[3] = {nullptr, nullptr, nullptr};
    double* stage_tap[3] = {nullptr, nullptr, nullptr};
    size_t stage_in_bytes = 0, stage_out_bytes = 0, stage_tap_bytes = 0;
    int64_t stage_stride = 0;
    cudaEvent_t ev_h2d[3] = {}, ev_comp[3] = {}, ev_d2h[3] = {};
};

// Device copy of a call's parameter rows as quantised doubles [rows][n_sliders] (k_params.cu).
struct RawRows {
    float* d_raw = nullptr; double* d_q = nullptr; size_t cap = 0;      // cap counts values
    cudaStream_t up = nullptr; cudaEvent_t ev_ready = nullptr, ev_consumed = nullptr;
    std::vector<double> last;                                           // host-quantised rows currently on the device
    ~RawRows();
    const double* upload(Instance& I, const float* raw_host_rows, size_t rows);          // raw floats -> device quantiser; I.stream waits for it
    const double* upload_quantised(Instance& I, const double* q, size_t count);          // host-quantised doubles (skipped when unchanged)
    void consumed(Instance& I);                                                          // call after the kernel that read the rows was launched
    void invalidate() { last.clear(); }
};

PluginImpl* make_cascade8();
PluginImpl* make_erbtilt();
PluginImpl* make_easyexpander();
PluginImpl* make_stft_mask2048();
PluginImpl* make_modtilt();

}  // namespace za
