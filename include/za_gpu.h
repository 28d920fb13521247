/*
 * za_gpu.h -- C ABI of the B200 batch renderer (libza_gpu.so).
 *
 * Drop-in boundary for ONE path of ZorakAudio-Experimental-Plugins: the per-sample / per-block
 * DSP of a plugin instance (reference: generated `jsfx_init/slider/block/sample/process_block`,
 * dsp_jsfx_aot.py:5713-5905, driven by JSFXJuceProcessor::prepareToPlay / processBlock,
 * src/JSFXJuceProcessor.cpp:3239-3342, 3435-3772; Faust: FaustJuceProcessor.cpp:430-482),
 * rendered for thousands of independent instances ("streams") at once.
 *
 * Plain pointers and sizes only; no C++/torch types; never throws.  Every function returns
 * ZA_OK (0) or a negative za_status; za_gpu_last_error() gives the text for the calling thread.
 * There is NO CPU fallback: without a CUDA device every compute entry point returns ZA_ERR_CUDA.
 *
 * Buffer layout (both directions):  planar float32, [stream][channel][frame]; element
 * (s, c, i) lives at base[(s * n_ch + c) * frame_stride + i].  This is the reference's
 * `const float* const* in / float* const* out` (one pointer per channel, dsp_jsfx_aot.py:5716-5719)
 * made dense over streams.  in == out (in-place) is legal when n_in == n_out, as JUCE bus buffers
 * alias in the reference (src/JSFXJuceProcessor.cpp:3456-3528).
 *
 * Parameter block: float32 HOST values as the reference's parameter atomics hold them
 * (src/JSFXJuceProcessor.cpp:5564-5566): the plain slider value, or the choice index for
 * `step{A,B,C}` sliders.  Column j = j-th declared slider in slider-index order.  The library
 * clamps and step-quantises exactly as hostParameterToJsfxSliderValue (:5556-5597) once per
 * host block, runs @slider only when a quantised value changed (:3545-3547), and mirrors alias
 * variables (:9347-9352).
 */
#ifndef ZA_GPU_H
#define ZA_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define ZA_API __attribute__((visibility("default")))
#else
#define ZA_API
#endif

typedef struct za_gpu_instance* za_gpu_handle;

typedef enum za_status {
    ZA_OK = 0,
    ZA_ERR_ARG = -1,        /* bad argument (null, negative size, misaligned for the requested mode) */
    ZA_ERR_PLUGIN = -2,     /* unknown plugin_key */
    ZA_ERR_CUDA = -3,       /* CUDA runtime/driver error or no device (no CPU fallback exists) */
    ZA_ERR_STATE = -4,      /* call order violated (process before prepare) */
    ZA_ERR_UNSUPPORTED = -5 /* valid request this build does not implement */
} za_status;

typedef enum za_precision { ZA_F32 = 0, ZA_F64 = 1 } za_precision;   /* arithmetic type; I/O is always float32 */
typedef enum za_memory { ZA_MEM_HOST = 0, ZA_MEM_DEVICE = 1 } za_memory;

typedef struct za_plugin_info {
    int32_t n_in;          /* DSPJSFX_INPUT_CHANNELS  (generated header, dsp_jsfx_aot.py:5924-6190) */
    int32_t n_out;         /* DSPJSFX_OUTPUT_CHANNELS */
    int32_t n_sliders;     /* declared numeric sliders = columns of a parameter block */
    int32_t kind;          /* 0 = DSP-JSFX leaf, 1 = Faust leaf, 2 = authored DSP-JSFX workload */
    int32_t latency;       /* frames of algorithmic latency the plugin reports (0 if none) */
} za_plugin_info;

typedef struct za_slider_info {    /* JsfxSliderDecl, src/JSFXJuceProcessor.cpp:520-545, numeric part */
    int32_t index0;        /* slider number - 1 */
    int32_t is_choice;
    float def, min, max, step;    /* stored as float like the reference (:531-534) */
    char alias[64];        /* `sliderN:alias=...` variable name, "" if none */
} za_slider_info;

/* ---- catalogue ---------------------------------------------------------------------------- */
ZA_API int za_gpu_plugin_count(void);
ZA_API const char* za_gpu_plugin_key(int i);                       /* plugin.json "slug" */
ZA_API int za_gpu_plugin_info(const char* plugin_key, za_plugin_info* out);
ZA_API int za_gpu_slider_info(const char* plugin_key, int column, za_slider_info* out);
/* parseJsfxSliderDecls (src/JSFXJuceProcessor.cpp:710-939) on arbitrary .jsfx text: fills up to
 * `cap` entries sorted by slider index, returns the count (>= 0) or a negative za_status. */
ZA_API int za_gpu_parse_sliders(const char* jsfx_text, za_slider_info* out, int cap);
/* hostParameterToJsfxSliderValue (src/JSFXJuceProcessor.cpp:5556-5597) for one value. */
ZA_API int za_gpu_quantise(const za_slider_info* slider, float host_value, double* out);

/* ---- instance life cycle ------------------------------------------------------------------ */
/* One handle = n_streams independent instances of one plugin on one device, all state in
 * structure-of-arrays HBM buffers owned by the library.  block_len = host block size the
 * reference would be driven with (samplesblock; parameters are sampled once per block). */
ZA_API int za_gpu_create(const char* plugin_key, int n_streams, double sample_rate, int block_len,
                         int precision, int device, za_gpu_handle* out);
/* The same with the reference's oversampling option (x1 / x2 / x4 / x8; src/JSFXJuceProcessor.cpp:3126-3146, 8704-8789): per host
 * block the input is up-sampled linearly inside the block, the plugin runs at sample_rate x oversampling on blocks of
 * block_len x oversampling frames, and the output is box-averaged back.  I/O stays at the host rate. */
ZA_API int za_gpu_create_os(const char* plugin_key, int n_streams, double sample_rate, int block_len, int precision, int device,
                            int oversampling, za_gpu_handle* out);
ZA_API int za_gpu_destroy(za_gpu_handle h);
/* Run on an existing CUDA stream (e.g. the caller's torch stream); NULL = the handle's own. */
ZA_API int za_gpu_set_stream(za_gpu_handle h, void* cuda_stream);

/* prepareToPlay order (src/JSFXJuceProcessor.cpp:3250-3318): reset state (keeping memory),
 * push param_block0 [n_streams][n_sliders] to the sliders, @init, alias re-apply, @slider. */
ZA_API int za_gpu_prepare(za_gpu_handle h, const float* param_block0);

/* processBlock for every stream over n_frames frames, split into host blocks of block_len
 * (last one short).  For block b the parameter row is param_block[min(b, n_param_blocks-1)]
 * ([n_param_blocks][n_streams][n_sliders], host memory); n_param_blocks == 0 keeps the current
 * sliders.  Per block: quantise+push sliders, @slider if changed, @block, then per frame @sample
 * (dsp_jsfx_aot.py:5766-5892).  State carries over between calls (time-chunked rendering).
 * mem = ZA_MEM_DEVICE: in/out are device pointers on the handle's device, work is enqueued on the
 * handle's stream and the call returns without synchronising.  mem = ZA_MEM_HOST: in/out are host
 * buffers (pinned or pageable); copies are pipelined with the kernels and the call returns when
 * `out` is complete. */
ZA_API int za_gpu_process(za_gpu_handle h, const float* param_block, int n_param_blocks,
                          const float* in, float* out, int64_t n_frames, int64_t frame_stride, int mem);

/* Same, additionally writing st.spl[] as double BEFORE the (float) store of
 * dsp_jsfx_aot.py:5866-5892 to tap64 [stream][out_ch][frame] (same stride, same memory kind).
 * Parity instrument for "fp64 mode within 1e-12"; requires precision == ZA_F64. */
ZA_API int za_gpu_process_tap64(za_gpu_handle h, const float* param_block, int n_param_blocks,
                                const float* in, float* out, double* tap64, int64_t n_frames,
                                int64_t frame_stride, int mem);

/* Read one named per-stream state variable (DSPJSFX_VARS name -> vars[] slot in the reference) as
 * double into out[n_streams] (host).  Integer state (counters, ring positions) is exact. */
ZA_API int za_gpu_read_state(za_gpu_handle h, const char* var_name, double* out);
ZA_API int za_gpu_sync(za_gpu_handle h);

/* Checkpoint / resume (the renderer's per-stream SoA state IS its checkpoint).  The blob holds
 * everything a continuing render depends on: current sliders, every per-stream DSP variable, mem[] heaps, rand state,
 * hop phases -- DSPJSFX_State for all streams.  A blob loads only into a handle of the same plugin, stream count,
 * block length, sample rate and precision (any device), so a 10-minute bounce can resume after a fault or a stream
 * range can move to another GPU; rendering after load continues bit-identically to the uninterrupted render. */
ZA_API int za_gpu_state_size(za_gpu_handle h, size_t* bytes);
ZA_API int za_gpu_state_save(za_gpu_handle h, void* host_blob, size_t bytes);
ZA_API int za_gpu_state_load(za_gpu_handle h, const void* host_blob, size_t bytes);

/* Page-locked host buffers for ZA_MEM_HOST calls (the staged copies of za_gpu_process reach PCIe rate only from pinned
 * memory).  Allocated on the calling thread's NUMA node: bind the thread to the GPU's node first on multi-socket boxes. */
ZA_API int za_gpu_host_alloc(size_t bytes, void** out);
ZA_API int za_gpu_host_free(void* p);

/* ---- introspection for benches ------------------------------------------------------------- */
ZA_API int za_gpu_launch_count(za_gpu_handle h, int64_t* kernels_launched);   /* our kernels, cumulative */
/* Device time (ms, CUDA events on the handle's stream) of the dominant kernel launches of the
 * most recent za_gpu_process call, and how many launches that covers. */
ZA_API int za_gpu_last_kernel_ms(za_gpu_handle h, double* ms, int* launches);
/* Tuning knobs for experiments: name = "force_generic_io" | "warps_per_stream" | "host_chunk_frames" |
 * "tile_buffers" (1|2, packed fp32 kernel) | "scalar_f32" (1 = fp32 on the scalar-FFMA kernel) |
 * "heap_slots" (translated leaves: use a smaller mem[] heap than the plugin was sized for; an access beyond it fails the
 * call -- ZA_ERR_STATE, "mem[] index beyond the heap" -- on that call (host memory) or on the next call / za_gpu_sync) |
 * "scalar_f32" also: 1 = the generic stream-per-thread kernel for the Faust leaves that have a tiled one (RED, VAR, ClickBeGoneSG;
 * 2 = the tiled one whatever the batch size, 0 = VAR switches to the generic kernel above ~160 streams per SM),
 * 1 / 2 = the older EasyExpander fp32 kernels | "unstaged" (translated leaves with a stage pipeline -- RTT, SOMA, ATTACK, ADS, SaliencePush: 1 = the
 * one-thread-per-stream kernel, 2 = the pipeline whatever the batch size, 0 = picked by batch size; same bits in float64) | "lanes_per_warp" (8 | 16: streams per warp of the stream-per-thread
 * kernels) | "role_off" (EasyExpander v3 role ablation; measurement only: outputs are not valid). */
ZA_API int za_gpu_set_option(za_gpu_handle h, const char* name, int value);

/* ---- one-shot form sketched in BASELINE.json north_star ------------------------------------ */
/* za_gpu_process(plugin_key, param_block, in, out, n_streams, n_frames): create + prepare +
 * process (host buffers, static parameters, 48 kHz, block 512, fp64) + destroy. */
ZA_API int za_gpu_render(const char* plugin_key, const float* param_block, const float* in, float* out,
                         int n_streams, int64_t n_frames);

/* ---- serial plugin chains on the device -------------------------------- */
/* n_plugins stages rendered one after the other for every stream, stage k's float32 output being stage k + 1's input
 * without leaving the GPU -- what a host does when it hands one AudioBuffer<float> from plugin to plugin
 * (src/JSFXJuceProcessor.cpp:3456-3528).  Routing: input channel c of stage k > 0 = output channel c of stage k - 1 where
 * that exists, else channel c of the CHAIN input (side-chain keys, e.g. RTT's channels 4-7,
 * plugins/Dynamics/RTT/src/Reverb Tail Tamer.jsfx:3-6), else silence.  The chain input has max(n_in of all stages)
 * channels, the chain output the last stage's n_out.  The call is cut into time chunks of `chunk_blocks` host blocks
 * (default 4) so the intermediates stay small; state carries across chunks and calls in every stage.
 * params[k]: stage k's automation rows [n_param_blocks[k]][n_streams][n_sliders of stage k] (host floats) or NULL. */
typedef struct za_chain* za_chain_handle;
ZA_API int za_chain_create(const char* const* plugin_keys, int n_plugins, int n_streams, double sample_rate, int block_len,
                           int precision, int device, za_chain_handle* out);
ZA_API int za_chain_destroy(za_chain_handle c);
ZA_API int za_chain_info(za_chain_handle c, int* n_in, int* n_out, int* n_stages);
ZA_API int za_chain_stage(za_chain_handle c, int k, za_gpu_handle* stage);    /* the stage's handle (read_state, options; owned by the chain) */
ZA_API int za_chain_set_chunk_blocks(za_chain_handle c, int blocks);
ZA_API int za_chain_prepare(za_chain_handle c, const float* const* param_block0);
ZA_API int za_chain_process(za_chain_handle c, const float* const* params, const int* n_param_blocks, const float* in, float* out,
                            int64_t n_frames, int64_t frame_stride, int mem);
ZA_API int za_chain_sync(za_chain_handle c);

/* ---- one call, several GPUs --------------------------------------------------- */
/* Streams shard as contiguous ranges over `devices` (stream s -> shard s * n_devices / n_streams), no collective on the data
 * path.  za_batch_process takes HOST buffers for the whole batch and renders every shard from its own host thread (bound to
 * its GPU's NUMA node) through that device's own copy / kernel pipeline.  za_batch_shard gives the per-device handle
 * (checkpoints, read_state, options). */
typedef struct za_batch* za_batch_handle;
ZA_API int za_batch_create(const char* plugin_key, int n_streams, double sample_rate, int block_len, int precision,
                           const int* devices, int n_devices, int oversampling, za_batch_handle* out);
ZA_API int za_batch_destroy(za_batch_handle b);
ZA_API int za_batch_shards(za_batch_handle b, int* n_shards, int* first_stream, int* n_streams, int* device, int cap);
ZA_API int za_batch_shard(za_batch_handle b, int shard, za_gpu_handle* out);
ZA_API int za_batch_prepare(za_batch_handle b, const float* param_block0);
ZA_API int za_batch_process(za_batch_handle b, const float* param_block, int n_param_blocks, const float* in, float* out,
                            int64_t n_frames, int64_t frame_stride);

/* ---- synthetic programme material + checksums ------------------- */
/* Bench / sampled-parity tooling beside the render path: the inputs of a large bounce are generated ON THE DEVICE from a
 * counter-based hash of (seed, first_stream + s, channel, t0 + frame), so no PCIe or host storage is on the path, and
 * any stream can be regenerated on the CPU bit for bit (za_gpu/workloads.py::synth; integer hashing, two host-made
 * tables, individually rounded float operations).  kind 0 = 0.25 x uniform noise + two sines (the cfg2 input),
 * kind 1 = noise bursts under a -60 .. -6 dBFS piecewise-linear envelope (the cfg3 "programme-like" input).
 * dev_out: device rows [stream][channel][frame_stride]; enqueued on cuda_stream (NULL = default), not synchronised. */
ZA_API int za_gpu_synth(int kind, uint32_t seed, int first_stream, int n_streams, int n_ch, int64_t n_frames,
                        int64_t frame_stride, int64_t t0, float* dev_out, int device, void* cuda_stream);
ZA_API int za_gpu_synth_tables(float* sine4096, float* amp256);      /* the generator's two host tables (no GPU needed) */
/* Per-row reduction of device rows: host_sum[r] = sum_i (bits(x[r][i]) + 1) * (mix32(i) | 1) mod 2^64 (order
 * independent, bit-exact) and host_energy[r] = sum_i x^2 (double).  Synchronises cuda_stream. */
ZA_API int za_gpu_checksum(const float* dev, int n_rows, int64_t n_frames, int64_t frame_stride, uint64_t* host_sum,
                           double* host_energy, int device, void* cuda_stream);
/* The same per-row reduction for callers that may not allocate device memory (CUDA graph capture around them, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Its accumulators live in a file-scope static device array of ZA_CHECKSUM_NOALLOC_MAX_ROWS rows (256 KB that
 * every GPU loading the library holds until the process exits); the call makes no cudaMalloc / cudaFree.  n_rows beyond
 * the maximum is refused with ZA_ERR_ARG and a message (split the rows over calls, or use za_gpu_checksum).  The array is
 * shared by every call on a device and keeps the previous call's contents: each call clears the cells it uses on
 * cuda_stream before its kernel, and calls that overlap on the GPU (different streams, different host threads) race --
 * serialise them.  Results are bit-identical to za_gpu_checksum's sums; energies agree to double rounding of the
 * atomics' order.  Synchronises cuda_stream. */
#define ZA_CHECKSUM_NOALLOC_MAX_ROWS 16384
ZA_API int za_gpu_checksum_noalloc(const float* dev, int n_rows, int64_t n_frames, int64_t frame_stride, uint64_t* host_sum,
                                   double* host_energy, int device, void* cuda_stream);

ZA_API const char* za_gpu_last_error(void);
ZA_API int za_gpu_device_count(void);
ZA_API const char* za_gpu_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ZA_GPU_H */
