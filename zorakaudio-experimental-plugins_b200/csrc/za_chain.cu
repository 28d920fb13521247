// za_chain.cu -- serial plugin chains rendered on the device: stage k's float32 output buffer is
// stage k + 1's input without leaving the GPU, the way a host passes one AudioBuffer<float> from plugin to plugin
// (JUCE bus buffers, src/JSFXJuceProcessor.cpp:3456-3528), including side-chain routing as in RTT's 8-channel contract
// (plugins/Dynamics/RTT/src/Reverb Tail Tamer.jsfx:3-6,229-236: wet 0/1, keys on channels 4-7).
//
// A chain is a list of ordinary handles (za_gpu_create) on one device sharing one CUDA stream.  The call is cut into TIME
// CHUNKS of whole host blocks; per chunk every stage renders all streams into a ping-pong intermediate, so the
// intermediates are small (chunk_blocks x block_len frames per row; with the default 4 blocks a 4096-stream stereo chain
// keeps them inside the 126 MB L2) and no stage ever waits for a host round trip.  State carries from chunk to chunk in
// each stage's own handle, exactly as in a time-chunked za_gpu_process render.
// Routing rule: input channel c of stage k > 0 is output channel c of stage k - 1 when that stage has such a channel,
// else channel c of the CHAIN input (side-chain keys ride along), else silence; a gather kernel builds the input rows
// only where the channel counts differ.
#include <cstring>
#include <string>
#include <vector>

#include "za_host.h"

using namespace za;

struct za_chain {
    std::vector<za_gpu_handle> stages;
    std::vector<za_plugin_info> info;
    int n_streams = 0, block_len = 512, device = 0, chunk_blocks = 4, n_in = 0, n_out = 0;
    cudaStream_t stream = nullptr;
    float* buf[2] = {nullptr, nullptr};     // ping-pong intermediates, [n_streams][max channels][chunk stride]
    float* route = nullptr;                 // gathered input rows of a stage whose channel count differs from its source
    float* h_in = nullptr; float* h_out = nullptr;      // device staging for ZA_MEM_HOST calls
    int64_t stride = 0;                     // frames per row of the intermediates
    int max_ch = 0;
    int64_t launches = 0;
};

namespace {

// dst[s][c][f] = c < n_prev ? prev[s][c][f] : (c < n_chain ? chain_in[s][c][f0 + f] : 0)
__global__ void __launch_bounds__(256) chain_route_kernel(float* __restrict__ dst, int n_dst, int64_t dst_stride, const float* __restrict__ prev, int n_prev,
                                                          int64_t prev_stride, const float* __restrict__ chain_in, int n_chain, int64_t chain_stride,
                                                          int64_t n_frames)
{
    const int row = blockIdx.y, s = row / n_dst, c = row - s * n_dst;
    float* d = dst + (size_t)row * dst_stride;
    const float* src = c < n_prev ? prev + ((size_t)s * n_prev + c) * prev_stride : (c < n_chain ? chain_in + ((size_t)s * n_chain + c) * chain_stride : nullptr);
    for (int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; f < n_frames; f += (int64_t)gridDim.x * blockDim.x) d[f] = src ? src[f] : 0.0f;
}

void chain_free(za_chain* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    for (auto h : c->stages) za_gpu_destroy(h);
    cudaFree(c->buf[0]); cudaFree(c->buf[1]); cudaFree(c->route); cudaFree(c->h_in); cudaFree(c->h_out);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

}  // namespace

extern "C" {

ZA_API int za_chain_create(const char* const* plugin_keys, int n_plugins, int n_streams, double sample_rate, int block_len, int precision,
                           int device, za_chain_handle* out)
{
    if (!out) return ZA_ERR_ARG;
    *out = nullptr;
    if (!plugin_keys || n_plugins <= 0 || n_plugins > 16 || n_streams <= 0 || block_len <= 0) { set_error("za_chain_create: bad argument"); return ZA_ERR_ARG; }
    za_chain* c = new za_chain();
    c->n_streams = n_streams; c->block_len = block_len; c->device = device;
    for (int k = 0; k < n_plugins; ++k) {
        za_plugin_info pi;
        int rc = za_gpu_plugin_info(plugin_keys[k], &pi);
        if (rc != ZA_OK) { chain_free(c); return rc; }
        za_gpu_handle h = nullptr;
        // Faust leaves compute in float32 whatever the chain's precision (the reference builds them without -double)
        rc = za_gpu_create(plugin_keys[k], n_streams, sample_rate, block_len, pi.kind == 1 ? ZA_F32 : precision, device, &h);
        if (rc != ZA_OK) { chain_free(c); return rc; }
        c->stages.push_back(h); c->info.push_back(pi);
        c->max_ch = std::max(c->max_ch, std::max(pi.n_in, pi.n_out));
    }
    c->n_in = c->info.front().n_in;
    for (const auto& pi : c->info) c->n_in = std::max(c->n_in, pi.n_in);       // side-chain channels of later stages come from the chain input
    c->n_out = c->info.back().n_out;
    if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice") || !cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "chain stream")) { chain_free(c); return ZA_ERR_CUDA; }
    for (auto h : c->stages) za_gpu_set_stream(h, c->stream);
    *out = c;
    return ZA_OK;
}

ZA_API int za_chain_destroy(za_chain_handle c) { chain_free(c); return ZA_OK; }

ZA_API int za_chain_info(za_chain_handle c, int* n_in, int* n_out, int* n_stages)
{
    if (!c) return ZA_ERR_ARG;
    if (n_in) *n_in = c->n_in;
    if (n_out) *n_out = c->n_out;
    if (n_stages) *n_stages = (int)c->stages.size();
    return ZA_OK;
}

ZA_API int za_chain_stage(za_chain_handle c, int k, za_gpu_handle* out)
{
    if (!c || !out || k < 0 || k >= (int)c->stages.size()) return ZA_ERR_ARG;
    *out = c->stages[k];
    return ZA_OK;
}

ZA_API int za_chain_set_chunk_blocks(za_chain_handle c, int blocks)
{
    if (!c || blocks <= 0) return ZA_ERR_ARG;
    c->chunk_blocks = blocks;
    return ZA_OK;
}

ZA_API int za_chain_prepare(za_chain_handle c, const float* const* param_block0)
{
    if (!c) return ZA_ERR_ARG;
    for (size_t k = 0; k < c->stages.size(); ++k) {
        const int rc = za_gpu_prepare(c->stages[k], param_block0 ? param_block0[k] : nullptr);
        if (rc != ZA_OK) return rc;
    }
    return ZA_OK;
}

ZA_API int za_chain_process(za_chain_handle c, const float* const* params, const int* n_param_blocks, const float* in, float* out,
                            int64_t n_frames, int64_t frame_stride, int mem)
{
    if (!c || !in || !out || n_frames < 0 || frame_stride < n_frames || (mem != ZA_MEM_HOST && mem != ZA_MEM_DEVICE)) { set_error("za_chain_process: bad argument"); return ZA_ERR_ARG; }
    if (!cuda_ok(cudaSetDevice(c->device), "cudaSetDevice")) return ZA_ERR_CUDA;
    const int64_t bl = c->block_len, cf = bl * c->chunk_blocks;
    const int64_t need = (cf + 31) / 32 * 32;
    const size_t rows = (size_t)c->n_streams * c->max_ch;
    if (need != c->stride) {
        cudaFree(c->buf[0]); cudaFree(c->buf[1]); cudaFree(c->route); cudaFree(c->h_in); cudaFree(c->h_out);
        c->buf[0] = c->buf[1] = c->route = c->h_in = c->h_out = nullptr; c->stride = 0;
        if (!cuda_ok(cudaMalloc(&c->buf[0], rows * need * 4), "chain intermediate") || !cuda_ok(cudaMalloc(&c->buf[1], rows * need * 4), "chain intermediate") ||
            !cuda_ok(cudaMalloc(&c->route, rows * need * 4), "chain routing buffer")) return ZA_ERR_CUDA;
        c->stride = need;
    }
    if (mem == ZA_MEM_HOST && !c->h_in) {
        if (!cuda_ok(cudaMalloc(&c->h_in, (size_t)c->n_streams * c->n_in * need * 4), "chain staging") ||
            !cuda_ok(cudaMalloc(&c->h_out, (size_t)c->n_streams * c->n_out * need * 4), "chain staging")) return ZA_ERR_CUDA;
    }
    const int ns = c->n_streams, n_st = (int)c->stages.size();
    for (int64_t f0 = 0; f0 < n_frames; f0 += cf) {
        const int64_t nf = std::min<int64_t>(cf, n_frames - f0);
        const int64_t b0 = f0 / bl;
        const float* cin; int64_t cin_stride;
        if (mem == ZA_MEM_HOST) {
            if (!cuda_ok(cudaMemcpy2DAsync(c->h_in, need * 4, in + f0, frame_stride * 4, nf * 4, (size_t)ns * c->n_in, cudaMemcpyHostToDevice, c->stream), "chain H2D")) return ZA_ERR_CUDA;
            cin = c->h_in; cin_stride = need;
        } else { cin = in + f0; cin_stride = frame_stride; }
        const float* prev = nullptr; int n_prev = 0; int64_t prev_stride = 0;
        for (int k = 0; k < n_st; ++k) {
            const za_plugin_info& pi = c->info[k];
            const float* src; int64_t src_stride;
            if (k == 0 && pi.n_in == c->n_in) { src = cin; src_stride = cin_stride; }
            else if (k > 0 && pi.n_in == n_prev) { src = prev; src_stride = prev_stride; }
            else {      // gather this stage's input rows from the previous output and the chain input's side-chain channels
                const dim3 grid((unsigned)std::min<int64_t>((nf + 255) / 256, 16), (unsigned)(ns * pi.n_in));
                chain_route_kernel<<<grid, 256, 0, c->stream>>>(c->route, pi.n_in, need, k == 0 ? cin : prev, k == 0 ? c->n_in : n_prev,
                                                                k == 0 ? cin_stride : prev_stride, cin, c->n_in, cin_stride, nf);
                if (!cuda_ok(cudaGetLastError(), "chain_route_kernel")) return ZA_ERR_CUDA;
                ++c->launches;
                src = c->route; src_stride = need;
            }
            float* dst; int64_t dst_stride;
            if (k == n_st - 1) {
                if (mem == ZA_MEM_HOST) { dst = c->h_out; dst_stride = need; } else { dst = out + f0; dst_stride = frame_stride; }
            } else { dst = c->buf[k & 1]; dst_stride = need; }
            // a stage renders with ONE row stride for input and output: rows of the narrower side use the same stride
            if (src_stride != dst_stride) {
                if (mem == ZA_MEM_DEVICE && (k == 0 || k == n_st - 1) && n_st > 1) {
                    // caller-stride buffer on one side, chunk-stride intermediate on the other: bring the caller side to the chunk stride
                    if (k == 0) {
                        const dim3 grid((unsigned)std::min<int64_t>((nf + 255) / 256, 16), (unsigned)(ns * pi.n_in));
                        chain_route_kernel<<<grid, 256, 0, c->stream>>>(c->route, pi.n_in, need, src, pi.n_in, src_stride, cin, c->n_in, cin_stride, nf);
                        src = c->route; src_stride = need; ++c->launches;
                    } else { dst = c->buf[k & 1]; dst_stride = need; }
                }
            }
            const float* p = nullptr; int np = 0;
            if (params && params[k] && n_param_blocks && n_param_blocks[k] > 0) {
                const int64_t row0 = std::min<int64_t>(b0, n_param_blocks[k] - 1);
                p = params[k] + (size_t)row0 * ns * pi.n_sliders; np = (int)(n_param_blocks[k] - row0);
            }
            const int rc = za_gpu_process(c->stages[k], p, np, src, dst, nf, dst_stride, ZA_MEM_DEVICE);
            if (rc != ZA_OK) return rc;
            if (k == n_st - 1 && mem == ZA_MEM_DEVICE && dst != out + f0) {      // last stage rendered at the chunk stride: copy rows out
                if (!cuda_ok(cudaMemcpy2DAsync(out + f0, frame_stride * 4, dst, dst_stride * 4, nf * 4, (size_t)ns * pi.n_out, cudaMemcpyDeviceToDevice, c->stream), "chain D2D")) return ZA_ERR_CUDA;
            }
            prev = dst; n_prev = pi.n_out; prev_stride = dst_stride;
        }
        if (mem == ZA_MEM_HOST) {
            if (!cuda_ok(cudaMemcpy2DAsync(out + f0, frame_stride * 4, c->h_out, need * 4, nf * 4, (size_t)ns * c->n_out, cudaMemcpyDeviceToHost, c->stream), "chain D2H")) return ZA_ERR_CUDA;
            if (!cuda_ok(cudaStreamSynchronize(c->stream), "chain sync")) return ZA_ERR_CUDA;   // h_in / h_out are reused by the next chunk
        }
    }
    if (mem == ZA_MEM_HOST && !cuda_ok(cudaStreamSynchronize(c->stream), "chain sync")) return ZA_ERR_CUDA;
    return ZA_OK;
}

ZA_API int za_chain_sync(za_chain_handle c)
{
    if (!c) return ZA_ERR_ARG;
    cudaSetDevice(c->device);
    return cuda_ok(cudaStreamSynchronize(c->stream), "chain sync") ? ZA_OK : ZA_ERR_CUDA;
}

}  // extern "C"
