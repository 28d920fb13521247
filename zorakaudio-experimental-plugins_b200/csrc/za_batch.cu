// za_batch.cu -- one call, several GPUs: streams are independent, so a batch shards as contiguous
// stream ranges over the devices of one box (stream s -> device s * G / n_streams) with NO collective on the data path.
// A za_batch handle owns one ordinary handle (za_gpu_create_os) per device; za_batch_process runs one HOST THREAD PER
// DEVICE, each rendering its shard's rows of the caller's host buffers through za_gpu_process(..., ZA_MEM_HOST) -- its own
// H2D / kernels / D2H pipeline on its own streams -- so the PCIe links of all devices work at once from one process.
// Each device thread first binds itself to the CPUs of its GPU's NUMA node (sysfs numa_node of the PCI function), which
// keeps the staging traffic of the pageable / pinned host rows it touches on the socket the GPU hangs off.
#include <sched.h>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

#include "za_host.h"

using namespace za;

struct za_batch {
    std::vector<za_gpu_handle> shard;
    std::vector<int> device, first, count;          // per shard: device index, first stream, streams
    int n_streams = 0, n_sliders = 0, n_in = 0, n_out = 0;
    std::vector<std::string> errors;
};

namespace {

void bind_to_gpu_node(int device)
{
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) return;
    for (char* p = bus; *p; ++p) *p = (char)std::tolower((unsigned char)*p);
    std::ifstream nf(std::string("/sys/bus/pci/devices/") + bus + "/numa_node");
    int node = -1;
    if (!(nf >> node) || node < 0) return;
    std::ifstream cf("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist");
    std::string list;
    if (!std::getline(cf, list)) return;
    cpu_set_t want, have;
    CPU_ZERO(&want);
    if (sched_getaffinity(0, sizeof have, &have) != 0) return;
    size_t i = 0;
    int any = 0;
    while (i < list.size()) {
        const int a = std::atoi(list.c_str() + i);
        size_t j = list.find_first_of(",-", i);
        int b = a;
        if (j != std::string::npos && list[j] == '-') { b = std::atoi(list.c_str() + j + 1); j = list.find(',', j); }
        for (int c = a; c <= b && c < CPU_SETSIZE; ++c) if (CPU_ISSET(c, &have)) { CPU_SET(c, &want); any = 1; }
        if (j == std::string::npos) break;
        i = j + 1;
    }
    if (any) sched_setaffinity(0, sizeof want, &want);
}

}  // namespace

extern "C" {

ZA_API int za_batch_create(const char* plugin_key, int n_streams, double sample_rate, int block_len, int precision, const int* devices,
                           int n_devices, int oversampling, za_batch_handle* out)
{
    if (!out) return ZA_ERR_ARG;
    *out = nullptr;
    if (!devices || n_devices <= 0 || n_devices > 64 || n_streams < n_devices) { set_error("za_batch_create: bad argument (at least one stream per device)"); return ZA_ERR_ARG; }
    za_plugin_info pi;
    int rc = za_gpu_plugin_info(plugin_key, &pi);
    if (rc != ZA_OK) return rc;
    za_batch* b = new za_batch();
    b->n_streams = n_streams; b->n_sliders = pi.n_sliders; b->n_in = pi.n_in; b->n_out = pi.n_out;
    for (int g = 0; g < n_devices; ++g) {
        const int s0 = (int)((int64_t)n_streams * g / n_devices), s1 = (int)((int64_t)n_streams * (g + 1) / n_devices);
        za_gpu_handle h = nullptr;
        rc = za_gpu_create_os(plugin_key, s1 - s0, sample_rate, block_len, precision, devices[g], oversampling, &h);
        if (rc != ZA_OK) { for (auto x : b->shard) za_gpu_destroy(x); delete b; return rc; }
        b->shard.push_back(h); b->device.push_back(devices[g]); b->first.push_back(s0); b->count.push_back(s1 - s0);
    }
    b->errors.resize(b->shard.size());
    *out = b;
    return ZA_OK;
}

ZA_API int za_batch_destroy(za_batch_handle b)
{
    if (!b) return ZA_OK;
    for (auto h : b->shard) za_gpu_destroy(h);
    delete b;
    return ZA_OK;
}

ZA_API int za_batch_shards(za_batch_handle b, int* n_shards, int* first_stream, int* n_streams, int* device, int cap)
{
    if (!b || !n_shards) return ZA_ERR_ARG;
    *n_shards = (int)b->shard.size();
    for (int g = 0; g < (int)b->shard.size() && g < cap; ++g) {
        if (first_stream) first_stream[g] = b->first[g];
        if (n_streams) n_streams[g] = b->count[g];
        if (device) device[g] = b->device[g];
    }
    return ZA_OK;
}

ZA_API int za_batch_shard(za_batch_handle b, int g, za_gpu_handle* out)
{
    if (!b || !out || g < 0 || g >= (int)b->shard.size()) return ZA_ERR_ARG;
    *out = b->shard[g];
    return ZA_OK;
}

ZA_API int za_batch_prepare(za_batch_handle b, const float* param_block0)
{
    if (!b) return ZA_ERR_ARG;
    for (size_t g = 0; g < b->shard.size(); ++g) {
        const int rc = za_gpu_prepare(b->shard[g], param_block0 ? param_block0 + (size_t)b->first[g] * b->n_sliders : nullptr);
        if (rc != ZA_OK) return rc;
    }
    return ZA_OK;
}

// HOST buffers [n_streams][channels][frame_stride]; params [n_param_blocks][n_streams][n_sliders] or null.
ZA_API int za_batch_process(za_batch_handle b, const float* param_block, int n_param_blocks, const float* in, float* out, int64_t n_frames,
                            int64_t frame_stride)
{
    if (!b || !out || n_frames < 0 || frame_stride < n_frames || n_param_blocks < 0 || (n_param_blocks > 0 && !param_block)) {
        set_error("za_batch_process: bad argument");
        return ZA_ERR_ARG;
    }
    const int G = (int)b->shard.size(), nd = b->n_sliders;
    std::vector<int> rcs((size_t)G, ZA_OK);
    std::vector<std::thread> th;
    for (int g = 0; g < G; ++g)
        th.emplace_back([&, g] {
            bind_to_gpu_node(b->device[g]);
            const int s0 = b->first[g], ns = b->count[g];
            std::vector<float> p;                                   // this shard's rows, repacked [block][its streams][sliders]
            if (n_param_blocks > 0) {
                p.resize((size_t)n_param_blocks * ns * nd);
                for (int blk = 0; blk < n_param_blocks; ++blk)
                    std::memcpy(&p[(size_t)blk * ns * nd], param_block + ((size_t)blk * b->n_streams + s0) * nd, (size_t)ns * nd * sizeof(float));
            }
            rcs[(size_t)g] = za_gpu_process(b->shard[g], n_param_blocks > 0 ? p.data() : nullptr, n_param_blocks,
                                            in ? in + (size_t)s0 * b->n_in * frame_stride : nullptr, out + (size_t)s0 * b->n_out * frame_stride,
                                            n_frames, frame_stride, ZA_MEM_HOST);
            if (rcs[(size_t)g] != ZA_OK) b->errors[(size_t)g] = za_gpu_last_error();
        });
    for (auto& t : th) t.join();
    for (int g = 0; g < G; ++g)
        if (rcs[(size_t)g] != ZA_OK) { set_error("device " + std::to_string(b->device[g]) + ": " + b->errors[(size_t)g]); return rcs[(size_t)g]; }
    return ZA_OK;
}

}  // extern "C"
