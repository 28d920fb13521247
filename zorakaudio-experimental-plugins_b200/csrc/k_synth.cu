// k_synth.cu -- counter-based synthetic programme material generated ON THE DEVICE, and per-row checksums.
//
// "inputs generated on device from per-stream counters-based PRNG (seed = stream id) ...
// outputs reduced to per-stream checksums + energy ... a random subset of streams re-rendered by the CPU oracle".
// Every sample is a pure function of (seed, stream, channel, absolute frame): integer hashing, exact int -> float
// conversions, table look-ups and individually rounded float operations only, so za_gpu/workloads.py::synth
// reproduces any stream on the CPU bit for bit (no libm on the device side; the two tables come from the host).
// This is bench / test tooling beside the render path, not part of it; it is exported through the C ABI so that the
// same generator serves bench.py, bench_cfg5.py and the bounce driver's --synth mode.
#include <cmath>
#include <mutex>

#include "za_host.h"

namespace za {

namespace {

__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__host__ __device__ __forceinline__ uint32_t stream_key(uint32_t seed, uint32_t stream, uint32_t ch)
{
    return mix32(seed * 0x9E3779B1u + stream * 0x85EBCA77u + ch * 0xC2B2AE3Du + 0x27D4EB2Fu);
}
__device__ __forceinline__ float uni(uint32_t h) { return __fadd_rn(__fmul_rn((float)(h >> 8), 1.1920928955078125e-07f), -1.0f); }   // [-1, 1), exact
__device__ __forceinline__ uint32_t frame_hash(uint32_t key, int64_t i)
{
    return mix32(((uint32_t)i * 0x9E3779B1u + (uint32_t)((uint64_t)i >> 32) * 0x632BE5ABu) ^ key);
}

constexpr int kSine = 4096, kAmp = 256, kKnot = 2400;

struct SynthTables { float sine[kSine]; float amp[kAmp]; };

const SynthTables& host_tables()
{
    static SynthTables t;
    static std::once_flag once;
    std::call_once(once, [] {
        for (int k = 0; k < kSine; ++k) t.sine[k] = (float)std::sin(6.283185307179586476925 * k / kSine);
        for (int j = 0; j < kAmp; ++j) t.amp[j] = (float)(0.7 * std::pow(10.0, (-60.0 + 54.0 * j / (kAmp - 1)) / 20.0));
    });
    return t;
}

// kind 0: 0.25 x uniform noise + two table sines at -18 dBFS each (the cfg2 input of SURVEY 8d; f1 60..400 Hz, f2 1..6 kHz
//         per stream from the key); kind 1: noise (common + 0.3 x per-channel) under an envelope that is piecewise linear
//         between 50 ms knots drawn from -60 .. -6 dBFS (cfg3 "programme-like bursts").
__global__ void __launch_bounds__(256) synth_kernel(float* __restrict__ out, int kind, uint32_t seed, int first_stream, int n_ch,
                                                   int64_t n_frames, int64_t stride, int64_t t0, const SynthTables* __restrict__ tab)
{
    const int row = blockIdx.y;                              // stream * n_ch + ch
    const int s = row / n_ch, c = row - s * n_ch;
    const uint32_t key = stream_key(seed, (uint32_t)(first_stream + s), (uint32_t)c);
    const uint32_t key_m = stream_key(seed, (uint32_t)(first_stream + s), 255u);
    float* o = out + (size_t)row * stride;
    uint32_t inc1 = 0, inc2 = 0;
    if (kind == 0) {
        inc1 = 5368709u + mix32(key_m ^ 0x1111u) % 30422000u;        // 60 .. 400 Hz as 2^32 / 48000 per Hz
        inc2 = 89478485u + mix32(key_m ^ 0x2222u) % 447392000u;      // 1 .. 6 kHz
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_frames; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = t0 + i;
        float v;
        if (kind == 0) {
            const float u = uni(frame_hash(key, t));
            const float s1 = tab->sine[((uint32_t)t * inc1) >> 20], s2 = tab->sine[((uint32_t)t * inc2) >> 20];
            v = __fadd_rn(__fadd_rn(__fmul_rn(0.25f, u), __fmul_rn(0.125f, s1)), __fmul_rn(0.125f, s2));
        } else {
            const int64_t k = t / kKnot;
            const int r = (int)(t - k * kKnot);
            const float a0 = tab->amp[mix32(key_m ^ 0xA5A5A5A5u ^ ((uint32_t)k * 0x9E3779B1u)) & (kAmp - 1)];
            const float a1 = tab->amp[mix32(key_m ^ 0xA5A5A5A5u ^ ((uint32_t)(k + 1) * 0x9E3779B1u)) & (kAmp - 1)];
            const float e = __fadd_rn(a0, __fmul_rn(__fsub_rn(a1, a0), __fmul_rn((float)r, 4.1666667675599456e-04f)));
            const float mono = uni(frame_hash(key_m, t)), side = uni(frame_hash(key, t));
            v = __fmul_rn(e, __fadd_rn(mono, __fmul_rn(0.3f, side)));
        }
        o[i] = v;
    }
}

// Order-independent checksum of a row: sum over frames of (bits(x[i]) + 1) * (mix32(i) | 1) in wrapping 64-bit arithmetic
// (commutative, so the atomics give the same value on every run), and the energy sum of squares in double.
__global__ void __launch_bounds__(256) checksum_kernel(const float* __restrict__ x, int64_t n_frames, int64_t stride,
                                                      unsigned long long* __restrict__ sum, double* __restrict__ energy)
{
    const int row = blockIdx.y;
    const float* p = x + (size_t)row * stride;
    unsigned long long a = 0;
    double e = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_frames; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = p[i];
        a += ((unsigned long long)__float_as_uint(v) + 1ull) * (unsigned long long)(mix32((uint32_t)i) | 1u);
        e += (double)v * (double)v;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, d);
        e += __shfl_xor_sync(0xffffffffu, e, d);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(sum + row, a); atomicAdd(energy + row, e); }
}

struct DevTables { SynthTables* d[64] = {}; std::mutex m; };
DevTables g_tabs;

const SynthTables* device_tables(int device)
{
    std::lock_guard<std::mutex> lk(g_tabs.m);
    if (device < 0 || device >= 64) return nullptr;
    if (!g_tabs.d[device]) {
        SynthTables* p = nullptr;
        if (!cuda_ok(cudaMalloc(&p, sizeof(SynthTables)), "synth tables") ||
            !cuda_ok(cudaMemcpy(p, &host_tables(), sizeof(SynthTables), cudaMemcpyHostToDevice), "synth tables upload")) return nullptr;
        g_tabs.d[device] = p;
    }
    return g_tabs.d[device];
}

// ---- scratch of za_gpu_checksum_noalloc: a file-scope device array, so that entry point never allocates -----------------
// Every GPU that loads the library holds these 256 KB until the process exits.  Layout: [0, kNoAllocMaxRows) the wrapping
// 64-bit row sums, [kNoAllocMaxRows, 2 kNoAllocMaxRows) the row energies (doubles, same 8-byte cells).
constexpr int kNoAllocMaxRows = 16384;
static __device__ __align__(256) unsigned long long g_checksum_scratch[2 * kNoAllocMaxRows];

// In host code `g_checksum_scratch` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
struct ScratchAddr { unsigned long long* d[64] = {}; std::mutex m; };
ScratchAddr g_scratch_addr;

unsigned long long* checksum_scratch(int device)       // the caller has made `device` current
{
    if (device < 0 || device >= 64) { set_error("za_gpu_checksum_noalloc: device index out of range"); return nullptr; }
    std::lock_guard<std::mutex> lk(g_scratch_addr.m);
    if (!g_scratch_addr.d[device]) {
        void* p = nullptr;
        if (!cuda_ok(cudaGetSymbolAddress(&p, g_checksum_scratch), "cudaGetSymbolAddress(checksum scratch)") || !p) return nullptr;
        g_scratch_addr.d[device] = (unsigned long long*)p;
    }
    return g_scratch_addr.d[device];
}

}  // namespace
}  // namespace za

using namespace za;

extern "C" {

ZA_API int za_gpu_synth_tables(float* sine4096, float* amp256)
{
    if (!sine4096 || !amp256) return ZA_ERR_ARG;
    const SynthTables& t = host_tables();
    for (int k = 0; k < kSine; ++k) sine4096[k] = t.sine[k];
    for (int j = 0; j < kAmp; ++j) amp256[j] = t.amp[j];
    return ZA_OK;
}

ZA_API int za_gpu_synth(int kind, uint32_t seed, int first_stream, int n_streams, int n_ch, int64_t n_frames, int64_t frame_stride,
                        int64_t t0, float* dev_out, int device, void* cuda_stream)
{
    if ((kind != 0 && kind != 1) || n_streams <= 0 || n_ch <= 0 || n_ch > 64 || n_frames < 0 || frame_stride < n_frames || !dev_out ||
        (int64_t)n_streams * n_ch > 65535) {
        set_error("za_gpu_synth: bad argument");
        return ZA_ERR_ARG;
    }
    if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice")) return ZA_ERR_CUDA;
    const SynthTables* tab = device_tables(device);
    if (!tab) return ZA_ERR_CUDA;
    if (n_frames == 0) return ZA_OK;
    const int bx = (int)std::min<int64_t>((n_frames + 255) / 256, 64);
    synth_kernel<<<dim3(bx, n_streams * n_ch), 256, 0, (cudaStream_t)cuda_stream>>>(dev_out, kind, seed, first_stream, n_ch, n_frames,
                                                                                   frame_stride, t0, tab);
    return cuda_ok(cudaGetLastError(), "synth_kernel") ? ZA_OK : ZA_ERR_CUDA;
}

ZA_API int za_gpu_checksum(const float* dev, int n_rows, int64_t n_frames, int64_t frame_stride, uint64_t* host_sum, double* host_energy,
                           int device, void* cuda_stream)
{
    if (!dev || n_rows <= 0 || n_rows > 65535 || n_frames < 0 || frame_stride < n_frames || !host_sum || !host_energy) {
        set_error("za_gpu_checksum: bad argument");
        return ZA_ERR_ARG;
    }
    if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice")) return ZA_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    unsigned long long* d_sum = nullptr;
    double* d_e = nullptr;
    if (!cuda_ok(cudaMalloc(&d_sum, (size_t)n_rows * 8), "checksum buffer") || !cuda_ok(cudaMalloc(&d_e, (size_t)n_rows * 8), "energy buffer")) {
        cudaFree(d_sum);
        return ZA_ERR_CUDA;
    }
    cudaMemsetAsync(d_sum, 0, (size_t)n_rows * 8, st);
    cudaMemsetAsync(d_e, 0, (size_t)n_rows * 8, st);
    if (n_frames > 0) {
        const int bx = (int)std::min<int64_t>((n_frames + 255) / 256, 64);
        checksum_kernel<<<dim3(bx, n_rows), 256, 0, st>>>(dev, n_frames, frame_stride, d_sum, d_e);
    }
    bool ok = cuda_ok(cudaGetLastError(), "checksum_kernel") &&
              cuda_ok(cudaMemcpyAsync(host_sum, d_sum, (size_t)n_rows * 8, cudaMemcpyDeviceToHost, st), "checksum read") &&
              cuda_ok(cudaMemcpyAsync(host_energy, d_e, (size_t)n_rows * 8, cudaMemcpyDeviceToHost, st), "energy read") &&
              cuda_ok(cudaStreamSynchronize(st), "checksum sync");
    cudaFree(d_sum); cudaFree(d_e);
    return ok ? ZA_OK : ZA_ERR_CUDA;
}

// The same reduction as za_gpu_checksum with the accumulators in the static device array above: no cudaMalloc / cudaFree,
// here or in anything it calls.  The array keeps the previous call's sums, so the cells this call uses are cleared first (on
// the caller's stream, ahead of the kernel).  All calls on a device share the array: calls that overlap on the GPU race.
ZA_API int za_gpu_checksum_noalloc(const float* dev, int n_rows, int64_t n_frames, int64_t frame_stride, uint64_t* host_sum,
                                   double* host_energy, int device, void* cuda_stream)
{
    if (!dev || n_rows <= 0 || n_frames < 0 || frame_stride < n_frames || !host_sum || !host_energy) {
        set_error("za_gpu_checksum_noalloc: bad argument");
        return ZA_ERR_ARG;
    }
    if (n_rows > kNoAllocMaxRows) {
        set_error("za_gpu_checksum_noalloc: " + std::to_string(n_rows) + " rows exceed the static scratch (" + std::to_string(kNoAllocMaxRows) +
                  " rows per call); split the rows over several calls or use za_gpu_checksum");
        return ZA_ERR_ARG;
    }
    if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice")) return ZA_ERR_CUDA;
    unsigned long long* scratch = checksum_scratch(device);
    if (!scratch) return ZA_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    unsigned long long* d_sum = scratch;
    double* d_e = reinterpret_cast<double*>(scratch + kNoAllocMaxRows);
    if (!cuda_ok(cudaMemsetAsync(d_sum, 0, (size_t)n_rows * 8, st), "checksum scratch clear") ||
        !cuda_ok(cudaMemsetAsync(d_e, 0, (size_t)n_rows * 8, st), "energy scratch clear")) return ZA_ERR_CUDA;
    if (n_frames > 0) {
        const int bx = (int)std::min<int64_t>((n_frames + 255) / 256, 64);
        checksum_kernel<<<dim3(bx, n_rows), 256, 0, st>>>(dev, n_frames, frame_stride, d_sum, d_e);
    }
    const bool ok = cuda_ok(cudaGetLastError(), "checksum_kernel") &&
                    cuda_ok(cudaMemcpyAsync(host_sum, d_sum, (size_t)n_rows * 8, cudaMemcpyDeviceToHost, st), "checksum read") &&
                    cuda_ok(cudaMemcpyAsync(host_energy, d_e, (size_t)n_rows * 8, cudaMemcpyDeviceToHost, st), "energy read") &&
                    cuda_ok(cudaStreamSynchronize(st), "checksum sync");
    return ok ? ZA_OK : ZA_ERR_CUDA;
}

}  // extern "C"
