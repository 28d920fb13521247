// p_modtilt.cu -- Faust leaf "ModTilt": host half.  mydsp::init() constants and the per-compute() slider-rate values
// (Faust's fSlow expressions, ModTilt.dsp:25-45) are evaluated here in float with the host libm, one row per host
// block and stream; the per-frame work is k_modtilt.cu.  Faust leaves compute in float32 only
// (cmake/plugin/CMakeLists.txt:144 passes no -double), so ZA_F64 is refused at create.
//
// Parameter API: FaustJuceProcessor collects the leaf's widgets in buildUserInterface order
// (src/FaustJuceProcessor.cpp:162-309) -> Tilt (dB), Pivot (Hz), Mix; the C ABI's param_block columns follow that
// order.  The layout is written in the library's slider-line form (default<min,max,step>label) so one parser and one
// grid quantiser serve every plugin; the quantised value is narrowed to the float zone (:467-471).
#include <cmath>
#include <cstring>

#include "za_host.h"
#include "modtilt_scan_core.h"

namespace za {

namespace {

const char* kSliderText =
    "slider1:tilt_db=0<-6,3,0.1>Tilt (dB)\n"           // ModTilt.dsp:9-11
    "slider2:pivot_hz=3<2,5,0.01>Pivot (Hz)\n"         // :13-15
    "slider3:mix=1<0,1,0.001>Mix\n";                   // :17-19

float a_from_hz(float hz, float fs)                    // :25-26
{
    const float pi = 3.14159265358979323846f;
    return 1.0f - std::exp(((-2.0f * pi) * std::fmax(hz, 0.001f)) / fs);
}

int piv_grid_index(double q_pivot, double mn, double step, int n)     // the pivot slider's grid index: row of the scan kernel's power table
{
    if (!(step > 0.0) || n <= 0) return 0;
    const long long k = std::llround((q_pivot - mn) / step);
    return (int)(k < 0 ? 0 : (k > n - 1 ? n - 1 : k));
}

void modtilt_slow(const double* q, float fs, MTCoef& c, int piv_index = 0)
{
    const float tilt = (float)q[0], pivot = (float)q[1], mix = (float)q[2];
    c.a_piv = a_from_hz(pivot, fs);                                // :37
    c.na_piv = 1.0f - c.a_piv;
    c.g_hi = std::pow(10.0f, (tilt * 0.5f) / 20.0f);               // :44  ba.db2linear(x) = pow(10, x/20)
    c.g_lo = std::pow(10.0f, ((-tilt) * 0.5f) / 20.0f);            // :45
    c.mix = mix; c.nmix = 1.0f - mix;
    std::memcpy(&c.piv_index, &piv_index, sizeof(float));
    c.pad1 = 0.0f;
}

// ---- device-side parameter path -----------------------------------------------------------------------------------
// A slider's quantised value is min + k * step for an integer k (the host grid, za_params.cpp): the slider-rate
// expressions are tabulated on the host with the host libm for every k, and a kernel maps the raw host floats of a
// whole automation trajectory to k with the quantiser's own double arithmetic (clamp, llround((v - min) / step)) and
// copies the table entries into coefficient rows.  Same bits as the host path, no per-row exp / pow on host cores.
struct MTGrid { double mn[3], mx[3], step[3]; int n[3]; };

__global__ void modtilt_param_kernel(const float* __restrict__ raw, MTCoef* __restrict__ coef, size_t rows, MTGrid g,
                                     const float2* __restrict__ t_tilt, const float2* __restrict__ t_piv, const float2* __restrict__ t_mix)
{
    for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (size_t)gridDim.x * blockDim.x) {
        int k[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double v = (double)raw[r * 3 + j];
            v = v < g.mn[j] ? g.mn[j] : (v > g.mx[j] ? g.mx[j] : v);
            long long q = llround(__ddiv_rn(__dsub_rn(v, g.mn[j]), g.step[j]));
            k[j] = (int)(q < 0 ? 0 : (q > g.n[j] - 1 ? g.n[j] - 1 : q));
        }
        const float2 a = t_tilt[k[0]], b = t_piv[k[1]], c = t_mix[k[2]];
        float4* o = reinterpret_cast<float4*>(coef + r);
        o[0] = make_float4(b.x, b.y, a.x, a.y);        // a_piv, na_piv, g_lo, g_hi
        o[1] = make_float4(c.x, c.y, __int_as_float(k[1]), 0.f);        // mix, nmix, pivot grid index
    }
}

class ModTilt : public PluginImpl {
    float* d_raw2[2] = {nullptr, nullptr}; MTCoef* d_coef2[2] = {nullptr, nullptr}; size_t raw_cap = 0;   // per-segment raw rows / coefficient rows, double buffered
    cudaStream_t up_stream = nullptr;
    cudaEvent_t ev_ready[2] = {}, ev_consumed[2] = {};
    float2* d_tab = nullptr;                          // tilt | pivot | mix tables
    MTGrid grid{};
    bool grid_ok = false;

    float* d_q = nullptr;                              // scan kernel: Q rows [4 fixed stages][32] then [pivot grid][32]
    mtscan::Consts scanK{};
    bool scan_ready = false;

    float* d_state = nullptr;
    MTCoef* d_coef = nullptr;
    size_t coef_cap = 0;
    std::vector<MTCoef> h_coef, h_coef_dev;
    bool coef_valid = false;
    int n_sm = 148;
    float fs = 48000.0f;

public:
    ~ModTilt() override
    {
        cudaFree(d_state); cudaFree(d_coef); cudaFree(d_tab); cudaFree(d_q);
        for (int k = 0; k < 2; ++k) {
            cudaFree(d_raw2[k]); cudaFree(d_coef2[k]);
            if (ev_ready[k]) cudaEventDestroy(ev_ready[k]);
            if (ev_consumed[k]) cudaEventDestroy(ev_consumed[k]);
        }
        if (up_stream) cudaStreamDestroy(up_stream);
    }

    bool build_tables(Instance& I)
    {
        std::vector<float2> tab;
        for (int j = 0; j < 3; ++j) {
            const SliderDecl& d = I.decls[j];
            grid.mn[j] = (double)d.min; grid.mx[j] = (double)d.max; grid.step[j] = (double)d.step;
            if (!(d.step > 0.0f) || d.is_choice) return true;                  // no grid: host path only
            const long long n = llround((grid.mx[j] - grid.mn[j]) / grid.step[j]) + 2;
            if (n > (1 << 20)) return true;
            grid.n[j] = (int)n;
            for (long long k = 0; k < n; ++k) {
                // every k the device can produce equals llround((v - min) / step) of a clamped v, so this is the
                // quantiser's own second half (za_params.cpp / JSFXJuceProcessor.cpp:5583-5594)
                double v = grid.mn[j] + (double)k * grid.step[j];
                v = v < grid.mn[j] ? grid.mn[j] : (v > grid.mx[j] ? grid.mx[j] : v);
                double q[3] = {0.0, 3.0, 1.0};
                q[j] = v;
                MTCoef c; modtilt_slow(q, fs, c);
                tab.push_back(j == 0 ? make_float2(c.g_lo, c.g_hi) : j == 1 ? make_float2(c.a_piv, c.na_piv) : make_float2(c.mix, c.nmix));
            }
        }
        if (!cuda_ok(cudaMalloc(&d_tab, tab.size() * sizeof(float2)), "modtilt tables") ||
            !cuda_ok(cudaMemcpy(d_tab, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice), "modtilt tables upload")) return false;
        grid_ok = true;
        // the scan kernel's powers: (na^8)^j, j = 0 .. 31, for the four fixed stages and for every pivot grid value
        const float a_env = a_from_hz(25.0f, fs), a_base = a_from_hz(1.0f, fs), a_trim = a_from_hz(0.2f, fs);
        scanK.env = mtscan::make_stage(a_env, 1.0f - a_env);
        scanK.base = mtscan::make_stage(a_base, 1.0f - a_base);
        scanK.ratio = mtscan::make_stage(0.05f, 1.0f - 0.05f);
        scanK.trim = mtscan::make_stage(a_trim, 1.0f - a_trim);
        std::vector<float> q((size_t)(4 + grid.n[1]) * 32);
        mtscan::power_row(scanK.env.na, &q[0]); mtscan::power_row(scanK.base.na, &q[32]);
        mtscan::power_row(scanK.ratio.na, &q[64]); mtscan::power_row(scanK.trim.na, &q[96]);
        for (int k = 0; k < grid.n[1]; ++k) mtscan::power_row(tab[(size_t)grid.n[0] + k].y, &q[(size_t)(4 + k) * 32]);
        if (!cuda_ok(cudaMalloc(&d_q, q.size() * sizeof(float)), "modtilt scan tables") ||
            !cuda_ok(cudaMemcpy(d_q, q.data(), q.size() * sizeof(float), cudaMemcpyHostToDevice), "modtilt scan tables upload")) return false;
        scan_ready = true;
        return true;
    }

    // Device-side parameter path, pipelined over TIME SEGMENTS of the call: the raw rows of segment k + 1 are uploaded
    // and turned into coefficient rows on a side stream (two buffers) while the render kernel of segment k runs, so a
    // 60 s call with per-block automation (5625 x 4096 rows, 276 MB of host floats) does not wait for its parameters
    // before the first frame (the one-shot upload was 8.4 ms of a 26.8 ms step at 4096 x 20 s).
    int run_blocks_raw(Instance& I, const float* params, int nbq, const float* in, float* out, double* tap,
                       int64_t n, int64_t stride) override
    {
        if (!grid_ok || tap || I.opt_force_generic) return 0;      // "force_generic_io": the host parameter path (tests compare the two)
        const int64_t bl = I.block_len;
        const int64_t n_blocks = (n + bl - 1) / bl;
        // >= 4 segments of >= 32 host blocks each (short calls: one segment)
        const int64_t seg_blocks = std::max<int64_t>(32, (n_blocks + 7) / 8);
        const size_t seg_rows = (size_t)std::min<int64_t>(seg_blocks, nbq) * I.n_streams;
        if (seg_rows > raw_cap) {
            for (int k = 0; k < 2; ++k) { cudaFree(d_raw2[k]); cudaFree(d_coef2[k]); d_raw2[k] = nullptr; d_coef2[k] = nullptr; }
            raw_cap = 0;
            for (int k = 0; k < 2; ++k)
                if (!cuda_ok(cudaMalloc(&d_raw2[k], seg_rows * 3 * sizeof(float)), "modtilt raw params") ||
                    !cuda_ok(cudaMalloc(&d_coef2[k], seg_rows * sizeof(MTCoef)), "modtilt coef")) return -1;
            raw_cap = seg_rows;
        }
        if (!up_stream) {
            if (!cuda_ok(cudaStreamCreateWithFlags(&up_stream, cudaStreamNonBlocking), "modtilt upload stream")) return -1;
            for (int k = 0; k < 2; ++k) {
                cudaEventCreateWithFlags(&ev_ready[k], cudaEventDisableTiming);
                cudaEventCreateWithFlags(&ev_consumed[k], cudaEventDisableTiming);
            }
        }
        coef_valid = false;
        int k = 0;
        for (int64_t b0 = 0; b0 < n_blocks; b0 += seg_blocks, ++k) {
            const int buf = k & 1;
            const int64_t b1 = std::min<int64_t>(n_blocks, b0 + seg_blocks);
            const int64_t r0 = std::min<int64_t>(b0, nbq - 1);                      // the last row persists beyond nbq
            const int rows_here = (int)std::max<int64_t>(1, std::min<int64_t>(b1, nbq) - r0);
            const size_t rows = (size_t)rows_here * I.n_streams;
            cudaStreamWaitEvent(up_stream, ev_consumed[buf], 0);                   // the kernel two segments back has read this buffer
            if (!cuda_ok(cudaMemcpyAsync(d_raw2[buf], params + (size_t)r0 * I.n_streams * 3, rows * 3 * sizeof(float), cudaMemcpyHostToDevice, up_stream),
                         "modtilt params upload")) return -1;
            const unsigned blocks = (unsigned)std::min<size_t>((rows + 255) / 256, (size_t)n_sm * 8);
            modtilt_param_kernel<<<blocks, 256, 0, up_stream>>>(d_raw2[buf], d_coef2[buf], rows, grid, d_tab, d_tab + grid.n[0], d_tab + grid.n[0] + grid.n[1]);
            if (!cuda_ok(cudaGetLastError(), "modtilt param kernel")) return -1;
            ++I.launches;
            cudaEventRecord(ev_ready[buf], up_stream);
            cudaStreamWaitEvent(I.stream, ev_ready[buf], 0);
            const int64_t f0 = b0 * bl, f1 = std::min<int64_t>(n, b1 * bl);
            if (!launch(I, d_coef2[buf], rows_here, in + f0, out + f0, f1 - f0, stride)) return -1;
            cudaEventRecord(ev_consumed[buf], I.stream);
        }
        return 1;
    }

    bool launch(Instance& I, const MTCoef* coef, int nbq, const float* in, float* out, int64_t n, int64_t stride)
    {
        MTArgs a{};
        a.in = in; a.out = out; a.row_stride = stride; a.n_frames = n;
        a.n_streams = I.n_streams; a.block_len = I.block_len; a.n_coef_blocks = nbq; a.coef = coef; a.state = d_state;
        a.a_env = a_from_hz(25.0f, fs); a.a_base = a_from_hz(1.0f, fs); a.a_trim = a_from_hz(0.2f, fs);   // :35-36,40
        // option "warps_per_stream" = streams per warp (4 / 8 / 16 / 32) + 100 x frames per run (default 8; 4 / 16 / 32 are measurement variants)
        const int ol = I.opt_warps % 100, run = I.opt_warps / 100 ? I.opt_warps / 100 : 8;
        const int lanes = (ol == 4 || ol == 8 || ol == 16 || ol == 32) ? ol : modtilt_pick_lanes(I.n_streams, n_sm);
        // Time-parallel scan kernel for the whole host blocks of the call when they are whole 256-frame tiles; the
        // stream-per-thread kernel renders the ragged tail (and every other shape).  "scalar_f32" = 1 forces the latter.
        int64_t n_main = 0;
        if (scan_ready && !I.opt_scalar_f32 && a.block_len % mtscan::kTile == 0 && n >= a.block_len) {
            MTScanArgs S{};
            S.a = a; S.a.n_frames = n / a.block_len * a.block_len;
            if (modtilt_scan_ok(S.a)) {
                S.K = scanK; S.qfix = d_q; S.qpiv = d_q + 4 * 32;
                const bool timed = I.time_begin();
                if (!modtilt_scan_launch(S, I.stream, I.opt_nbuf == 2)) return false;     // option "tile_buffers" = 2: the correctly rounded maps (measurement / cross-check)
                if (timed) I.time_end();
                ++I.launches;
                n_main = S.a.n_frames;
            }
        }
        if (n_main == n) return true;
        if (n_main > 0) {                                   // the tail starts on a host-block boundary: advance the rows
            const int64_t b0 = n_main / a.block_len;
            const int skip = (int)std::min<int64_t>(b0, nbq - 1);
            a.in += n_main; a.out += n_main; a.n_frames = n - n_main;
            a.coef += (size_t)skip * I.n_streams; a.n_coef_blocks = nbq - skip;
        }
        const bool timed = n_main == 0 && I.time_begin();
        if (!modtilt_launch(a, lanes, I.stream, run)) return false;
        if (timed) I.time_end();
        ++I.launches;
        return true;
    }

    bool create(Instance& I) override
    {
        if (I.precision != ZA_F32) { set_error("ModTilt: Faust leaves compute in float32 (no -double in the reference build); create it with ZA_F32"); I.create_status = ZA_ERR_UNSUPPORTED; return false; }
        cudaDeviceProp prop;
        if (!cuda_ok(cudaGetDeviceProperties(&prop, I.device), "device properties")) return false;
        n_sm = prop.multiProcessorCount;
        fs = std::fmin(192000.0f, std::fmax(1.0f, (float)(int)I.srate));   // ma.SR
        if (I.decls.size() != 3) { set_error("ModTilt: three parameters expected"); return false; }
        return cuda_ok(cudaMalloc(&d_state, (size_t)kMTState * I.n_streams * sizeof(float)), "modtilt state") && build_tables(I);
    }

    bool reset(Instance& I) override   // mydsp::init -> instanceClear: every recursive signal starts at 0
    {
        return cuda_ok(cudaMemsetAsync(d_state, 0, (size_t)kMTState * I.n_streams * sizeof(float), I.stream), "modtilt reset");
    }

    bool sliders_changed(Instance&, int, int, const std::vector<int>&) override { return true; }
    bool takes_block_params() const override { return true; }

    bool run(Instance& I, int s0, int ns, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        return run_blocks(I, s0, ns, &I.sliders[(size_t)s0 * I.decls.size()], 1, in, out, tap, n, stride);
    }

    bool run_blocks(Instance& I, int s0, int ns, const double* q, int nbq, const float* in, float* out, double* tap,
                    int64_t n, int64_t stride) override
    {
        if (s0 != 0 || ns != I.n_streams) { set_error("ModTilt: stream sub-ranges are not supported"); return false; }
        if (tap) { set_error("ModTilt: no double tap (float32 leaf)"); return false; }
        const int nd = (int)I.decls.size();
        h_coef.resize((size_t)nbq * ns);
        parallel_for(ns, 64, [&](int64_t sb, int64_t se) {
            for (int b = 0; b < nbq; ++b)
                for (int64_t s = sb; s < se; ++s) {
                    const double* sl = q + ((size_t)b * ns + s) * nd;
                    MTCoef& c = h_coef[(size_t)b * ns + s];
                    if (b > 0 && std::memcmp(sl, sl - (size_t)ns * nd, sizeof(double) * nd) == 0) c = h_coef[(size_t)(b - 1) * ns + s];
                    else modtilt_slow(sl, fs, c, piv_grid_index(sl[1], grid.mn[1], grid.step[1], grid_ok ? grid.n[1] : 0));
                }
        });
        const bool same = coef_valid && h_coef.size() == h_coef_dev.size() &&
                          std::memcmp(h_coef.data(), h_coef_dev.data(), h_coef.size() * sizeof(MTCoef)) == 0;
        if (!same) {
            coef_valid = false;
            if (h_coef.size() > coef_cap) {
                cudaFree(d_coef);
                if (!cuda_ok(cudaMalloc(&d_coef, h_coef.size() * sizeof(MTCoef)), "modtilt coef")) { coef_cap = 0; d_coef = nullptr; return false; }
                coef_cap = h_coef.size();
            }
            if (!cuda_ok(cudaMemcpyAsync(d_coef, h_coef.data(), h_coef.size() * sizeof(MTCoef), cudaMemcpyHostToDevice, I.stream), "modtilt coef upload"))
                return false;
            coef_valid = h_coef.size() <= (1u << 20);
            if (coef_valid) h_coef_dev = h_coef;
        }
        return launch(I, d_coef, nbq, in, out, n, stride);
    }

    void state_io(Instance& I, StateIO& io) override { io.dev(d_state, (size_t)kMTState * I.n_streams * sizeof(float)); }
    bool after_state_load(Instance&) override { coef_valid = false; return true; }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        static const char* const names[kMTState] = {"env2", "base", "lp_piv", "ratio", "mean_rdb"};
        int slot = -1;
        for (int k = 0; k < kMTState; ++k) if (!std::strcmp(names[k], name)) slot = k;
        if (slot < 0) { set_error(std::string("ModTilt: unknown state variable ") + name); return false; }
        std::vector<float> t((size_t)I.n_streams);
        if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync") ||
            !cuda_ok(cudaMemcpy(t.data(), d_state + (size_t)slot * I.n_streams, t.size() * sizeof(float), cudaMemcpyDeviceToHost), "state read"))
            return false;
        for (int s = 0; s < I.n_streams; ++s) out[s] = (double)t[s];
        return true;
    }
};

}  // namespace

PluginImpl* make_modtilt() { return new ModTilt(); }
const char* modtilt_slider_text() { return kSliderText; }

}  // namespace za
