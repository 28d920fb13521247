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

void modtilt_slow(const double* q, float fs, MTCoef& c)
{
    const float tilt = (float)q[0], pivot = (float)q[1], mix = (float)q[2];
    c.a_piv = a_from_hz(pivot, fs);                                // :37
    c.na_piv = 1.0f - c.a_piv;
    c.g_hi = std::pow(10.0f, (tilt * 0.5f) / 20.0f);               // :44  ba.db2linear(x) = pow(10, x/20)
    c.g_lo = std::pow(10.0f, ((-tilt) * 0.5f) / 20.0f);            // :45
    c.mix = mix; c.nmix = 1.0f - mix;
    c.pad0 = c.pad1 = 0.0f;
}

class ModTilt : public PluginImpl {
    float* d_state = nullptr;
    MTCoef* d_coef = nullptr;
    size_t coef_cap = 0;
    std::vector<MTCoef> h_coef, h_coef_dev;
    bool coef_valid = false;
    int n_sm = 148;
    float fs = 48000.0f;

public:
    ~ModTilt() override { cudaFree(d_state); cudaFree(d_coef); }

    bool create(Instance& I) override
    {
        if (I.precision != ZA_F32) { set_error("ModTilt: Faust leaves compute in float32 (no -double in the reference build); create it with ZA_F32"); return false; }
        cudaDeviceProp prop;
        if (!cuda_ok(cudaGetDeviceProperties(&prop, I.device), "device properties")) return false;
        n_sm = prop.multiProcessorCount;
        fs = std::fmin(192000.0f, std::fmax(1.0f, (float)(int)I.srate));   // ma.SR
        return cuda_ok(cudaMalloc(&d_state, (size_t)kMTState * I.n_streams * sizeof(float)), "modtilt state");
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
                    else modtilt_slow(sl, fs, c);
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
        MTArgs a{};
        a.in = in; a.out = out; a.row_stride = stride; a.n_frames = n;
        a.n_streams = ns; a.block_len = I.block_len; a.n_coef_blocks = nbq; a.coef = d_coef; a.state = d_state;
        a.a_env = a_from_hz(25.0f, fs); a.a_base = a_from_hz(1.0f, fs); a.a_trim = a_from_hz(0.2f, fs);   // :35-36,40
        const int lanes = (I.opt_warps == 8 || I.opt_warps == 16 || I.opt_warps == 32) ? I.opt_warps : modtilt_pick_lanes(ns, n_sm);
        const bool timed = I.time_begin();
        if (!modtilt_launch(a, lanes, I.stream)) return false;
        if (timed) I.time_end();
        ++I.launches;
        return true;
    }

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
