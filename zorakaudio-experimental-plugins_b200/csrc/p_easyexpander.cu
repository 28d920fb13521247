// p_easyexpander.cu -- plugin "EasyExpander": host half (@init / @slider with glibc, so every
// coefficient is bit-identical to the compiled reference, which calls the same libm:
// dsp_jsfx_aot.py:3746-3754) + the launch plan.  Per-sample work is in k_easyexpander.cu.
#include <cmath>
#include <cstring>

#include "za_host.h"

namespace za {

namespace {

// plugins/Dynamics/EasyExpander/src/EasyExpander.jsfx:13-21 (the plugin's parameter API)
const char* kSliderText =
    "slider1:thresh_db=-40<-80,0,0.1>Threshold (dB)\n"
    "slider2:depth_db=24<0,60,1>Depth (dB)\n"
    "slider3:contour=50<0,100,1>Contour (0=gentle, 100=hard)\n"
    "slider4:det_hpf_hz=0<0,20000,1:log=1000>Detector HPF (Hz, 0=Off)\n"
    "slider5:det_lpf_hz=20000<20,20000,1:log=1000>Detector LPF (Hz, 20k=Off)\n";

const double kPi = 3.141592653589793;      // $pi, dsp_jsfx_aot.py:4283-4293

double onepole_coeff(double tau, double srate)      // EasyExpander.jsfx:51-54
{
    return tau <= 0.000001 ? 1.0 : 1.0 - std::exp(-1.0 / (tau * srate));
}

// @init (:27-171): band centres on an ERB-number grid, TPT-SVF g/k, normalised weights
void ee_init(double srate, EEConsts& c)
{
    const double ln10 = 2.302585092994046, eps = 0.000000000001;
    const int N = 8;
    c.band_c = onepole_coeff(0.030, srate);
    c.open_c = onepole_coeff(0.008, srate);
    c.close_fast_c = onepole_coeff(0.060, srate);
    c.close_slow_c = onepole_coeff(0.250, srate);
    c.slow_c = onepole_coeff(0.200, srate);
    const double erb_min = 21.4 * (std::log(0.00437 * 80.0 + 1.0) / ln10);
    const double erb_max = 21.4 * (std::log(0.00437 * 8000.0 + 1.0) / ln10);
    double sumw = 0.0;
    for (int b = 0; b < N; ++b) {
        const double e = erb_min + (erb_max - erb_min) * ((double)b / (double)(N - 1));
        const double fc = (std::pow(10.0, e / 21.4) - 1.0) / 0.00437;
        c.k[b] = 1.0 / 1.0;
        c.g[b] = std::tan(kPi * fc / srate);
        const double w = fc < 200.0 ? 0.50 : fc < 800.0 ? 0.80 : fc < 4000.0 ? 1.00 : 0.70;
        sumw += w;
        c.w[b] = w;
    }
    for (int b = 0; b < N; ++b) {
        c.w[b] = c.w[b] / (sumw > eps ? sumw : eps);
        c.den[b] = 1.0 + c.g[b] * (c.g[b] + c.k[b]);        // loop-invariant part of svf_bp (:83)
    }
}

// @slider (:154-170) for one stream; sl = thresh_db, depth_db, contour, det_hpf_hz, det_lpf_hz
void ee_slider(const double* sl, double srate, EECoef& c)
{
    c.thresh_db = sl[0]; c.depth_db = sl[1];
    const double cont = sl[2] / 100.0;
    double ratio_base = 1.5 + 2.5 * cont;
    if (ratio_base < 1.0) ratio_base = 1.0;
    c.ratio_base = ratio_base;
    c.hpf_on = sl[3] > 0.0;
    c.hpf_c = c.hpf_on ? (1.0 - std::exp(-2.0 * kPi * sl[3] / srate)) : 0.0;
    c.lpf_on = sl[4] < 20000.0;
    c.lpf_c = c.lpf_on ? (1.0 - std::exp(-2.0 * kPi * sl[4] / srate)) : 0.0;
}

class EasyExpander : public PluginImpl {
    EEConsts k_{};
    double* d_state = nullptr;
    EECoef* d_coef = nullptr;
    size_t coef_cap = 0;
    int n_sm = 148;
    std::vector<EECoef> h_coef;

public:
    ~EasyExpander() override { cudaFree(d_state); cudaFree(d_coef); }

    bool create(Instance& I) override
    {
        ee_init(I.srate, k_);
        cudaDeviceProp prop;
        if (!cuda_ok(cudaGetDeviceProperties(&prop, I.device), "device properties")) return false;
        n_sm = prop.multiProcessorCount;
        return cuda_ok(cudaMalloc(&d_state, (size_t)I.n_streams * kEEState * sizeof(double)), "EasyExpander state");
    }

    bool reset(Instance& I) override   // @init: all state zero except det_db_m = -120 (:139-150)
    {
        std::vector<double> h((size_t)I.n_streams * kEEState, 0.0);
        for (int s = 0; s < I.n_streams; ++s) h[(size_t)29 * I.n_streams + s] = -120.0;
        return cuda_ok(cudaMemcpyAsync(d_state, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, I.stream), "EasyExpander reset") &&
               cuda_ok(cudaStreamSynchronize(I.stream), "EasyExpander reset sync");
    }

    // @slider only derives coefficients from the sliders; they are recomputed per block in run_blocks
    bool sliders_changed(Instance&, int, int, const std::vector<int>&) override { return true; }

    bool takes_block_params() const override { return true; }

    bool run(Instance& I, int s0, int ns, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        return run_blocks(I, s0, ns, &I.sliders[(size_t)s0 * I.decls.size()], 1, in, out, tap, n, stride);
    }

    bool run_blocks(Instance& I, int s0, int ns, const double* q, int nbq, const float* in, float* out, double* tap,
                    int64_t n, int64_t stride) override
    {
        if (s0 != 0 || ns != I.n_streams) { set_error("EasyExpander: stream sub-ranges are not supported"); return false; }
        const int nd = (int)I.decls.size();
        h_coef.resize((size_t)nbq * ns);
        parallel_for(ns, 64, [&](int64_t sb, int64_t se) {
            for (int64_t s = sb; s < se; ++s) {
                // exp() only when the detector-filter sliders moved (it is the expensive part of @slider)
                for (int b = 0; b < nbq; ++b) {
                    const double* sl = q + ((size_t)b * ns + s) * nd;
                    EECoef& c = h_coef[(size_t)b * ns + s];
                    if (b > 0 && sl[3] == sl[3 - (size_t)ns * nd] && sl[4] == sl[4 - (size_t)ns * nd]) {
                        const EECoef& p = h_coef[(size_t)(b - 1) * ns + s];
                        c = p;
                        c.thresh_db = sl[0]; c.depth_db = sl[1];
                        double rb = 1.5 + 2.5 * (sl[2] / 100.0);
                        c.ratio_base = rb < 1.0 ? 1.0 : rb;
                    } else {
                        ee_slider(sl, I.srate, c);
                    }
                }
            }
        });
        if (h_coef.size() > coef_cap) {
            cudaFree(d_coef);
            if (!cuda_ok(cudaMalloc(&d_coef, h_coef.size() * sizeof(EECoef)), "EasyExpander coef")) { coef_cap = 0; d_coef = nullptr; return false; }
            coef_cap = h_coef.size();
        }
        if (!cuda_ok(cudaMemcpyAsync(d_coef, h_coef.data(), h_coef.size() * sizeof(EECoef), cudaMemcpyHostToDevice, I.stream), "EasyExpander coef upload"))
            return false;
        EEArgs a{};
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n;
        a.n_streams = ns; a.block_len = I.block_len; a.n_coef_blocks = nbq; a.coef = d_coef; a.state = d_state; a.k = k_;
        const bool timed = I.time_begin();
        if (!easyexpander_launch(a, I.precision, n_sm, I.stream)) return false;
        if (timed) I.time_end();
        ++I.launches;
        // h_coef is pageable: the async copy was staged by the runtime before returning
        return true;
    }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        static const char* names[7] = {"hpf_lp", "lpf_y", "det_slow", "expanding", "gain_db_s", "det_db_m", "gr_db_m"};
        int idx = -1;
        for (int i = 0; i < 7; ++i) if (std::strcmp(name, names[i]) == 0) idx = 24 + i;
        if (idx < 0) { set_error(std::string("EasyExpander: unknown state variable ") + name); return false; }
        if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync")) return false;
        return cuda_ok(cudaMemcpy(out, d_state + (size_t)idx * I.n_streams, (size_t)I.n_streams * sizeof(double), cudaMemcpyDeviceToHost), "state read");
    }
};

}  // namespace

PluginImpl* make_easyexpander() { return new EasyExpander(); }
const char* easyexpander_slider_text() { return kSliderText; }

}  // namespace za
