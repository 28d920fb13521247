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

// ---- device-side parameter path ------------------------------------------------------------------------------
struct SliderQ { double mn, mx, step; int is_choice; };          // quantise_slider's inputs (za_params.cpp)

__device__ __forceinline__ double quantise_dev(const SliderQ& d, float host_value)   // = za::quantise_slider, same operations
{
    double v = (double)host_value;
    if (d.is_choice) v = __dadd_rn(d.mn, __dmul_rn((double)llround(v), d.step));
    v = v < d.mn ? d.mn : (v > d.mx ? d.mx : v);
    if (!d.is_choice && d.step > 0.0) {
        const double q = (double)llround(__ddiv_rn(__dsub_rn(v, d.mn), d.step));
        v = __dadd_rn(d.mn, __dmul_rn(q, d.step));
        v = v < d.mn ? d.mn : (v > d.mx ? d.mx : v);
    }
    return v;
}

struct EEParamArgs {
    const float* params;       // [nb][ns][5] raw host values (device copy)
    const EECoef* per_stream;  // [ns]: hpf/lpf part of @slider, computed on the host with glibc (constant over the call)
    EECoef* out;               // [nb][ns]
    int nb, ns;
    SliderQ q[5];
    const double* tab_hpf;     // 1 - exp(-2 pi f / srate) for every grid value of the detector HPF slider (host libm), or null
    const double* tab_lpf;     // the same for the detector LPF slider
    int n_hpf, n_lpf;
};

// One thread per (block, stream): quantise threshold / depth / contour exactly as the host does and derive the
// @slider values that need no libm (EasyExpander.jsfx:154-160).
__global__ void ee_param_kernel(const EEParamArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)a.nb * a.ns) return;
    const int s = (int)(i % a.ns);
    const float* p = a.params + i * 5;
    EECoef c = a.per_stream[s];
    c.thresh_db = quantise_dev(a.q[0], p[0]);
    c.depth_db = quantise_dev(a.q[1], p[1]);
    const double cont = __ddiv_rn(quantise_dev(a.q[2], p[2]), 100.0);
    const double rb = __dadd_rn(1.5, __dmul_rn(2.5, cont));
    c.ratio_base = rb < 1.0 ? 1.0 : rb;
    if (a.tab_hpf) {
        // the detector-filter sliders may move inside the call: a quantised value is mn + k step for an integer k, and the
        // table holds the host libm's coefficient for exactly that double (:169-175), so this equals the host path bit for bit
        const double fh = quantise_dev(a.q[3], p[3]), fl = quantise_dev(a.q[4], p[4]);
        int kh = (int)llround(__ddiv_rn(__dsub_rn(fh, a.q[3].mn), a.q[3].step)), kl = (int)llround(__ddiv_rn(__dsub_rn(fl, a.q[4].mn), a.q[4].step));
        kh = kh < 0 ? 0 : (kh >= a.n_hpf ? a.n_hpf - 1 : kh); kl = kl < 0 ? 0 : (kl >= a.n_lpf ? a.n_lpf - 1 : kl);
        c.hpf_on = fh > 0.0; c.hpf_c = c.hpf_on ? a.tab_hpf[kh] : 0.0;
        c.lpf_on = fl < 20000.0; c.lpf_c = c.lpf_on ? a.tab_lpf[kl] : 0.0;
    }
    a.out[i] = c;
}

class EasyExpander : public PluginImpl {
    EEConsts k_{};
    double* d_state = nullptr;
    EECoef* d_coef = nullptr;
    size_t coef_cap = 0;
    float* d_params = nullptr;        // raw automation rows of the current call
    size_t params_cap = 0;
    EECoef* d_per_stream = nullptr;
    double* d_tab_hpf = nullptr; double* d_tab_lpf = nullptr;   // detector-filter coefficients per slider grid value (host libm)
    int n_tab_hpf = 0, n_tab_lpf = 0;
    cudaEvent_t ev_params = nullptr, ev_consumed = nullptr;   // upload done / previous rows consumed
    cudaStream_t up_stream = nullptr; // uploads overlap the previous call's kernel
    int n_sm = 148;
    std::vector<EECoef> h_coef, h_coef_dev;   // host rows of this call / rows currently in d_coef
    bool coef_valid = false;

public:
    ~EasyExpander() override { cudaFree(d_state); cudaFree(d_coef); cudaFree(d_params); cudaFree(d_per_stream); cudaFree(d_tab_hpf); cudaFree(d_tab_lpf); if (ev_params) cudaEventDestroy(ev_params); if (ev_consumed) cudaEventDestroy(ev_consumed); if (up_stream) cudaStreamDestroy(up_stream); }

    bool create(Instance& I) override
    {
        ee_init(I.srate, k_);
        cudaDeviceProp prop;
        if (!cuda_ok(cudaGetDeviceProperties(&prop, I.device), "device properties")) return false;
        n_sm = prop.multiProcessorCount;
        if (I.decls.size() == 5 && !build_filter_tables(I)) return false;
        return cuda_ok(cudaMalloc(&d_state, (size_t)I.n_streams * kEEState * sizeof(double)), "EasyExpander state");
    }

    // 1 - exp(-2 pi f / srate) for every value the detector HPF / LPF sliders can take after quantisation (20 001 + 19 981
    // values), from the host's libm, so the device parameter path serves calls in which those sliders move (it used to hand
    // the whole batch to the host path as soon as one stream moved one of them)
    bool build_filter_tables(Instance& I)
    {
        double* dst[2] = {nullptr, nullptr};
        int cnt[2] = {0, 0};
        for (int t = 0; t < 2; ++t) {
            const SliderDecl& d = I.decls[3 + t];
            const double mn = (double)d.min, mx = (double)d.max, step = (double)(d.step > 0.0f ? d.step : 1.0f);
            const int64_t n = (int64_t)std::llround((mx - mn) / step) + 1;
            if (d.is_choice || n < 1 || n > (1 << 20)) return true;          // unusual declaration: keep the per-stream host path
            std::vector<double> tab((size_t)n);
            for (int64_t k = 0; k < n; ++k) {
                double v = mn + (double)k * step;                            // quantise_slider's own expression
                v = v < mn ? mn : (v > mx ? mx : v);
                double sl[5] = {0.0, 0.0, 0.0, t == 0 ? v : 0.0, t == 1 ? v : 20000.0};
                EECoef c;
                ee_slider(sl, I.srate, c);
                tab[(size_t)k] = t == 0 ? c.hpf_c : c.lpf_c;
            }
            if (!cuda_ok(cudaMalloc(&dst[t], tab.size() * sizeof(double)), "EasyExpander filter table") ||
                !cuda_ok(cudaMemcpy(dst[t], tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice), "EasyExpander filter table upload")) {
                cudaFree(dst[0]); cudaFree(dst[1]);
                return false;
            }
            cnt[t] = (int)n;
        }
        d_tab_hpf = dst[0]; d_tab_lpf = dst[1]; n_tab_hpf = cnt[0]; n_tab_lpf = cnt[1];
        return true;
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
        // unchanged coefficient rows stay on the device (a pageable upload would synchronise the stream first)
        const bool same = coef_valid && h_coef.size() == h_coef_dev.size() &&
                          std::memcmp(h_coef.data(), h_coef_dev.data(), h_coef.size() * sizeof(EECoef)) == 0;
        if (!same) {
            coef_valid = false;
            if (h_coef.size() > coef_cap) {
                cudaFree(d_coef);
                if (!cuda_ok(cudaMalloc(&d_coef, h_coef.size() * sizeof(EECoef)), "EasyExpander coef")) { coef_cap = 0; d_coef = nullptr; return false; }
                coef_cap = h_coef.size();
            }
            if (!cuda_ok(cudaMemcpyAsync(d_coef, h_coef.data(), h_coef.size() * sizeof(EECoef), cudaMemcpyHostToDevice, I.stream), "EasyExpander coef upload"))
                return false;
            coef_valid = h_coef.size() <= (1u << 20);      // cache small (static) tables only
            if (coef_valid) h_coef_dev = h_coef;
        }
        EEArgs a{};
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n;
        a.n_streams = ns; a.block_len = I.block_len; a.n_coef_blocks = nbq; a.coef = d_coef; a.state = d_state; a.k = k_; a.role_off = I.opt_role_off;
        const bool timed = I.time_begin();
        if (!easyexpander_launch(a, I.precision, n_sm, I.stream, I.opt_scalar_f32)) return false;
        if (timed) I.time_end();
        ++I.launches;
        // h_coef is pageable: the async copy was staged by the runtime before returning
        return true;
    }

    // Automation on the device: upload the raw float rows, quantise + derive there.  The detector filter sliders (the only
    // @slider inputs that go through exp()) take their coefficients from per-grid-value tables made with the host's glibc
    // (build_filter_tables), so every value stays bit-identical to the host path whether or not they move inside the call.
    int run_blocks_raw(Instance& I, const float* params, int nb, const float* in, float* out, double* tap,
                       int64_t n, int64_t stride) override
    {
        const int ns = I.n_streams, nd = (int)I.decls.size();
        if (nd != 5) return 0;
        std::vector<uint8_t> moved(ns, 0);
        if (!d_tab_hpf) parallel_for(ns, 64, [&](int64_t sb, int64_t se) {
            for (int64_t s = sb; s < se; ++s) {
                const float h = params[(size_t)s * nd + 3], l = params[(size_t)s * nd + 4];
                for (int b = 1; b < nb; ++b) {
                    const float* p = params + ((size_t)b * ns + s) * nd;
                    if (p[3] != h || p[4] != l) { moved[s] = 1; break; }
                }
            }
        });
        if (!d_tab_hpf)
            for (int s = 0; s < ns; ++s) if (moved[s]) return 0;
        std::vector<EECoef> ps(ns);
        for (int s = 0; s < ns; ++s) {
            double sl[5];
            for (int j = 0; j < 5; ++j) sl[j] = quantise_slider(I.decls[j], params[(size_t)s * nd + j]);
            ee_slider(sl, I.srate, ps[s]);
        }
        const size_t rows = (size_t)nb * ns;
        coef_valid = false;                                // d_coef is rewritten by the parameter kernel
        if (rows * nd > params_cap) {
            cudaFree(d_params); d_params = nullptr; params_cap = 0;
            if (!cuda_ok(cudaMalloc(&d_params, rows * nd * sizeof(float)), "EasyExpander params")) return -1;
            params_cap = rows * nd;
        }
        if (rows > coef_cap) {
            cudaFree(d_coef); d_coef = nullptr; coef_cap = 0;
            if (!cuda_ok(cudaMalloc(&d_coef, rows * sizeof(EECoef)), "EasyExpander coef")) return -1;
            coef_cap = rows;
        }
        if (!d_per_stream && !cuda_ok(cudaMalloc(&d_per_stream, (size_t)ns * sizeof(EECoef)), "EasyExpander per-stream coef")) return -1;
        if (!up_stream) {
            if (!cuda_ok(cudaStreamCreateWithFlags(&up_stream, cudaStreamNonBlocking), "param upload stream")) return -1;
            cudaEventCreateWithFlags(&ev_params, cudaEventDisableTiming);
            cudaEventCreateWithFlags(&ev_consumed, cudaEventDisableTiming);
        } else {
            cudaStreamWaitEvent(up_stream, ev_consumed, 0);      // the previous call's parameter kernel has read the buffers
        }
        if (!cuda_ok(cudaMemcpyAsync(d_per_stream, ps.data(), ps.size() * sizeof(EECoef), cudaMemcpyHostToDevice, up_stream), "per-stream coef upload"))
            return -1;
        // Long calls are cut into time segments of whole host blocks: the rows of segment k + 1 upload (and, from pageable
        // memory, are staged by the host) while segment k's kernel runs, so only the first segment's upload is exposed
        // (cfg3: 460 MB of rows per 60 s call, ~9 ms at PCIe rate, was all in front of the one launch).  State carries from
        // launch to launch in the state rows like between calls.
        const int64_t total_blocks = (n + I.block_len - 1) / I.block_len;
        int nseg = 1;
        if (nb >= 512 && nb >= total_blocks) nseg = (int)std::min<int64_t>(8, nb / 256);
        const int seg_blocks = (nb + nseg - 1) / nseg;
        EEParamArgs pa{};
        pa.per_stream = d_per_stream; pa.ns = ns;
        pa.tab_hpf = d_tab_hpf; pa.tab_lpf = d_tab_lpf; pa.n_hpf = n_tab_hpf; pa.n_lpf = n_tab_lpf;
        for (int j = 0; j < 5; ++j) {
            const SliderDecl& d = I.decls[j];
            pa.q[j] = SliderQ{(double)d.min, (double)d.max, (double)(d.step > 0.0f ? d.step : 1.0f), d.is_choice ? 1 : 0};
        }
        for (int k = 0; k < nseg; ++k) {
            const int b0 = k * seg_blocks, b1 = std::min(nb, b0 + seg_blocks);
            if (b0 >= b1) break;
            const size_t r0 = (size_t)b0 * ns, nr = (size_t)(b1 - b0) * ns;
            if (!cuda_ok(cudaMemcpyAsync(d_params + r0 * nd, params + r0 * nd, nr * nd * sizeof(float), cudaMemcpyHostToDevice, up_stream), "params upload"))
                return -1;
            cudaEventRecord(ev_params, up_stream);
            cudaStreamWaitEvent(I.stream, ev_params, 0);
            pa.params = d_params + r0 * nd; pa.out = d_coef + r0; pa.nb = b1 - b0;
            ee_param_kernel<<<(unsigned)((nr + 255) / 256), 256, 0, I.stream>>>(pa);
            if (!cuda_ok(cudaGetLastError(), "ee_param_kernel launch")) return -1;
            if (k == nseg - 1 || b1 == nb) cudaEventRecord(ev_consumed, I.stream);
            ++I.launches;
            const int64_t f0 = (int64_t)b0 * I.block_len;
            const int64_t f1 = (k == nseg - 1 || b1 == nb) ? n : std::min<int64_t>(n, (int64_t)b1 * I.block_len);
            if (f1 <= f0) continue;
            EEArgs a{};
            a.in = in + f0; a.out = out + f0; a.tap = tap ? tap + f0 : nullptr; a.row_stride = stride; a.n_frames = f1 - f0;
            a.n_streams = ns; a.block_len = I.block_len; a.n_coef_blocks = b1 - b0; a.coef = d_coef + r0; a.state = d_state; a.k = k_;
            a.role_off = I.opt_role_off;
            const bool timed = I.time_begin();
            if (!easyexpander_launch(a, I.precision, n_sm, I.stream, I.opt_scalar_f32)) return -1;
            if (timed) I.time_end();
            ++I.launches;
        }
        return 1;
    }

    void state_io(Instance& I, StateIO& io) override { io.dev(d_state, (size_t)I.n_streams * kEEState * sizeof(double)); }
    bool after_state_load(Instance&) override { coef_valid = false; return true; }

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
