// p_stft.cu -- plugin "stft_mask2048": host half (@init tables and @slider with glibc, in the script's
// expression order) + the launch plan.  Per-sample / per-hop work is in k_stft.cu.
#include <cmath>
#include <cstring>

#include "za_host.h"

namespace za {

namespace {

// plugins/Workloads/stft_mask2048/src/stft_mask2048.jsfx (the plugin's parameter API)
const char* kSliderText =
    "slider1:amount=60<0,100,1>Mask amount (%)\n"
    "slider2:thresh=2<1,8,0.01>Peak over neighbours threshold (ratio)\n"
    "slider3:tau_ms=250<20,2000,1>Baseline time constant (ms)\n"
    "slider4:floor_db=-24<-60,0,0.1>Gain floor (dB)\n"
    "slider5:mix=100<0,100,1>Wet mix (%)\n";

// @slider for one stream; sl = amount, thresh, tau_ms, floor_db, mix
void stft_slider(const double* sl, double srate, StftCoef& c)
{
    c.amt = sl[0] / 100.0;
    c.thresh = sl[1];
    const double d = sl[2] * 0.001 * srate;
    c.ab = std::exp(-(double)kStftHop / (1.0 > d ? 1.0 : d));        // exp(-HOP / max(1, tau_ms*0.001*srate))
    c.bb = 1.0 - c.ab;
    c.floor_g = std::exp(sl[3] * 0.115129254649702);
    c.mixv = sl[4] / 100.0;
}

class StftMask : public PluginImpl {
    StftState* d_state = nullptr;
    StftCoef* d_coef = nullptr;
    double* d_tab = nullptr;          // win[2048] | syn[2048] | tw1[15][128][2] | tw2[15][8][2]
    int n_sm = 148;
    size_t coef_cap = 0;
    std::vector<StftCoef> h_coef, h_coef_dev;   // host rows of this call / rows currently in d_coef
    bool coef_valid = false;
    int64_t t_abs = 0;                // frames rendered since prepare

public:
    ~StftMask() override { cudaFree(d_state); cudaFree(d_coef); cudaFree(d_tab); }

    bool create(Instance& I) override
    {
        // @init: symmetric Hann, synthesis window = w / sum of w^2 over the 4 overlaps (Sample.jsfx:3503-3528)
        std::vector<double> tab(2 * kStftN + 2 * 15 * 128 + 2 * 15 * 8);
        double* win = tab.data(); double* syn = win + kStftN; double* tw1 = syn + kStftN; double* tw2 = tw1 + 2 * 15 * 128;
        for (int i = 0; i < kStftN; ++i) win[i] = 0.5 - 0.5 * std::cos(6.283185307179586 * i / (kStftN - 1));
        const int overlaps = kStftN / kStftHop;
        for (int i = 0; i < kStftN; ++i) {
            double sum = 0; int idx = i;
            for (int j = 0; j < overlaps; ++j) { sum += win[idx] * win[idx]; idx += kStftHop; if (idx >= kStftN) idx -= kStftN; }
            syn[i] = sum > 0.000000000001 ? win[i] / sum : win[i];
        }
        // the script's `x * scale * syn[i]` with scale = 1/2048: a power of two, so folding it into the table is exact
        for (int i = 0; i < kStftN; ++i) syn[i] *= 1.0 / kStftN;
        // twiddles of the 16 x 16 x 8 decomposition (k_stft.cu / stft_core.h)
        for (int q = 1; q < 16; ++q) {
            for (int t = 0; t < 128; ++t) {
                const double ang = 6.283185307179586 * (double)(t * q) / 2048.0;
                tw1[2 * ((q - 1) * 128 + t)] = std::cos(ang); tw1[2 * ((q - 1) * 128 + t) + 1] = -std::sin(ang);
            }
            for (int j = 0; j < 8; ++j) {
                const double ang = 6.283185307179586 * (double)(j * q) / 128.0;
                tw2[2 * ((q - 1) * 8 + j)] = std::cos(ang); tw2[2 * ((q - 1) * 8 + j) + 1] = -std::sin(ang);
            }
        }
        cudaDeviceProp prop;
        if (!cuda_ok(cudaGetDeviceProperties(&prop, I.device), "device properties")) return false;
        n_sm = prop.multiProcessorCount;
        return cuda_ok(cudaMalloc(&d_state, (size_t)I.n_streams * sizeof(StftState)), "stft state") &&
               cuda_ok(cudaMalloc(&d_tab, tab.size() * sizeof(double)), "stft tables") &&
               cuda_ok(cudaMemcpy(d_tab, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice), "stft tables upload");
    }

    bool reset(Instance& I) override   // @init: rings, accumulators and per-bin state zero; counters zero
    {
        t_abs = 0;
        return cuda_ok(cudaMemsetAsync(d_state, 0, (size_t)I.n_streams * sizeof(StftState), I.stream), "stft reset");
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
        if (s0 != 0 || ns != I.n_streams) { set_error("stft_mask2048: stream sub-ranges are not supported"); return false; }
        const int nd = (int)I.decls.size();
        h_coef.resize((size_t)nbq * ns);
        parallel_for(ns, 64, [&](int64_t sb, int64_t se) {
            for (int b = 0; b < nbq; ++b)
                for (int64_t s = sb; s < se; ++s) {
                    const double* sl = q + ((size_t)b * ns + s) * nd;
                    StftCoef& c = h_coef[(size_t)b * ns + s];
                    if (b > 0 && std::memcmp(sl, sl - (size_t)ns * nd, sizeof(double) * nd) == 0) c = h_coef[(size_t)(b - 1) * ns + s];
                    else stft_slider(sl, I.srate, c);
                }
        });
        // unchanged coefficient rows stay on the device (a pageable upload would synchronise the stream first)
        const bool same = coef_valid && h_coef.size() == h_coef_dev.size() &&
                          std::memcmp(h_coef.data(), h_coef_dev.data(), h_coef.size() * sizeof(StftCoef)) == 0;
        if (!same) {
            coef_valid = false;
            if (h_coef.size() > coef_cap) {
                cudaFree(d_coef);
                if (!cuda_ok(cudaMalloc(&d_coef, h_coef.size() * sizeof(StftCoef)), "stft coef")) { coef_cap = 0; d_coef = nullptr; return false; }
                coef_cap = h_coef.size();
            }
            if (!cuda_ok(cudaMemcpyAsync(d_coef, h_coef.data(), h_coef.size() * sizeof(StftCoef), cudaMemcpyHostToDevice, I.stream), "stft coef upload"))
                return false;
            coef_valid = h_coef.size() <= (1u << 20);      // cache small (static) tables only
            if (coef_valid) h_coef_dev = h_coef;
        }
        StftArgs a{};
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n; a.t0 = t_abs;
        a.n_streams = ns; a.block_len = I.block_len; a.n_coef_blocks = nbq; a.coef = d_coef; a.state = d_state;
        a.win = d_tab; a.syn = d_tab + kStftN; a.tw1 = d_tab + 2 * kStftN; a.tw2 = a.tw1 + 2 * 15 * 128;
        const bool timed = I.time_begin();
        if (!stft_launch(a, I.precision, n_sm, I.stream)) return false;
        if (timed) I.time_end();
        ++I.launches;
        t_abs += n;
        return true;
    }

    void state_io(Instance& I, StateIO& io) override
    {
        io.dev(d_state, (size_t)I.n_streams * sizeof(StftState));
        io.host(&t_abs, sizeof t_abs);                       // the hop phase (frames rendered since prepare)
    }
    bool after_state_load(Instance&) override { coef_valid = false; return true; }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        // integer state of the script: hop_ctr, ringpos, frames_done follow from the frame count
        double v;
        if (!std::strcmp(name, "hop_ctr")) v = (double)(t_abs % kStftHop);
        else if (!std::strcmp(name, "ringpos")) v = (double)(t_abs & 8191);
        else if (!std::strcmp(name, "frames_done")) v = (double)(t_abs / kStftHop);
        else { set_error(std::string("stft_mask2048: unknown state variable ") + name); return false; }
        for (int s = 0; s < I.n_streams; ++s) out[s] = v;
        return true;
    }
};

}  // namespace

PluginImpl* make_stft_mask2048() { return new StftMask(); }
const char* stft_mask2048_slider_text() { return kSliderText; }

}  // namespace za
