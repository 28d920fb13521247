// p_erbtilt.cu -- plugin "ERBTilt" (plugins/Spectral/ERBTilt/src/ERB Tilt.jsfx): host half of the hand-written scan kernel
// (k_erbtilt.cu / erbtilt_core.h).  @init runs here with the host libm (erbtilt_tables.h: the recursion coefficients are
// the reference's bits), @slider / @block / @sample run on the device.  The machine translation of the same script stays
// in the library as "ERBTilt.translated" (csrc/gen/g_ERBTilt.cu) and the tests compare the two.
// Not kept: the script's GUI snapshot arrays and `u.*` variables written in @block (:339-378) -- they feed @gfx only.
#include <cmath>
#include <cstring>

#include "za_host.h"
#include "erbtilt_args.h"
#include "erbtilt_tables.h"

namespace za {

namespace {

const char* kSliderText =
    "slider1:0<-18,18,0.1:log>Tilt (dB)\n"                  // ERB Tilt.jsfx:10-16
    "slider2:1000<20,20000,1:log>Pivot (Hz)\n"
    "slider3:100<0,100,1:sqr>Comp (Loudness) %\n"
    "slider4:50<0,100,1:sqr>Roughness Guard %\n";

class ErbTilt : public PluginImpl {
    erb::Init init_{};
    double* d_state = nullptr;
    void* d_tb_main = nullptr; void* d_tb_tail = nullptr;
    RawRows rows_;

    template <class T, int C>
    bool upload_tables()
    {
        erb::Tables<T, C> m{}; erb::Tables<T, 1> t{};
        erb::build_tables<T, C>(init_, m); erb::build_tables<T, 1>(init_, t);
        return cuda_ok(cudaMalloc(&d_tb_main, sizeof m), "ERBTilt tables") && cuda_ok(cudaMalloc(&d_tb_tail, sizeof t), "ERBTilt tables") &&
               cuda_ok(cudaMemcpy(d_tb_main, &m, sizeof m, cudaMemcpyHostToDevice), "ERBTilt tables upload") &&
               cuda_ok(cudaMemcpy(d_tb_tail, &t, sizeof t, cudaMemcpyHostToDevice), "ERBTilt tables upload");
    }

    bool launch(Instance& I, const double* d_q, int n_rows, const float* in, float* out, double* tap, int64_t n, int64_t stride)
    {
        ErbArgs a{};
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n;
        a.n_streams = I.n_streams; a.block_len = I.block_len; a.n_rows = n_rows; a.q = d_q; a.state = d_state;
        a.tb_main = d_tb_main; a.tb_tail = d_tb_tail; a.bc = init_.bc;
        const double den = init_.gain_tau * I.srate + 0.000001;                     // :335  alpha = exp(-samplesblock / (gain_tau*srate + 1e-6))
        a.alpha_full = std::exp(-(double)I.block_len / den);
        a.alpha_last = std::exp(-(double)(n % I.block_len ? n % I.block_len : I.block_len) / den);
        const bool timed = I.time_begin();
        if (!erbtilt_launch(a, I.precision, I.stream)) return false;
        if (timed) I.time_end();
        ++I.launches;
        rows_.consumed(I);
        return true;
    }

public:
    ~ErbTilt() override { cudaFree(d_state); cudaFree(d_tb_main); cudaFree(d_tb_tail); }

    bool create(Instance& I) override
    {
        if (I.decls.size() != 4) { set_error("ERBTilt: four sliders expected"); return false; }
        init_ = erb::build_filterbank(I.srate);
        const bool ok = I.precision ? upload_tables<double, kErbChunkF64>() : upload_tables<float, kErbChunkF32>();
        return ok && cuda_ok(cudaMalloc(&d_state, (size_t)kErbSlots * I.n_streams * sizeof(double)), "ERBTilt state");
    }

    bool reset(Instance& I) override   // @init: every mem[] array zero, global_gain = 1 (:209)
    {
        std::vector<double> h((size_t)kErbSlots * I.n_streams, 0.0);
        for (int s = 0; s < I.n_streams; ++s) h[(size_t)kErbGlobalGain * I.n_streams + s] = 1.0;
        rows_.invalidate();
        return cuda_ok(cudaMemcpyAsync(d_state, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, I.stream), "ERBTilt reset") &&
               cuda_ok(cudaStreamSynchronize(I.stream), "ERBTilt reset sync");
    }

    bool sliders_changed(Instance&, int, int, const std::vector<int>&) override { return true; }   // @slider runs on the device
    bool takes_block_params() const override { return true; }

    bool run(Instance& I, int s0, int ns, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        return run_blocks(I, s0, ns, &I.sliders[(size_t)s0 * I.decls.size()], 1, in, out, tap, n, stride);
    }

    bool run_blocks(Instance& I, int s0, int ns, const double* q, int nbq, const float* in, float* out, double* tap,
                    int64_t n, int64_t stride) override
    {
        if (s0 != 0 || ns != I.n_streams) { set_error("ERBTilt: stream sub-ranges are not supported"); return false; }
        const double* d_q = rows_.upload_quantised(I, q, (size_t)nbq * ns * 4);
        return d_q && launch(I, d_q, nbq, in, out, tap, n, stride);
    }

    int run_blocks_raw(Instance& I, const float* params, int nbq, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        if (I.opt_force_generic) return 0;                      // "force_generic_io": the host quantiser path (tests compare the two)
        rows_.invalidate();
        const double* d_q = rows_.upload(I, params, (size_t)nbq * I.n_streams);
        if (!d_q) return -1;
        return launch(I, d_q, nbq, in, out, tap, n, stride) ? 1 : -1;
    }

    void state_io(Instance& I, StateIO& io) override { io.dev(d_state, (size_t)kErbSlots * I.n_streams * sizeof(double)); }
    bool after_state_load(Instance&) override { rows_.invalidate(); return true; }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        static const char* const names[12] = {"global_gain", "tilt_db", "pivot_hz", "erb_pivot", "slope", "comp_mix", "guard", "Lpre", "Lpost",
                                              "target", "blend", "alpha"};
        int slot = -1;
        for (int k = 0; k < 12; ++k) if (!std::strcmp(names[k], name)) slot = kErbGlobalGain + k;
        // mem[] arrays by name and index: "lpL[3]", "preE[15]", "gBand[0]", ...
        static const struct { const char* n; int base, len; } arrays[] = {{"lpL", kErbLpL, 15}, {"lpR", kErbLpR, 15}, {"envF", kErbEnvF, 16},
            {"envS", kErbEnvS, 16}, {"preE", kErbPreE, 16}, {"postE", kErbPostE, 16}, {"gBand", kErbGBand, 16}};
        for (const auto& a : arrays) {
            const size_t ln = std::strlen(a.n);
            if (!std::strncmp(name, a.n, ln) && name[ln] == '[') {
                const int i = std::atoi(name + ln + 1);
                if (i >= 0 && i < a.len) slot = a.base + i;
            }
        }
        if (slot < 0) { set_error(std::string("ERBTilt: unknown state variable ") + name); return false; }
        if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync")) return false;
        return cuda_ok(cudaMemcpy(out, d_state + (size_t)slot * I.n_streams, (size_t)I.n_streams * sizeof(double), cudaMemcpyDeviceToHost), "state read");
    }
};

}  // namespace

PluginImpl* make_erbtilt() { return new ErbTilt(); }
const char* erbtilt_slider_text() { return kSliderText; }

}  // namespace za
