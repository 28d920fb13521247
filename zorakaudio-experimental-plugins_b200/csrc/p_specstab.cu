// p_specstab.cu -- plugin "SpectralStabilizer" (plugins/Spectral/SpectralStabilizer/src/Spectral Stabilizer.jsfx): host half of
// the hand-written scan kernel (k_specstab.cu / specstab_core.h).  @init runs here with the host libm (specstab_tables.h:
// the biquad coefficients are the reference's bits), @slider / @block / @sample run on the device.  The machine translation
// of the same script stays in the library as "SpectralStabilizer.translated" and the tests compare the two.
#include <cmath>
#include <cstring>

#include "za_host.h"
#include "specstab_args.h"
#include "specstab_tables.h"

namespace za {

namespace {

const char* kSliderText =
    "slider1:1.20<0.70,2.00,0.01:log=1.20>Sigma (spectral smoothness)\n"      // Spectral Stabilizer.jsfx:13-17
    "slider2:35<0,100,1:sqr=2>Depth (%)\n"
    "slider3:55<0,100,1:sqr=2>MotionBias (%)\n";

class SpecStab : public PluginImpl {
    sstab::Init init_{};
    double* d_state = nullptr;
    void* d_tb_main = nullptr; void* d_tb_tail = nullptr; void* d_pwA = nullptr; void* d_pwR = nullptr;
    RawRows rows_;

    template <class T, int C>
    bool upload_tables(int block_len)
    {
        sstab::Tables<T, C> m{}; sstab::Tables<T, 1> t{};
        sstab::build_tables<T, C>(init_, m); sstab::build_tables<T, 1>(init_, t);
        const std::vector<T> pa = sstab::power_table<T>(init_.k.atk_a, block_len), pr = sstab::power_table<T>(init_.k.rel_a, block_len);
        return cuda_ok(cudaMalloc(&d_tb_main, sizeof m), "SpectralStabilizer tables") && cuda_ok(cudaMalloc(&d_tb_tail, sizeof t), "SpectralStabilizer tables") &&
               cuda_ok(cudaMalloc(&d_pwA, pa.size() * sizeof(T)), "SpectralStabilizer tables") && cuda_ok(cudaMalloc(&d_pwR, pr.size() * sizeof(T)), "SpectralStabilizer tables") &&
               cuda_ok(cudaMemcpy(d_tb_main, &m, sizeof m, cudaMemcpyHostToDevice), "tables upload") &&
               cuda_ok(cudaMemcpy(d_tb_tail, &t, sizeof t, cudaMemcpyHostToDevice), "tables upload") &&
               cuda_ok(cudaMemcpy(d_pwA, pa.data(), pa.size() * sizeof(T), cudaMemcpyHostToDevice), "tables upload") &&
               cuda_ok(cudaMemcpy(d_pwR, pr.data(), pr.size() * sizeof(T), cudaMemcpyHostToDevice), "tables upload");
    }

    bool launch(Instance& I, const double* d_q, int n_rows, const float* in, float* out, double* tap, int64_t n, int64_t stride)
    {
        SsArgs a{};
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n;
        a.n_streams = I.n_streams; a.block_len = I.block_len; a.n_rows = n_rows; a.q = d_q; a.state = d_state;
        a.tb_main = d_tb_main; a.tb_tail = d_tb_tail; a.pwA = d_pwA; a.pwR = d_pwR; a.k = init_.k;
        const bool timed = I.time_begin();
        if (!specstab_launch(a, I.precision, I.stream)) return false;
        if (timed) I.time_end();
        ++I.launches;
        rows_.consumed(I);
        return true;
    }

public:
    ~SpecStab() override { cudaFree(d_state); cudaFree(d_tb_main); cudaFree(d_tb_tail); cudaFree(d_pwA); cudaFree(d_pwR); }

    bool create(Instance& I) override
    {
        if (I.decls.size() != 3) { set_error("SpectralStabilizer: three sliders expected"); return false; }
        init_ = sstab::build_init(I.srate);
        const bool ok = I.precision ? upload_tables<double, kSsChunkF64>(I.block_len) : upload_tables<float, kSsChunkF32>(I.block_len);
        return ok && cuda_ok(cudaMalloc(&d_state, (size_t)kSsSlots * I.n_streams * sizeof(double)), "SpectralStabilizer state");
    }

    bool reset(Instance& I) override   // @init (:158-180): every per-band state zero, gt = gs = 1
    {
        std::vector<double> h((size_t)kSsSlots * I.n_streams, 0.0);
        for (int b = 0; b < sstab::kBands; ++b)
            for (int s = 0; s < I.n_streams; ++s) {
                h[(size_t)(kSsCarry + 9 * sstab::kBands + b) * I.n_streams + s] = 1.0;
                h[(size_t)(kSsGt + b) * I.n_streams + s] = 1.0;
            }
        rows_.invalidate();
        return cuda_ok(cudaMemcpyAsync(d_state, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, I.stream), "SpectralStabilizer reset") &&
               cuda_ok(cudaStreamSynchronize(I.stream), "SpectralStabilizer reset sync");
    }

    bool sliders_changed(Instance&, int, int, const std::vector<int>&) override { return true; }
    bool takes_block_params() const override { return true; }

    bool run(Instance& I, int s0, int ns, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        return run_blocks(I, s0, ns, &I.sliders[(size_t)s0 * I.decls.size()], 1, in, out, tap, n, stride);
    }
    bool run_blocks(Instance& I, int s0, int ns, const double* q, int nbq, const float* in, float* out, double* tap, int64_t n,
                    int64_t stride) override
    {
        if (s0 != 0 || ns != I.n_streams) { set_error("SpectralStabilizer: stream sub-ranges are not supported"); return false; }
        const double* d_q = rows_.upload_quantised(I, q, (size_t)nbq * ns * 3);
        return d_q && launch(I, d_q, nbq, in, out, tap, n, stride);
    }
    int run_blocks_raw(Instance& I, const float* params, int nbq, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        if (I.opt_force_generic) return 0;
        rows_.invalidate();
        const double* d_q = rows_.upload(I, params, (size_t)nbq * I.n_streams);
        if (!d_q) return -1;
        return launch(I, d_q, nbq, in, out, tap, n, stride) ? 1 : -1;
    }

    void state_io(Instance& I, StateIO& io) override { io.dev(d_state, (size_t)kSsSlots * I.n_streams * sizeof(double)); }
    bool after_state_load(Instance&) override { rows_.invalidate(); return true; }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        static const char* const scal[4] = {"depth", "mb", "low_min_gain", "action_amt"};
        int slot = -1;
        for (int k = 0; k < 4; ++k) if (!std::strcmp(scal[k], name)) slot = kSsScalars + k;
        const double* konst = nullptr;
        if (!std::strcmp(name, "atk_a")) konst = &init_.k.atk_a; else if (!std::strcmp(name, "rel_a")) konst = &init_.k.rel_a;
        else if (!std::strcmp(name, "env_a")) konst = &init_.k.env_a; else if (!std::strcmp(name, "mot_a")) konst = &init_.k.mot_a;
        else if (!std::strcmp(name, "dbfac")) konst = &init_.k.dbfac; else if (!std::strcmp(name, "min_gain_lin")) konst = &init_.k.min_gain_lin;
        if (konst) { for (int s = 0; s < I.n_streams; ++s) out[s] = *konst; return true; }
        // mem[] arrays by the script's offsets' names: "z1L[3]", "envlog[11]", "mot[0]", "gs[5]", "gt[2]", "res[7]", "gn[10]", "gw[0]"
        static const struct { const char* n; int base, len; } arrays[] = {
            {"z1L", kSsCarry, 12}, {"z2L", kSsCarry + 12, 12}, {"z1R", kSsCarry + 24, 12}, {"z2R", kSsCarry + 36, 12}, {"envlin", kSsCarry + 48, 12},
            {"envlog", kSsCarry + 60, 12}, {"d1", kSsCarry + 72, 12}, {"d2", kSsCarry + 84, 12}, {"mot", kSsCarry + 96, 12}, {"gs", kSsCarry + 108, 12},
            {"gn", kSsGn, 11}, {"gw", kSsGw, 11}, {"res", kSsRes, 12}, {"gt", kSsGt, 12}};
        for (const auto& a : arrays) {
            const size_t ln = std::strlen(a.n);
            if (!std::strncmp(name, a.n, ln) && name[ln] == '[') {
                const int i = std::atoi(name + ln + 1);
                if (i >= 0 && i < a.len) slot = a.base + i;
            }
        }
        if (slot < 0) { set_error(std::string("SpectralStabilizer: unknown state variable ") + name); return false; }
        if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync")) return false;
        return cuda_ok(cudaMemcpy(out, d_state + (size_t)slot * I.n_streams, (size_t)I.n_streams * sizeof(double), cudaMemcpyDeviceToHost), "state read");
    }
};

}  // namespace

PluginImpl* make_specstab() { return new SpecStab(); }
const char* specstab_slider_text() { return kSliderText; }

}  // namespace za
