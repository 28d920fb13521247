// p_cascade8.cu -- plugin "cascade8": host half (@init / @slider) + device launch plan.
//
// @slider runs on the host with glibc's sin/cos/exp so the coefficients are bit-identical to what
// the reference's compiled @slider produces (it calls the same libm, dsp_jsfx_aot.py:3746-3754);
// a 1-ulp difference in a1/a2 of a 40 Hz biquad moves its response by ~1e-11, which would already
// break the 1e-12 fp64 criterion.  Everything per-sample runs in k_cascade8.cu.
#include <cmath>
#include <cstring>

#include "za_host.h"

namespace za {

namespace {

const char* kSliderText =
    "slider1:fc1=80<20,20000,0.01>S1 HP fc (Hz)\n"
    "slider2:q1=0.707<0.1,8,0.001>S1 HP Q\n"
    "slider3:fc2=12000<20,20000,0.01>S2 LP fc (Hz)\n"
    "slider4:q2=0.707<0.1,8,0.001>S2 LP Q\n"
    "slider5:fc3=1000<20,20000,0.01>S3 BP(peak) fc (Hz)\n"
    "slider6:q3=1<0.1,8,0.001>S3 BP(peak) Q\n"
    "slider7:fc4=3000<20,20000,0.01>S4 BP(skirt) fc (Hz)\n"
    "slider8:q4=1<0.1,8,0.001>S4 BP(skirt) Q\n"
    "slider9:fc5=8000<20,20000,0.01>S5 LP fc (Hz)\n"
    "slider10:q5=0.707<0.1,8,0.001>S5 LP Q\n"
    "slider11:fc6=40<20,20000,0.01>S6 HP fc (Hz)\n"
    "slider12:q6=0.707<0.1,8,0.001>S6 HP Q\n"
    "slider13:fc7=16000<20,20000,0.01>S7 one-pole LP fc (Hz)\n"
    "slider14:fc8=18000<20,20000,0.01>S8 one-pole LP fc (Hz)\n"
    "slider15:mix=1<0,1,0.001>S3/S4 band mix\n";

inline double jmin(double a, double b) { return a < b ? a : b; }   // olt/ogt selects, dsp_jsfx_aot.py:5223-5232
inline double jmax(double a, double b) { return a > b ? a : b; }

// cascade8.jsfx @slider, expression order kept
void c8_slider(const double* sl, double srate, double* c /* kC8Coefs */)
{
    const double two_pi = 6.283185307179586;
    static const int kind[6] = {0, 1, 2, 3, 1, 0};    // HP LP BPpeak BPskirt LP HP
    for (int k = 0; k < 6; ++k) {
        const double f = jmin(jmax(sl[2 * k], 10), srate * 0.45), q = jmax(sl[2 * k + 1], 0.1);
        const double w0 = two_pi * f / srate, cw = std::cos(w0), sw = std::sin(w0), alpha = sw / (2.0 * q);
        const double aa0 = 1.0 + alpha;
        double* b = c + 5 * k;
        switch (kind[k]) {
        case 0: b[0] = ((1.0 + cw) / 2.0) / aa0; b[1] = (-(1.0 + cw)) / aa0; b[2] = ((1.0 + cw) / 2.0) / aa0; break;
        case 1: b[0] = ((1.0 - cw) / 2.0) / aa0; b[1] = (1.0 - cw) / aa0;    b[2] = ((1.0 - cw) / 2.0) / aa0; break;
        case 2: b[0] = alpha / aa0;      b[1] = 0.0; b[2] = (-alpha) / aa0;    break;
        default: b[0] = (sw / 2.0) / aa0; b[1] = 0.0; b[2] = (-sw / 2.0) / aa0; break;
        }
        b[3] = (-2.0 * cw) / aa0; b[4] = (1.0 - alpha) / aa0;
    }
    const double a7 = std::exp(-two_pi * jmin(sl[12], srate * 0.45) / srate);
    c[30] = a7; c[31] = 1 - a7;
    c[32] = 1.0 - std::exp(-two_pi * jmin(sl[13], srate * 0.45) / srate);
    c[33] = sl[14]; c[34] = 0; c[35] = 0;
}

class Cascade8 : public PluginImpl {
    std::vector<double> h_coef;
    double* d_coef = nullptr;
    void* d_tables = nullptr;
    void* d_tables_x2 = nullptr;     // fp32 mode: tables of the packed-FP32 kernel
    double* d_state = nullptr;

public:
    ~Cascade8() override { cudaFree(d_coef); cudaFree(d_tables); cudaFree(d_tables_x2); cudaFree(d_state); cudaFree(d_rows[0]); cudaFree(d_rows[1]); }

    bool create(Instance& I) override
    {
        h_coef.assign((size_t)I.n_streams * kC8Coefs, 0.0);
        if (!I.precision &&
            !cuda_ok(cudaMalloc(&d_tables_x2, (size_t)I.n_streams * cascade8_x2_tables_bytes()), "cascade8 x2 tables"))
            return false;
        return cuda_ok(cudaMalloc(&d_coef, h_coef.size() * sizeof(double)), "cascade8 coef") &&
               cuda_ok(cudaMalloc(&d_tables, (size_t)I.n_streams * cascade8_tables_bytes(I.precision)), "cascade8 tables") &&
               cuda_ok(cudaMalloc(&d_state, (size_t)I.n_streams * kC8State * sizeof(double)), "cascade8 state");
    }

    bool reset(Instance& I) override   // @init: all filter state zero
    {
        return cuda_ok(cudaMemsetAsync(d_state, 0, (size_t)I.n_streams * kC8State * sizeof(double), I.stream), "cascade8 reset");
    }

    bool sliders_changed(Instance& I, int s0, int ns, const std::vector<int>& streams) override
    {
        const int nd = (int)I.decls.size();
        for (int s : streams) c8_slider(&I.sliders[(size_t)s * nd], I.srate, &h_coef[(size_t)s * kC8Coefs]);
        // The copy source must stay valid until the copy ran: h_coef is pageable, so the async copy
        // is staged by the runtime before returning.
        if (!cuda_ok(cudaMemcpyAsync(d_coef + (size_t)s0 * kC8Coefs, &h_coef[(size_t)s0 * kC8Coefs],
                                     (size_t)ns * kC8Coefs * sizeof(double), cudaMemcpyHostToDevice, I.stream), "cascade8 coef upload"))
            return false;
        ++I.launches;
        if (d_tables_x2) {
            ++I.launches;
            if (!cascade8_x2_build_tables(d_coef + (size_t)s0 * kC8Coefs,
                                          (char*)d_tables_x2 + (size_t)s0 * cascade8_x2_tables_bytes(), ns, I.stream))
                return false;
        }
        return cascade8_build_tables(d_coef + (size_t)s0 * kC8Coefs,
                                     (char*)d_tables + (size_t)s0 * cascade8_tables_bytes(I.precision), ns, I.precision, I.stream);
    }

    // Per-block automation: the whole trajectory in one launch per 256-block segment (k_cascade8_auto.cu) instead of one
    // launch, one upload and two table kernels per host block in which a slider moved.  @slider runs on the host with the host
    // libm for every (block, stream) whose quantised sliders differ from the previous block's (other rows are copies, which is
    // what tells the kernel to keep its tables).  Returns 0 (not taken) when no slider moves inside the call: the static
    // kernels are faster then.
    double* d_rows[2] = {nullptr, nullptr};
    size_t rows_cap = 0;
    std::vector<double> h_rows, h_q;

    int run_blocks_raw(Instance& I, const float* params, int nb, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        const int ns = I.n_streams, nd = (int)I.decls.size();
        if (nd != 15 || I.opt_force_generic || nb < 2) return 0;
        // does anything move?  (raw host values: equal floats quantise equally)
        bool moves = false;
        for (int b = 1; b < nb && !moves; ++b)
            moves = std::memcmp(params + (size_t)b * ns * nd, params, (size_t)ns * nd * sizeof(float)) != 0;
        if (!moves) return 0;
        constexpr int kSeg = 256;
        const size_t seg_rows = (size_t)std::min(nb, kSeg) * ns;
        if (seg_rows * kC8Coefs > rows_cap) {
            cudaFree(d_rows[0]); cudaFree(d_rows[1]); d_rows[0] = d_rows[1] = nullptr; rows_cap = 0;
            if (!cuda_ok(cudaMalloc(&d_rows[0], seg_rows * kC8Coefs * sizeof(double)), "cascade8 coefficient rows") ||
                !cuda_ok(cudaMalloc(&d_rows[1], seg_rows * kC8Coefs * sizeof(double)), "cascade8 coefficient rows")) return -1;
            rows_cap = seg_rows * kC8Coefs;
        }
        h_rows.resize(seg_rows * kC8Coefs);
        h_q.assign((size_t)ns * nd, 0.0);                 // quantised sliders of the previous block, per stream
        std::vector<uint8_t> have((size_t)ns, 0);
        const int64_t total_blocks = (n + I.block_len - 1) / I.block_len;
        int seg = 0;
        for (int b0 = 0; b0 < nb; b0 += kSeg, ++seg) {
            const int b1 = std::min(nb, b0 + kSeg);
            parallel_for(ns, 16, [&](int64_t sa, int64_t sb) {
                for (int64_t s = sa; s < sb; ++s) {
                    double* prev = &h_q[(size_t)s * nd];
                    for (int b = b0; b < b1; ++b) {
                        const float* p = params + ((size_t)b * ns + s) * nd;
                        double q[15];
                        bool ch = !have[s];
                        for (int j = 0; j < nd; ++j) { q[j] = quantise_slider(I.decls[j], p[j]); ch = ch || q[j] != prev[j]; }
                        double* row = &h_rows[((size_t)(b - b0) * ns + s) * kC8Coefs];
                        if (ch) {
                            c8_slider(q, I.srate, row);
                            std::memcpy(prev, q, sizeof(double) * nd);
                            have[s] = 1;
                        } else if (b > b0) {
                            std::memcpy(row, row - (size_t)ns * kC8Coefs, sizeof(double) * kC8Coefs);
                        } else {
                            std::memcpy(row, &h_coef[(size_t)s * kC8Coefs], sizeof(double) * kC8Coefs);    // first block of a later segment: the last row of the one before
                        }
                        if (b == b1 - 1) std::memcpy(&h_coef[(size_t)s * kC8Coefs], row, sizeof(double) * kC8Coefs);
                    }
                }
            });
            // h_rows is pageable: the copy is staged before cudaMemcpyAsync returns, so the buffer can be refilled; the device
            // buffers alternate, and the stream order keeps segment k's kernel ahead of the upload that reuses its buffer
            double* dr = d_rows[seg & 1];
            if (!cuda_ok(cudaMemcpyAsync(dr, h_rows.data(), (size_t)(b1 - b0) * ns * kC8Coefs * sizeof(double), cudaMemcpyHostToDevice, I.stream), "cascade8 rows upload"))
                return -1;
            ++I.launches;
            const int64_t f0 = (int64_t)b0 * I.block_len;
            const int64_t f1 = (b1 == nb) ? n : std::min<int64_t>(n, (int64_t)b1 * I.block_len);
            if (f1 <= f0) break;
            const bool timed = I.time_begin();
            if (!cascade8_auto_launch(in + f0, out + f0, tap ? tap + f0 : nullptr, stride, f1 - f0, ns, I.block_len, b1 - b0, dr, d_state, I.precision, I.stream))
                return -1;
            if (timed) I.time_end();
            ++I.launches;
            if (b1 >= total_blocks) break;
        }
        // the static kernels' coefficients and tables follow the last row, for whatever call comes next
        if (!cuda_ok(cudaMemcpyAsync(d_coef, h_coef.data(), h_coef.size() * sizeof(double), cudaMemcpyHostToDevice, I.stream), "cascade8 coef upload")) return -1;
        if (d_tables_x2 && !cascade8_x2_build_tables(d_coef, d_tables_x2, ns, I.stream)) return -1;
        if (!cascade8_build_tables(d_coef, d_tables, ns, I.precision, I.stream)) return -1;
        return 1;
    }

    bool run(Instance& I, int s0, int ns, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        // cascade8 has no @block section and its @sample does not read samplesblock, so a run of
        // host blocks with unchanged sliders is one contiguous span.
        C8Launch L;
        L.row_stride = stride; L.n_streams = ns; L.precision = I.precision; L.stream = I.stream;
        L.tables = (char*)d_tables + (size_t)s0 * cascade8_tables_bytes(I.precision);
        L.state = d_state + (size_t)s0 * kC8State;
        L.warps = I.opt_warps ? I.opt_warps : cascade8_pick_warps(ns, I.precision, I.device);
        const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0 && (stride % 4) == 0;
        const int64_t n_main = (aligned && !I.opt_force_generic) ? n - (n % kL_host()) : 0;
        if (n_main > 0) {
            L.in = in; L.out = out; L.tap = tap; L.n_frames = n_main; L.use_tma = true;
            const bool packed = !I.precision && !I.opt_scalar_f32;
            if (packed) {
                L.tables_x2 = (char*)d_tables_x2 + (size_t)s0 * cascade8_x2_tables_bytes();
                cascade8_x2_pick(ns, I.device, &L.warps, &L.nbuf);
                if (I.opt_warps) L.warps = I.opt_warps;
                if (I.opt_nbuf) L.nbuf = I.opt_nbuf;
            }
            const bool timed = I.time_begin();
            if (!(packed ? cascade8_x2_launch(L) : cascade8_launch(L))) return false;
            if (timed) I.time_end();
            ++I.launches;
        }
        if (n > n_main) {
            L.in = in + n_main; L.out = out + n_main; L.tap = tap ? tap + n_main : nullptr;
            L.n_frames = n - n_main; L.use_tma = false;
            L.warps = I.opt_warps ? I.opt_warps : cascade8_pick_warps(ns, I.precision, I.device);
            const bool timed = (n_main == 0) && I.time_begin();
            if (!cascade8_launch(L)) return false;
            if (timed) I.time_end();
            ++I.launches;
        }
        return true;
    }

    void state_io(Instance& I, StateIO& io) override { io.dev(d_state, (size_t)I.n_streams * kC8State * sizeof(double)); }
    bool after_state_load(Instance& I) override      // coefficients and scan tables are functions of the restored sliders
    {
        std::vector<int> all(I.n_streams);
        for (int s = 0; s < I.n_streams; ++s) all[s] = s;
        return sliders_changed(I, 0, I.n_streams, all);
    }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        static const char* names[14] = {"z11", "z12", "z21", "z22", "z31", "z32", "z41", "z42", "z51", "z52", "z61", "z62", "lp7", "lp8"};
        int idx = -1;
        for (int i = 0; i < 14; ++i) if (std::strcmp(name, names[i]) == 0) idx = i;
        if (idx < 0) { set_error(std::string("cascade8: unknown state variable ") + name); return false; }
        std::vector<double> h((size_t)I.n_streams * kC8State);
        if (!cuda_ok(cudaStreamSynchronize(I.stream), "sync")) return false;
        if (!cuda_ok(cudaMemcpy(h.data(), d_state, h.size() * sizeof(double), cudaMemcpyDeviceToHost), "state read")) return false;
        for (int s = 0; s < I.n_streams; ++s) out[s] = h[(size_t)s * kC8State + idx];
        return true;
    }

    static int kL_host() { return 32; }
};

}  // namespace

PluginImpl* make_cascade8() { return new Cascade8(); }
const char* cascade8_slider_text() { return kSliderText; }

}  // namespace za
