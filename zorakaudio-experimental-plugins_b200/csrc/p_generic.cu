// p_generic.cu -- host half of every translated DSP-JSFX leaf (csrc/gen/g_*.cu): device state sized at create
// (vars / sliders / mem[] heap / MT19937 columns, all stream-minor structure of arrays), prepareToPlay and
// processBlock as one kernel launch each (jsfx_rt.cuh).  The quantised automation rows of a call are uploaded
// as doubles; @init / @slider / @block / @sample all run on the device.
#include <cmath>
#include <cstring>
#include <string>

#include "za_host.h"
#include "gen/gen_list.h"

namespace za {

namespace jsfx {
#define ZA_X(IDENT) const GenDesc* desc_##IDENT();
ZA_GEN_PLUGINS(ZA_X)
#undef ZA_X
}  // namespace jsfx

namespace {

class GenericPlugin : public PluginImpl {
    const GenDesc* d_;
    void* d_state = nullptr; void* d_mem = nullptr; unsigned* d_mt = nullptr; int* d_err = nullptr;
    double* d_q = nullptr; size_t q_cap = 0;
    int* d_index0 = nullptr;
    double* d_tw = nullptr;         // FFT twiddles e^{-2 pi i k / 32768}, k < 16384 (host libm), for leaves that call the FFT hooks
    int* h_err = nullptr;           // pinned mirror of d_err, refreshed by an async copy behind every launch
    size_t rsize = 8;

    GenLaunch base(Instance& I) const
    {
        GenLaunch a{};
        a.mem_slots = I.opt_heap_slots > 0 ? std::min<int64_t>(I.opt_heap_slots, d_->mem_slots) : d_->mem_slots;   // (tests shrink the heap to provoke the overflow path)
        a.n_streams = I.n_streams; a.block_len = I.block_len; a.n_decl = (int)I.decls.size();
        a.decl_index0 = d_index0; a.state = d_state; a.mem = d_mem; a.mt = d_mt; a.err = d_err; a.srate = I.srate;
        a.fft_scratch = d_->fft_scratch; a.tw = d_tw; a.lanes = I.opt_lanes; a.unstaged = I.opt_unstaged;
        return a;
    }
    std::vector<double> q_last;     // rows already on the device: static parameters are not uploaded again (a pageable
                                    // cudaMemcpyAsync synchronises the stream first, which would serialise leaves that
                                    // render concurrently on their own streams)
    bool upload_q(Instance& I, const double* q, size_t count)
    {
        if (count == q_last.size() && count > 0 && std::memcmp(q, q_last.data(), count * sizeof(double)) == 0) return true;
        q_last.assign(q, q + count);
        if (count > q_cap) {
            cudaFree(d_q); d_q = nullptr; q_cap = 0;
            if (!cuda_ok(cudaMalloc(&d_q, count * sizeof(double)), "slider rows")) return false;
            q_cap = count;
        }
        return cuda_ok(cudaMemcpyAsync(d_q, q, count * sizeof(double), cudaMemcpyHostToDevice, I.stream), "slider rows upload");
    }
    bool check_err(Instance& I)
    {
        int e = 0;
        if (!cuda_ok(cudaMemcpyAsync(&e, d_err, sizeof e, cudaMemcpyDeviceToHost, I.stream), "err read") ||
            !cuda_ok(cudaStreamSynchronize(I.stream), "sync")) return false;
        if (e) { set_error(std::string(d_->key) + ": mem[] index beyond the heap this plugin was sized for"); return false; }
        return true;
    }

public:
    explicit GenericPlugin(const GenDesc* d) : d_(d) {}
    ~GenericPlugin() override { cudaFree(d_state); cudaFree(d_mem); cudaFree(d_mt); cudaFree(d_err); cudaFree(d_q); cudaFree(d_index0); cudaFree(d_tw); cudaFreeHost(h_err); }

    bool create(Instance& I) override
    {
        // float32 mode is offered only where it meets the north_star criterion (residual below -100 dBFS against the
        // reference render).  DDT and BedRock keep long phase / delay-position accumulators that lose it in float
        // (-50 / -53 dBFS measured on B200, tests/test_generated_gpu.py), so ZA_F32 is refused for them.
        if (I.precision == ZA_F32 && (!std::strcmp(d_->key, "DDT") || !std::strcmp(d_->key, "BedRock"))) {
            set_error(std::string(d_->key) + ": float32 mode does not meet the -100 dBFS criterion for this leaf; create it with ZA_F64");
            I.create_status = ZA_ERR_UNSUPPORTED;
            return false;
        }
        rsize = I.precision ? 8 : 4;
        const size_t ns = (size_t)I.n_streams;
        std::vector<int> idx;
        for (const auto& s : I.decls) idx.push_back(s.index0);
        bool same = (int)idx.size() == d_->n_decl;     // the kernels address sliders by compile-time slot
        for (int j = 0; same && j < d_->n_decl; ++j) same = idx[j] == d_->decl_index0[j];
        if (!same) { set_error(std::string(d_->key) + ": slider declarations differ from the ones the generated code was built for"); return false; }
        return cuda_ok(cudaMalloc(&d_state, (size_t)(d_->n_vars + 65) * ns * rsize), "state") &&
               cuda_ok(cudaMalloc(&d_mem, (size_t)heap_slots() * ns * rsize), "mem[] heap") /* + the spare slot out-of-range accesses land in + FFT scratch */ &&
               upload_twiddles() &&
               cuda_ok(cudaMalloc(&d_mt, 625 * ns * sizeof(unsigned)), "rand state") &&
               cuda_ok(cudaMalloc(&d_err, sizeof(int)), "err flag") &&
               cuda_ok(cudaHostAlloc(&h_err, sizeof(int), cudaHostAllocDefault), "err mirror") && ((*h_err = 0), true) &&
               cuda_ok(cudaMalloc(&d_index0, std::max<size_t>(idx.size(), 1) * sizeof(int)), "slider index") &&
               cuda_ok(cudaMemcpy(d_index0, idx.data(), idx.size() * sizeof(int), cudaMemcpyHostToDevice), "slider index upload");
    }

    size_t heap_slots() const { return (size_t)(d_->mem_slots + 1 + d_->fft_scratch); }
    bool upload_twiddles()
    {
        if (d_->fft_scratch <= 0) return true;
        std::vector<double> t(2 * 16384);
        for (int k = 0; k < 16384; ++k) { const double a = 6.283185307179586476925 * k / 32768.0; t[2 * k] = std::cos(a); t[2 * k + 1] = -std::sin(a); }
        return cuda_ok(cudaMalloc(&d_tw, t.size() * sizeof(double)), "FFT twiddles") &&
               cuda_ok(cudaMemcpy(d_tw, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice), "FFT twiddles upload");
    }
    bool heap_zeroed = false;
    bool reset(Instance& I) override   // resetStateStructOnly, then sliders, @init, alias re-apply, @slider
    {
        const size_t ns = (size_t)I.n_streams;
        // The heap is zero-filled when it is first allocated (jsfx_ensure_mem, JSFXJuceProcessor.cpp:1452-1491) and KEPT
        // across prepareToPlay (resetStateStructOnly, :8972-8980, preserves st.mem / st.memN): a second prepare on a
        // handle sees the delay lines the previous render left, exactly as the reference instance would.
        if (!heap_zeroed) {
            if (!cuda_ok(cudaMemsetAsync(d_mem, 0, heap_slots() * ns * rsize, I.stream), "heap reset")) return false;
            heap_zeroed = true;
        }
        if (!cuda_ok(cudaMemsetAsync(d_state, 0, (size_t)(d_->n_vars + 65) * ns * rsize, I.stream), "state reset") ||
            !cuda_ok(cudaMemsetAsync(d_mt, 0, 625 * ns * sizeof(unsigned), I.stream), "rand reset") ||
            !cuda_ok(cudaMemsetAsync(d_err, 0, sizeof(int), I.stream), "err reset"))
            return false;
        if (!upload_q(I, I.sliders.data(), I.sliders.size())) return false;
        GenLaunch a = base(I);
        a.q = d_q; a.n_rows = 1;
        if (!d_->prepare(a, I.precision, I.stream)) return false;
        ++I.launches;
        if (!check_err(I)) return false;
        *h_err = 0;
        return true;
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
        if (s0 != 0 || ns != I.n_streams) { set_error("translated plugins: stream sub-ranges are not supported"); return false; }
        if (!upload_q(I, q, (size_t)nbq * ns * I.decls.size())) return false;
        GenLaunch a = base(I);
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n; a.q = d_q; a.n_rows = nbq;
        const bool timed = I.time_begin();
        if (!d_->process(a, I.precision, I.stream)) return false;
        if (timed) I.time_end();
        ++I.launches;
        return cuda_ok(cudaMemcpyAsync(h_err, d_err, sizeof(int), cudaMemcpyDeviceToHost, I.stream), "err mirror copy");
    }

    // Whole-call automation without a host pass: the raw float rows are uploaded on a side stream and quantised by
    // quantise_rows_kernel (k_params.cu) with the host quantiser's own operations.
    RawRows rows_;
    int run_blocks_raw(Instance& I, const float* params, int nbq, const float* in, float* out, double* tap, int64_t n, int64_t stride) override
    {
        if (I.opt_force_generic) return 0;                      // "force_generic_io": the host quantiser path (tests compare the two)
        const double* dq = rows_.upload(I, params, (size_t)nbq * I.n_streams);
        if (!dq) return -1;
        q_last.clear();
        GenLaunch a = base(I);
        a.in = in; a.out = out; a.tap = tap; a.row_stride = stride; a.n_frames = n; a.q = dq; a.n_rows = nbq;
        const bool timed = I.time_begin();
        if (!d_->process(a, I.precision, I.stream)) return -1;
        if (timed) I.time_end();
        ++I.launches;
        rows_.consumed(I);
        return cuda_ok(cudaMemcpyAsync(h_err, d_err, sizeof(int), cudaMemcpyDeviceToHost, I.stream), "err mirror copy") ? 1 : -1;
    }

    bool deferred_error(Instance&) override
    {
        if (!h_err || *(volatile int*)h_err == 0) return false;
        set_error(std::string(d_->key) + ": mem[] index beyond the heap this plugin was sized for");
        return true;
    }

    void state_io(Instance& I, StateIO& io) override        // DSPJSFX_State: vars + sliders + lastSlidersValid, mem[], randMT + index
    {
        const size_t ns = (size_t)I.n_streams;
        io.dev(d_state, (size_t)(d_->n_vars + 65) * ns * rsize);
        io.dev(d_mem, heap_slots() * ns * rsize);
        io.dev(d_mt, 625 * ns * sizeof(unsigned));
        io.host(&heap_zeroed, sizeof heap_zeroed);
    }
    bool after_state_load(Instance&) override { q_last.clear(); return true; }

    bool read_state(Instance& I, const char* name, double* out) override
    {
        int slot = -1;
        for (int k = 0; d_->var_names[k]; ++k) if (!std::strcmp(d_->var_names[k], name)) { slot = k; break; }
        if (slot < 0) { set_error(std::string(d_->key) + ": unknown state variable " + name); return false; }
        if (!check_err(I)) return false;
        const size_t ns = (size_t)I.n_streams;
        if (I.precision) return cuda_ok(cudaMemcpy(out, (const double*)d_state + slot * ns, ns * 8, cudaMemcpyDeviceToHost), "state read");
        std::vector<float> t(ns);
        if (!cuda_ok(cudaMemcpy(t.data(), (const float*)d_state + slot * ns, ns * 4, cudaMemcpyDeviceToHost), "state read")) return false;
        for (size_t s = 0; s < ns; ++s) out[s] = (double)t[s];
        return true;
    }
};

}  // namespace

#define ZA_X(IDENT) static PluginImpl* make_gen_##IDENT() { return new GenericPlugin(jsfx::desc_##IDENT()); }
ZA_GEN_PLUGINS(ZA_X)
#undef ZA_X

void append_generated_plugins(std::vector<PluginDesc>& t)
{
    // A leaf that also has a hand-written kernel class keeps its machine translation under "<key>.translated" (cross-check)
    static std::vector<std::string> names;
    names.reserve(64);
#define ZA_X(IDENT) { const GenDesc* d = jsfx::desc_##IDENT(); bool taken = false; for (const auto& e : t) taken = taken || !std::strcmp(e.key, d->key); \
                      names.push_back(std::string(d->key) + (taken ? ".translated" : "")); \
                      t.push_back(PluginDesc{names.back().c_str(), d->n_in, d->n_out, 0, 0, d->slider_text, &make_gen_##IDENT}); }
    ZA_GEN_PLUGINS(ZA_X)
#undef ZA_X
}

}  // namespace za
