// k_params.cu -- device side of the parameter path shared by the plugins that take per-block automation as quantised
// slider rows: the raw host floats of a call [rows][n_sliders] are uploaded on a side stream (overlapping the previous
// call's render kernel) and clamped + step-quantised to doubles by a kernel with exactly the operations of
// hostParameterToJsfxSliderValue (src/JSFXJuceProcessor.cpp:5556-5597; host twin: za_params.cpp::quantise_slider), so no
// host core touches the [blocks][streams][sliders] trajectory (1.8e8 quantiser calls per cfg3-shaped step for an 8-slider
// leaf when done on the host).
#include "za_host.h"

namespace za {

namespace {

struct QDecl { double mn, mx, step; int is_choice; };
struct QArgs { const float* raw; double* q; size_t n; int nd; QDecl d[64]; };

__global__ void __launch_bounds__(256) quantise_rows_kernel(const QArgs a)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (size_t)gridDim.x * blockDim.x) {
        const QDecl d = a.d[i % a.nd];
        double v = (double)a.raw[i];
        if (d.is_choice) v = __dadd_rn(d.mn, __dmul_rn((double)llround(v), d.step));
        v = v < d.mn ? d.mn : (v > d.mx ? d.mx : v);
        if (!d.is_choice && d.step > 0.0) {
            const double k = (double)llround(__ddiv_rn(__dsub_rn(v, d.mn), d.step));
            v = __dadd_rn(d.mn, __dmul_rn(k, d.step));
            v = v < d.mn ? d.mn : (v > d.mx ? d.mx : v);
        }
        a.q[i] = v;
    }
}

}  // namespace

RawRows::~RawRows()
{
    cudaFree(d_raw); cudaFree(d_q);
    if (ev_ready) cudaEventDestroy(ev_ready);
    if (ev_consumed) cudaEventDestroy(ev_consumed);
    if (up) cudaStreamDestroy(up);
}

const double* RawRows::upload(Instance& I, const float* params, size_t rows)
{
    const int nd = (int)I.decls.size();
    if (nd <= 0 || nd > 64) { set_error("parameter rows: 1 .. 64 sliders expected"); return nullptr; }
    const size_t n = rows * nd;
    if (n > cap) {
        cudaFree(d_raw); cudaFree(d_q); d_raw = nullptr; d_q = nullptr; cap = 0;
        if (!cuda_ok(cudaMalloc(&d_raw, n * sizeof(float)), "raw parameter rows") || !cuda_ok(cudaMalloc(&d_q, n * sizeof(double)), "quantised parameter rows"))
            return nullptr;
        cap = n;
    }
    if (!up) {
        if (!cuda_ok(cudaStreamCreateWithFlags(&up, cudaStreamNonBlocking), "parameter upload stream")) return nullptr;
        cudaEventCreateWithFlags(&ev_ready, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ev_consumed, cudaEventDisableTiming);
    } else {
        cudaStreamWaitEvent(up, ev_consumed, 0);              // the previous call's render kernel has read d_q
    }
    if (!cuda_ok(cudaMemcpyAsync(d_raw, params, n * sizeof(float), cudaMemcpyHostToDevice, up), "parameter rows upload")) return nullptr;
    QArgs a{};
    a.raw = d_raw; a.q = d_q; a.n = n; a.nd = nd;
    for (int j = 0; j < nd; ++j) {
        const SliderDecl& d = I.decls[j];
        a.d[j] = QDecl{(double)d.min, (double)d.max, (double)(d.step > 0.0f ? d.step : 1.0f), d.is_choice ? 1 : 0};
    }
    quantise_rows_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 8), 256, 0, up>>>(a);
    if (!cuda_ok(cudaGetLastError(), "quantise_rows_kernel")) return nullptr;
    ++I.launches;
    cudaEventRecord(ev_ready, up);
    cudaStreamWaitEvent(I.stream, ev_ready, 0);
    return d_q;
}

const double* RawRows::upload_quantised(Instance& I, const double* q, size_t count)
{
    if (count == last.size() && count > 0 && std::memcmp(q, last.data(), count * sizeof(double)) == 0) return d_q;   // static rows already there
    if (count > cap) {
        cudaFree(d_raw); cudaFree(d_q); d_raw = nullptr; d_q = nullptr; cap = 0;
        if (!cuda_ok(cudaMalloc(&d_q, count * sizeof(double)), "quantised parameter rows")) return nullptr;
        cap = count;
    }
    last.assign(q, q + count);
    if (!cuda_ok(cudaMemcpyAsync(d_q, q, count * sizeof(double), cudaMemcpyHostToDevice, I.stream), "quantised rows upload")) return nullptr;
    return d_q;
}

void RawRows::consumed(Instance& I)
{
    if (ev_consumed) cudaEventRecord(ev_consumed, I.stream);
}

}  // namespace za
