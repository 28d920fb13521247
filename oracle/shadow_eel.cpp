/*
 * oracle/shadow_eel.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product): the SECONDARY oracle.
 *
 * A JUCE-free restatement of the reference's "shadow EEL2" runtime (src/JSFXCorrectnessCheck.h:111-720 `ShadowVm`,
 * :722-1613 `Runtime`; section extraction and the portable-EEL shim of src/YSFXGfxInterpreter.h:117-432, no-op host
 * stubs :962-1035): the plugin's .jsfx source is split into @init / @slider / @block / @sample, each section compiled by
 * the reference's own WDL EEL2 interpreter (/root/reference/src/WDL/eel2, EEL_TARGET_PORTABLE, compiled where it lies by
 * oracle/build_ref.py into oracle/_ref/libshadow_eel.so) and driven exactly as the correctness build drives it
 * (src/JSFXJuceProcessor.cpp:3536-3658): sliders pushed and aliases mirrored, @slider on change, samplesblock, @block,
 * per frame spl[] <- float input, @sample, spl[] read back as double.
 *
 * Purpose (BASELINE.json configs[0]): a differential check that oracle/ref_shim.c drives the
 * reference's compiled DSP the way the host does -- compiled path vs shadow at the tolerances the reference ships
 * (1e-5 audio / 1e-8 scalar, JSFXCorrectnessCheck.h:34-35).  The reference's own ShadowVm cannot be used verbatim: it
 * inherits jsfx_gfx::GfxVm, which needs juce::Colour / Font / String (YSFXGfxInterpreter.h:507-598).
 * Known semantic gaps against the compiled path: EEL2 compares `==` / `!=` and truthiness with a 1e-5
 * epsilon, the compiled path exactly; EEL2's rand() state is global, the compiled one per instance.
 */
#define EELSCRIPT_NO_NET
#define EELSCRIPT_NO_LICE
#define EELSCRIPT_NO_FILE
#define EELSCRIPT_NO_STDIO
#define EELSCRIPT_NO_MDCT
#define EELSCRIPT_NO_PREPROC
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <xmmintrin.h>
#include <pmmintrin.h>

#include <mutex>
#include <string>

#include "WDL/eel2/ns-eel.h"
#include "WDL/eel2/eelscript.h"

extern "C" void NSEEL_HOSTSTUB_EnterMutex() {}
extern "C" void NSEEL_HOSTSTUB_LeaveMutex() {}

#define ZSH_API extern "C" __attribute__((visibility("default")))

namespace {

struct ZrefSliders {             /* same layout as oracle/ref_shim.c / za_oracle.SliderLayout */
    int32_t n;
    int32_t index0[64];
    int32_t alias[64];           /* unused here: the shadow binds aliases by NAME */
    float vmin[64], vmax[64], vstep[64];
    int32_t is_choice[64];
};

double quantise(const ZrefSliders* sl, int j, float host_value)   /* src/JSFXJuceProcessor.cpp:5556-5597 */
{
    double v = (double)host_value;
    const double mn = (double)sl->vmin[j], mx = (double)sl->vmax[j];
    if (sl->is_choice[j]) v = mn + (double)(int64_t)llround(v) * (double)sl->vstep[j];
    v = v < mn ? mn : (v > mx ? mx : v);
    if (!sl->is_choice[j]) {
        const double step = (double)(sl->vstep[j] > 0.0f ? sl->vstep[j] : 0.0f);
        if (step > 0.0) {
            v = mn + (double)llround((v - mn) / step) * step;
            v = v < mn ? mn : (v > mx ? mx : v);
        }
    }
    return v;
}

/* ---- section extraction (YSFXGfxInterpreter.h:117-218) ------------------------------------------------------ */
struct Sections { std::string init, slider, block, sample; };

bool starts_with_section(const std::string& s, const char* sec)
{
    const size_t n = strlen(sec);
    if (s.size() < n + 1 || s[0] != '@') return false;
    for (size_t i = 0; i < n; ++i)
        if (tolower((unsigned char)s[i + 1]) != tolower((unsigned char)sec[i])) return false;
    return true;
}

Sections extract_sections(const char* text)
{
    Sections out;
    std::string* cur = nullptr;
    const char* p = text;
    while (*p) {
        const char* start = p;
        while (*p && *p != '\n') ++p;
        std::string line(start, p);
        if (*p == '\n') ++p;
        const size_t f = line.find_first_not_of(" \t\r");
        const std::string lt = f == std::string::npos ? std::string() : line.substr(f);
        if (!lt.empty() && lt[0] == '@') {
            if (starts_with_section(lt, "init")) cur = &out.init;
            else if (starts_with_section(lt, "slider")) cur = &out.slider;
            else if (starts_with_section(lt, "block")) cur = &out.block;
            else if (starts_with_section(lt, "sample")) cur = &out.sample;
            else cur = nullptr;                     /* @gfx, @serialize, unknown: not part of the DSP path */
            continue;
        }
        if (cur) { cur->append(line); cur->push_back('\n'); }
    }
    return out;
}

/* ---- portable-EEL shim (YSFXGfxInterpreter.h:220-432): `slider(i) = v;` / `spl(i) = v;` -> `slider(i, v);`, and spl() ->
 * dsp_spl() (JSFXCorrectnessCheck.h:457-462).  Text transform outside comments and strings. ------------------------------ */
bool ident_char(char c) { return isalnum((unsigned char)c) || c == '_'; }

std::string preprocess(const std::string& in)
{
    std::string out;
    out.reserve(in.size() + 64);
    bool line_c = false, block_c = false, str = false;
    char q = 0;
    for (size_t i = 0; i < in.size();) {
        const char c = in[i], nx = i + 1 < in.size() ? in[i + 1] : '\0';
        if (line_c) { out.push_back(c); if (c == '\n') line_c = false; ++i; continue; }
        if (block_c) { out.push_back(c); if (c == '*' && nx == '/') { out.push_back(nx); i += 2; block_c = false; } else ++i; continue; }
        if (str) { out.push_back(c); if (c == '\\' && i + 1 < in.size()) { out.push_back(in[i + 1]); i += 2; continue; } if (c == q) str = false; ++i; continue; }
        if (c == '/' && nx == '/') { out += "//"; i += 2; line_c = true; continue; }
        if (c == '/' && nx == '*') { out += "/*"; i += 2; block_c = true; continue; }
        if (c == '"' || c == '\'') { str = true; q = c; out.push_back(c); ++i; continue; }
        bool done = false;
        for (const char* name : {"slider", "spl"}) {
            const size_t nl = strlen(name);
            if (in.compare(i, nl, name) != 0 || (i > 0 && ident_char(in[i - 1])) || i + nl >= in.size() || in[i + nl] != '(') continue;
            size_t p = i + nl + 1;
            int depth = 1;
            while (p < in.size() && depth > 0) { if (in[p] == '(') ++depth; else if (in[p] == ')') --depth; ++p; }
            if (depth != 0) break;
            const std::string fn = strcmp(name, "spl") == 0 ? "dsp_spl" : "slider";
            const std::string arg = in.substr(i + nl + 1, p - 1 - (i + nl + 1));
            size_t a = p;
            while (a < in.size() && isspace((unsigned char)in[a])) ++a;
            if (a < in.size() && in[a] == '=' && !(a + 1 < in.size() && in[a + 1] == '=')) {
                size_t r = a + 1;
                int par = 0;
                while (r < in.size() && !(in[r] == ';' && par == 0) && !(in[r] == ')' && par == 0)) { if (in[r] == '(') ++par; else if (in[r] == ')') --par; ++r; }
                out += fn + "(" + preprocess(arg) + ", " + preprocess(in.substr(a + 1, r - (a + 1))) + ")";
                i = r;
            } else {
                out += fn + "(" + preprocess(arg) + ")";
                i = p;
            }
            done = true;
            break;
        }
        if (done) continue;
        out.push_back(c);
        ++i;
    }
    return out;
}

struct Shadow;
EEL_F NSEEL_CGEN_CALL fn_noop(void*, INT_PTR, EEL_F**) { return 0.0; }
EEL_F NSEEL_CGEN_CALL fn_slider(void* opaque, INT_PTR np, EEL_F** parms);
EEL_F NSEEL_CGEN_CALL fn_spl(void* opaque, INT_PTR np, EEL_F** parms);
EEL_F NSEEL_CGEN_CALL fn_memset(void* opaque, INT_PTR np, EEL_F** parms);

struct Shadow : public eelScriptInst {
    ZrefSliders sl{};
    std::string alias_names[64];
    Sections sec;
    std::string error;
    EEL_F* slider_ptr[64] = {};
    EEL_F* spl_ptr[64] = {};
    EEL_F* alias_ptr[64] = {};
    EEL_F* srate_var = nullptr;
    EEL_F* samplesblock_var = nullptr;
    NSEEL_CODEHANDLE c_init = nullptr, c_slider = nullptr, c_block = nullptr, c_sample = nullptr;
    double last[64] = {};
    bool last_valid = false;
    unsigned ram = 0;

    void ensure_ram(int64_t needed)
    {
        if (needed > (int64_t)ram) { NSEEL_VM_setramsize(m_vm, (unsigned)needed); ram = (unsigned)needed; }
    }
    bool compile()
    {
        for (int i = 0; i < 64; ++i) {
            char nm[32];
            snprintf(nm, sizeof nm, "slider%d", i + 1); slider_ptr[i] = NSEEL_VM_regvar(m_vm, nm);
            snprintf(nm, sizeof nm, "spl%d", i); spl_ptr[i] = NSEEL_VM_regvar(m_vm, nm);
            if (spl_ptr[i]) *spl_ptr[i] = 0.0;
        }
        srate_var = NSEEL_VM_regvar(m_vm, "srate");
        samplesblock_var = NSEEL_VM_regvar(m_vm, "samplesblock");
        for (int j = 0; j < sl.n; ++j)
            if (!alias_names[j].empty()) alias_ptr[sl.index0[j]] = NSEEL_VM_regvar(m_vm, alias_names[j].c_str());
        NSEEL_CODEHANDLE* dst[4] = {&c_init, &c_slider, &c_block, &c_sample};
        const std::string* src[4] = {&sec.init, &sec.slider, &sec.block, &sec.sample};
        static const char* nm[4] = {"@init", "@slider", "@block", "@sample"};
        for (int k = 0; k < 4; ++k) {
            const std::string pre = preprocess(src[k]->empty() ? std::string("0;") : *src[k]);
            const char* err = nullptr;
            *dst[k] = compile_code(pre.c_str(), &err);
            if (!*dst[k]) { error = std::string(nm[k]) + ": " + (err ? err : "EEL compile error"); return false; }
        }
        return true;
    }
    void sync_sliders(const double* v)              /* syncHostSlidersAndAliases (JSFXCorrectnessCheck.h:194-210) */
    {
        for (int i = 0; i < 64; ++i) if (slider_ptr[i]) *slider_ptr[i] = v[i];
        sync_aliases();
    }
    void sync_aliases()
    {
        for (int i = 0; i < 64; ++i) if (alias_ptr[i] && slider_ptr[i]) *alias_ptr[i] = *slider_ptr[i];
    }
    void run(NSEEL_CODEHANDLE h) { if (h) NSEEL_code_execute(h); }
};

EEL_F NSEEL_CGEN_CALL fn_slider(void* opaque, INT_PTR np, EEL_F** parms)
{
    Shadow* s = (Shadow*)opaque;
    if (!s || np < 1) return 0.0;
    const int i = (int)(int64_t)((double)*parms[0] + 1.0e-5) - 1;
    if (i < 0 || i >= 64 || !s->slider_ptr[i]) return np >= 2 ? *parms[1] : 0.0;
    if (np >= 2) *s->slider_ptr[i] = *parms[1];
    return *s->slider_ptr[i];
}
EEL_F NSEEL_CGEN_CALL fn_spl(void* opaque, INT_PTR np, EEL_F** parms)        /* eel_dsp_spl, JSFXCorrectnessCheck.h:505-527 */
{
    Shadow* s = (Shadow*)opaque;
    if (!s || np < 1) return 0.0;
    const int i = (int)(int64_t)((double)*parms[0] + 1.0e-5);
    if (i < 0 || i >= 64 || !s->spl_ptr[i]) return np >= 2 ? *parms[1] : 0.0;
    if (np >= 2) *s->spl_ptr[i] = *parms[1];
    return *s->spl_ptr[i];
}
EEL_F NSEEL_CGEN_CALL fn_memset(void* opaque, INT_PTR np, EEL_F** parms)     /* eel_memset, JSFXCorrectnessCheck.h:663-707 */
{
    Shadow* s = (Shadow*)opaque;
    if (!s || np < 3) return 0.0;
    int64_t dst = (int64_t)((double)*parms[0] + 1.0e-5), len = (int64_t)((double)*parms[2] + 1.0e-5);
    if (dst < 0) dst = 0;
    if (len <= 0) return (EEL_F)dst;
    s->ensure_ram(dst + len);
    const EEL_F value = *parms[1];
    int64_t pos = dst, left = len;
    while (left > 0) {
        int valid = 0;
        EEL_F* p = NSEEL_VM_getramptr(s->m_vm, (unsigned)pos, &valid);
        if (!p || valid <= 0) break;
        const int64_t n = left < valid ? left : valid;
        for (int64_t k = 0; k < n; ++k) p[k] = value;
        pos += n; left -= n;
    }
    return (EEL_F)dst;
}

void register_once()
{
    static std::once_flag once;
    std::call_once(once, [] {
        eelScriptInst::init();
        /* host-only builtins a DSP section may mention (scripts define @gfx helpers inside @init): evaluate to 0
         * (YSFXGfxInterpreter.h:962-1035 registers the same names) */
        static const char* const noop[] = {"gfx_set", "gfx_rect", "gfx_rectto", "gfx_circle", "gfx_roundrect", "gfx_arc", "gfx_triangle",
            "gfx_line", "gfx_lineto", "gfx_drawstr", "gfx_printf", "gfx_setfont", "gfx_measurestr", "gfx_getchar", "gfx_showmenu",
            "gfx_showmenu_nb_open", "gfx_showmenu_nb_poll", "gfx_showmenu_nb_cancel", "gfx_blit", "gfx_gradrect", "gfx_setpixel", "gfx_getpixel",
            "gfx_drawnumber", "gfx_drawchar", "gfx_blurto", "gfx_setimgdim", "gfx_getimgdim", "gfx_loadimg", "gfx_muladdrect", "gfx_deltablit",
            "gfx_transformblit", "gfx_blitext", "gfx_clienttoscreen", "gfx_screentoclient",
            "sliderchange", "slider_automate", "slider_show", "slider_next_chg", "track_name", "host_track_name", "track_name_available",
            "host_track_name_available", "track_name_seq", "host_track_name_seq", "freembuf", "midisend", "midirecv", "midisyx",
            "file_open", "file_close", "file_rewind", "file_seek", "file_avail", "file_text", "file_riff", "file_var", "file_mem"};
        for (const char* n : noop) NSEEL_addfunc_varparm_ex(n, 0, 0, NSEEL_PProc_THIS, &fn_noop, nullptr);
        NSEEL_addfunc_varparm_ex("slider", 1, 0, NSEEL_PProc_THIS, &fn_slider, nullptr);
        NSEEL_addfunc_varparm_ex("dsp_spl", 1, 0, NSEEL_PProc_THIS, &fn_spl, nullptr);
        NSEEL_addfunc_varparm_ex("memset", 3, 1, NSEEL_PProc_THIS, &fn_memset, nullptr);
    });
}

struct Handle {
    std::string text;
    ZrefSliders sl{};
    std::string alias_names[64];
    Shadow* vm = nullptr;
    std::string error;
    double srate = 48000.0;
};

bool fresh_vm(Handle* h)          /* Runtime::resetAndPrime builds a new ShadowVm (JSFXCorrectnessCheck.h:732-750) */
{
    delete h->vm;
    h->vm = new Shadow();
    h->vm->sl = h->sl;
    for (int j = 0; j < 64; ++j) h->vm->alias_names[j] = h->alias_names[j];
    h->vm->sec = extract_sections(h->text.c_str());
    if (!h->vm->m_vm || !h->vm->compile()) { h->error = h->vm->error.empty() ? "EEL VM not initialised" : h->vm->error; return false; }
    return true;
}

void ftz_on()
{
    _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);
    _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
}

}  // namespace

ZSH_API void* zsh_new(const char* jsfx_text, const void* sliders, const char* const* alias_names)
{
    register_once();
    Handle* h = new Handle();
    h->text = jsfx_text ? jsfx_text : "";
    if (sliders) h->sl = *(const ZrefSliders*)sliders;
    for (int j = 0; j < h->sl.n && alias_names; ++j) if (alias_names[j]) h->alias_names[j] = alias_names[j];
    fresh_vm(h);
    return h;
}
ZSH_API const char* zsh_error(void* hv) { return ((Handle*)hv)->error.c_str(); }
ZSH_API void zsh_free(void* hv) { Handle* h = (Handle*)hv; if (h) { delete h->vm; delete h; } }

ZSH_API int zsh_prepare(void* hv, double srate, const float* host_values)
{
    Handle* h = (Handle*)hv;
    ftz_on();
    if (!fresh_vm(h)) return -1;
    Shadow* s = h->vm;
    h->srate = srate;
    s->ensure_ram(8 << 20);          /* the reference grows the shadow's RAM with the compiled path's memN (beginBlock, :258); 8 Mi slots = the
                                        compiled path's default __memtop, so no access a leaf makes is ever out of range */
    double v[64] = {};
    for (int j = 0; j < s->sl.n; ++j) { v[s->sl.index0[j]] = quantise(&s->sl, j, host_values[j]); s->last[s->sl.index0[j]] = v[s->sl.index0[j]]; }
    s->last_valid = true;
    s->sync_sliders(v);
    if (s->srate_var) *s->srate_var = srate;
    if (s->samplesblock_var) *s->samplesblock_var = 0.0;
    s->run(s->c_init);
    s->sync_aliases();
    s->run(s->c_slider);
    return 0;
}

/* One host block: src/JSFXJuceProcessor.cpp:3536-3556 (sliders, @slider on change), :3593-3658 (beginBlock, @block, frames). */
ZSH_API int zsh_process_block(void* hv, const float* host_values, const float* const* in, float* const* out, double* const* out_d,
                              int n_in, int n_out, int n)
{
    Handle* h = (Handle*)hv;
    Shadow* s = h->vm;
    if (!s) return -1;
    ftz_on();
    int changed = 0;
    if (host_values) {
        double v[64];
        for (int i = 0; i < 64; ++i) v[i] = s->slider_ptr[i] ? *s->slider_ptr[i] : 0.0;
        for (int j = 0; j < s->sl.n; ++j) {
            const int i = s->sl.index0[j];
            const double qv = quantise(&s->sl, j, host_values[j]);
            if (!s->last_valid || qv != s->last[i]) { changed = 1; s->last[i] = qv; }
            v[i] = qv;
        }
        s->last_valid = true;
        if (changed) { s->sync_sliders(v); s->run(s->c_slider); }
    }
    if (s->srate_var) *s->srate_var = h->srate;
    if (s->samplesblock_var) *s->samplesblock_var = (EEL_F)n;
    s->run(s->c_block);
    const int nch = n_in > n_out ? n_in : n_out;
    for (int i = 0; i < n; ++i) {
        for (int c = 0; c < nch && c < 64; ++c) if (s->spl_ptr[c]) *s->spl_ptr[c] = c < n_in ? (EEL_F)in[c][i] : 0.0;
        s->run(s->c_sample);
        for (int c = 0; c < n_out && c < 64; ++c) {
            const double y = s->spl_ptr[c] ? (double)*s->spl_ptr[c] : 0.0;
            if (out_d) out_d[c][i] = y;
            if (out) out[c][i] = (float)y;
        }
    }
    return changed;
}

ZSH_API int zsh_var(void* hv, const char* name, double* out)
{
    Handle* h = (Handle*)hv;
    if (!h->vm || !name || !out) return -1;
    EEL_F* p = NSEEL_VM_regvar(h->vm->m_vm, name);
    if (!p) return -1;
    *out = (double)*p;
    return 0;
}
