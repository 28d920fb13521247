// za_params.cpp -- the host parameter path of the reference, as C++ without JUCE:
//   slider declaration parser      src/JSFXJuceProcessor.cpp:710-939  (parseJsfxSliderDecls)
//   parameter -> slider quantiser  src/JSFXJuceProcessor.cpp:5556-5597 (hostParameterToJsfxSliderValue)
// Same grammar, same float storage of min/max/step (:531-534, 2492-2502), same llround grid.
#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdlib>
#include <regex>

#include "za_host.h"

namespace za {

static std::string trim(std::string s)
{
    while (!s.empty() && std::isspace((unsigned char)s.front())) s.erase(s.begin());
    while (!s.empty() && std::isspace((unsigned char)s.back())) s.pop_back();
    return s;
}

static bool parse_float(const std::string& s, float& out)   // strtod prefix parse, then narrowed (:664-672)
{
    char* end = nullptr;
    const double v = std::strtod(s.c_str(), &end);
    if (end == s.c_str()) return false;
    out = (float)v;
    return true;
}

static std::vector<std::string> split_top_level_commas(const std::string& s)   // `1{A,B,C}` keeps its commas
{
    std::vector<std::string> parts;
    std::string cur;
    int depth = 0;
    for (char c : s) {
        if (c == '{') ++depth;
        else if (c == '}' && depth > 0) --depth;
        if (c == ',' && depth == 0) { parts.push_back(trim(cur)); cur.clear(); }
        else cur.push_back(c);
    }
    parts.push_back(trim(cur));
    return parts;
}

static std::string lower(std::string s)
{
    for (auto& c : s) c = (char)std::tolower((unsigned char)c);
    return s;
}

std::vector<SliderDecl> parse_slider_decls(const char* text_c)
{
    std::vector<SliderDecl> out;
    if (!text_c) return out;
    static const std::regex re(R"(^\s*slider\s*([0-9]{1,2})\s*:\s*([^<\r\n;]+)\s*(?:<\s*([^>]+)\s*>)?\s*(.*)$)",
                               std::regex::ECMAScript);
    const std::string text(text_c);
    size_t start = 0;
    while (start < text.size()) {
        size_t end = text.find_first_of("\r\n", start);
        if (end == std::string::npos) end = text.size();
        const std::string line = text.substr(start, end - start);
        size_t next = end;
        while (next < text.size() && (text[next] == '\r' || text[next] == '\n')) ++next;
        start = next;

        std::smatch m;
        if (!std::regex_match(line, m, re)) continue;
        const int n = std::atoi(m[1].str().c_str());
        if (n < 1 || n > 64) continue;

        SliderDecl d;
        d.index0 = n - 1;
        std::string tok = trim(m[2].str()), var, deftok = tok;
        const auto eq = tok.rfind('=');
        if (eq != std::string::npos) { var = trim(tok.substr(0, eq)); deftok = trim(tok.substr(eq + 1)); }
        float def = 0.0f;
        if (!parse_float(deftok, def)) def = 0.0f;
        d.def = def;
        d.alias = var;

        bool is_string = false;
        if (m[3].matched) {
            const std::string kind = lower(trim(m[3].str()));
            if (kind == "string" || kind == "str" || kind == "text") is_string = true;
        }
        if (!is_string && !var.empty() && var[0] == '#') is_string = true;
        if (is_string) continue;      // string sliders are host-only (file/text slots)

        if (m[3].matched) {
            const auto parts = split_top_level_commas(m[3].str());
            float vmin = 0.0f, vmax = 1.0f, vstep = 0.001f;
            if (parts.size() >= 2) {
                if (!parse_float(parts[0], vmin)) vmin = 0.0f;
                if (!parse_float(parts[1], vmax)) vmax = 1.0f;
            }
            if (parts.size() >= 3) {
                std::string st = parts[2];
                const auto bp = st.find('{');
                if (bp != std::string::npos) {
                    const auto cp = st.find('}', bp + 1);
                    if (cp != std::string::npos) {
                        const auto labels = split_top_level_commas(st.substr(bp + 1, cp - bp - 1));
                        for (const auto& l : labels)
                            if (!trim(l).empty()) { d.is_choice = true; break; }
                    }
                    st = trim(st.substr(0, bp));
                }
                const auto colon = st.find(':');          // ":log", ":sqr=2": host-GUI curve only
                if (colon != std::string::npos) st = trim(st.substr(0, colon));
                if (st.empty()) vstep = 1.0f;
                else if (!parse_float(st, vstep)) vstep = 1.0f;
            }
            if (vmax < vmin) std::swap(vmax, vmin);
            d.min = vmin; d.max = vmax;
            d.step = vstep > 0.0f ? vstep : 0.001f;
            d.def = std::min(std::max(d.def, d.min), d.max);
        }
        out.push_back(d);
    }
    std::stable_sort(out.begin(), out.end(), [](const SliderDecl& a, const SliderDecl& b) { return a.index0 < b.index0; });
    out.erase(std::unique(out.begin(), out.end(), [](const SliderDecl& a, const SliderDecl& b) { return a.index0 == b.index0; }),
              out.end());
    return out;
}

double quantise_slider(const SliderDecl& d, float host_value)
{
    double v = (double)host_value;
    const double mn = (double)d.min, mx = (double)d.max;
    const float stepf = d.step > 0.0f ? d.step : 1.0f;                 // SliderParamInfo.step (:2784)
    if (d.is_choice) {
        const int64_t idx = (int64_t)std::llround(v);
        v = mn + (double)idx * (double)stepf;
    }
    v = v < mn ? mn : (v > mx ? mx : v);
    if (!d.is_choice) {
        const double step = (double)stepf;
        if (step > 0.0) {
            const double q = (double)std::llround((v - mn) / step);
            v = mn + q * step;
            v = v < mn ? mn : (v > mx ? mx : v);
        }
    }
    return v;
}

}  // namespace za
