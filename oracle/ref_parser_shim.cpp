// oracle/ref_parser_shim.cpp -- TEST INFRASTRUCTURE.  Puts the reference's OWN slider-declaration parser behind a C entry
// point so that tests can pin the product's parser (csrc/za_params.cpp) against it, not against a restatement.
//
// Nothing of the reference is copied into this repository: oracle/build_ref.py::build_parser() extracts, at build time and
// from where it lies (/root/reference/src/JSFXJuceProcessor.cpp), the text of `struct JsfxSliderDecl`, `trimAscii`,
// `splitTopLevelCommas`, `parseFloat`, `parseJsfxStringDefaultToken` and `parseJsfxSliderDecls` (:526-556, :567, :633-704,
// :710-939) into a temporary oracle/_ref/ref_parser_extract.inc, compiles this file around it and deletes the extract.  The only thing the
// extract needs from JUCE is a string class, a string list and jlimit: the minimal stand-ins below (my code, ASCII only --
// slider lines are ASCII) implement just the members the extract calls, with JUCE's documented behaviour.
#include <algorithm>
#include <cctype>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <regex>
#include <string>
#include <vector>

namespace juce {

class String {
    std::string s_;
    static bool ws(char c) { return std::isspace((unsigned char)c) != 0; }

public:
    String() = default;
    String(const char* t) : s_(t ? t : "") {}
    String(const std::string& t) : s_(t) {}
    explicit String(int v) : s_(std::to_string(v)) {}
    static String fromUTF8(const char* t) { return String(t); }
    String trimStart() const { size_t a = 0; while (a < s_.size() && ws(s_[a])) ++a; return String(s_.substr(a)); }
    String trimEnd() const { size_t b = s_.size(); while (b > 0 && ws(s_[b - 1])) --b; return String(s_.substr(0, b)); }
    String trim() const { return trimStart().trimEnd(); }
    String toLowerCase() const { std::string t = s_; for (char& c : t) c = (char)std::tolower((unsigned char)c); return String(t); }
    String substring(int start) const { return String(start >= (int)s_.size() ? std::string() : s_.substr((size_t)std::max(start, 0))); }
    bool isEmpty() const { return s_.empty(); }
    bool startsWithChar(char c) const { return !s_.empty() && s_[0] == c; }
    void clear() { s_.clear(); }
    String& operator<<(const String& o) { s_ += o.s_; return *this; }
    String& operator<<(const char* o) { s_ += o; return *this; }
    bool operator==(const char* o) const { return s_ == o; }
    bool operator==(const String& o) const { return s_ == o.s_; }
    friend String operator+(const char* a, const String& b) { return String(std::string(a) + b.s_); }
    friend String operator+(const String& a, const String& b) { return String(a.s_ + b.s_); }
    const std::string& str() const { return s_; }
};

class StringArray {
    std::vector<String> v_;

public:
    // addTokens(text, breakCharacters, quoteCharacters): split at any break character (no quote characters are used by the caller)
    int addTokens(const String& text, const char* breaks, const char*)
    {
        const std::string& t = text.str();
        int n = 0;
        size_t a = 0;
        if (t.empty()) return 0;
        for (;;) {
            const size_t b = t.find_first_of(breaks, a);
            v_.push_back(String(t.substr(a, b == std::string::npos ? std::string::npos : b - a)));
            ++n;
            if (b == std::string::npos) break;
            a = b + 1;
        }
        return n;
    }
    void trim() { for (String& s : v_) s = s.trim(); }
    void removeEmptyStrings() { v_.erase(std::remove_if(v_.begin(), v_.end(), [](const String& s) { return s.isEmpty(); }), v_.end()); }
    int size() const { return (int)v_.size(); }
    void clear() { v_.clear(); }
    const String& operator[](int i) const { return v_[(size_t)i]; }
};

template <typename T> T jlimit(T lo, T hi, T v) { return v < lo ? lo : (hi < v ? hi : v); }

}  // namespace juce

#include "_ref/ref_parser_extract.inc"

extern "C" {

struct RefSliderDecl {
    int32_t index0, is_choice, n_choices, is_string, hidden, shape;
    float def, min, max, step, shape_mod;
    char var[64], name[128];
};

__attribute__((visibility("default"))) int ref_parse_sliders(const char* text, RefSliderDecl* out, int cap)
{
    const std::vector<JsfxSliderDecl> v = parseJsfxSliderDecls(text);
    int n = 0;
    for (const JsfxSliderDecl& d : v) {
        if (n >= cap) break;
        RefSliderDecl& o = out[n++];
        std::memset(&o, 0, sizeof o);
        o.index0 = d.index0; o.is_choice = d.isChoice ? 1 : 0; o.n_choices = d.choices.size(); o.is_string = d.isString ? 1 : 0;
        o.hidden = d.hidden ? 1 : 0; o.shape = (int)d.shape;
        o.def = d.def; o.min = d.min; o.max = d.max; o.step = d.step; o.shape_mod = d.shapeModifier;
        std::strncpy(o.var, d.varName.str().c_str(), sizeof o.var - 1);
        std::strncpy(o.name, d.name.str().c_str(), sizeof o.name - 1);
    }
    return (int)v.size();
}

}  // extern "C"
