// za_bounce.cpp -- offline bounce driver over the C ABI of include/za_gpu.h.
//
// Every input WAV file is one stream (one plugin instance / bounce job).  The files are rendered together through
// ONE za_gpu handle in time chunks (state is carried by the library between chunks), the way an offline host would
// have looped `prepareToPlay(sr, block); for each block: processBlock(...)` over every instance
// (src/JSFXJuceProcessor.cpp:3239, :3435), and each result is written as 24-bit PCM WAV next to a report.json -- the
// format of the reference's own exporter (src/JSFXCorrectnessCheck.h:1177-1255).
//
// Host C++ only: no CUDA, no torch; links libza_gpu.so.  There is no CPU rendering path -- without a CUDA device the
// library's create call fails and so does this program.
#include <sys/stat.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <map>

#include "../../include/za_gpu.h"
#include "za_wav.h"

namespace {

int usage(const char* argv0)
{
    std::fprintf(stderr,
                 "usage: %s (--plugin KEY | --plugin-json plugin.json) [--precision f32|f64] [--block N] [--chunk-seconds S]\n"
                 "          [--tail-seconds S] [--preset STATE.xml] [--set SLIDER=VALUE ...] [--save-preset OUT.xml]\n"
                 "          [--device D | --devices 0,1,..|0-7] [--oversampling 1|2|4|8] [--state-in CKPT] [--state-out CKPT]\n"
                 "          [--out-dir DIR] IN.wav [IN2.wav ...]\n"
                 "       %s --list                 plugin keys, channel counts and sliders the library renders\n"
                 "       %s --wav-info FILE.wav    header + checksum of the decoded samples (no GPU needed)\n",
                 argv0, argv0, argv0);
    return 2;
}

int fail(const std::string& what)
{
    std::fprintf(stderr, "za_bounce: %s\n", what.c_str());
    return 1;
}

std::string base_name(const std::string& p)
{
    const size_t s = p.find_last_of('/');
    std::string b = s == std::string::npos ? p : p.substr(s + 1);
    const size_t d = b.find_last_of('.');
    return d == std::string::npos ? b : b.substr(0, d);
}

std::string json_escape(const std::string& s)
{
    std::string o;
    for (char c : s) {
        if (c == '"' || c == '\\') { o += '\\'; o += c; }
        else if ((unsigned char)c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += c;
    }
    return o;
}

bool slurp(const std::string& path, std::string& out)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char buf[4096];
    size_t n;
    out.clear();
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, n);
    std::fclose(f);
    return true;
}

// value of attribute `name` inside the tag text [b, e)
bool attr(const std::string& t, size_t b, size_t e, const char* name, std::string& out)
{
    const std::string key = std::string(name) + "=\"";
    const size_t k = t.find(key, b);
    if (k == std::string::npos || k >= e) return false;
    const size_t q = t.find('"', k + key.size());
    if (q == std::string::npos || q > e) return false;
    out = t.substr(k + key.size(), q - k - key.size());
    return true;
}

// APVTS state as the reference saves it (getStateInformation, src/JSFXJuceProcessor.cpp:3777-3880: the ValueTree's XML,
// one <PARAM id="sliderN" value="..."/> per parameter, ids from :775): slider number -> host value.  Other children
// (FILES, FILE_UI, SMART_IDLE, string sliders) are host-only state and are ignored.
bool read_preset(const std::string& path, std::map<int, float>& out, std::string& err)
{
    std::string t;
    if (!slurp(path, t)) { err = "cannot open " + path; return false; }
    size_t pos = 0;
    int found = 0;
    while ((pos = t.find("<PARAM", pos)) != std::string::npos) {
        const size_t e = t.find('>', pos);
        if (e == std::string::npos) break;
        std::string id, val;
        if (attr(t, pos, e, "id", id) && attr(t, pos, e, "value", val) && id.compare(0, 6, "slider") == 0) {
            out[std::atoi(id.c_str() + 6)] = (float)std::atof(val.c_str());
            ++found;
        }
        pos = e;
    }
    if (!found) { err = path + ": no <PARAM id=\"sliderN\" value=...> elements"; return false; }
    return true;
}

// plugin.json of a catalogue leaf: the fields scripts/pluginlib.py:105-212 requires, checked the way load_plugin does (the
// packaging-only checks -- README present, location under plugins/<Category>/<Key>/ -- are not this tool's business).
bool json_string(const std::string& t, const char* key, std::string& out)
{
    const size_t k = t.find(std::string("\"") + key + "\"");
    const size_t c = k == std::string::npos ? k : t.find(':', k);
    if (c == std::string::npos) return false;
    size_t q0 = c + 1;
    while (q0 < t.size() && std::isspace((unsigned char)t[q0])) ++q0;
    if (q0 >= t.size() || t[q0] != '"') return false;
    const size_t q1 = t.find('"', q0 + 1);
    if (q1 == std::string::npos) return false;
    out = t.substr(q0 + 1, q1 - q0 - 1);
    return true;
}
bool ends_with(const std::string& a, const char* suf) { const size_t n = std::strlen(suf); return a.size() >= n && a.compare(a.size() - n, n, suf) == 0; }

bool read_plugin_json(const std::string& path, std::string& slug, std::string& err)
{
    std::string t;
    if (!slurp(path, t)) { err = "cannot open " + path; return false; }
    std::string name, code, mcode, type, entry, tmp;
    if (!json_string(t, "slug", slug) || slug.empty()) { err = path + ": Missing 'slug'"; return false; }
    if (!json_string(t, "pluginCode", code) || code.empty()) { err = path + ": Missing 'pluginCode'"; return false; }
    if (code.size() != 4) { err = path + ": pluginCode must be exactly 4 characters"; return false; }
    if (json_string(t, "manufacturerCode", mcode) && mcode.size() != 4) { err = path + ": manufacturerCode must be exactly 4 characters"; return false; }
    if (t.find("\"installPath\"") != std::string::npos) { err = path + ": installPath overrides are no longer supported"; return false; }
    json_string(t, "entry", entry);
    if (json_string(t, "pluginType", type)) for (auto& ch : type) ch = (char)std::tolower((unsigned char)ch);
    if (type.empty()) type = ends_with(entry, ".dsp") ? "faust" : (ends_with(entry, ".jsfx") ? "jsfx" : "");
    if (type != "faust" && type != "jsfx") { err = path + ": Invalid pluginType (expected 'faust' or 'jsfx')"; return false; }
    if (!entry.empty() && type == "faust" && !ends_with(entry, ".dsp")) { err = path + ": Faust plugin entry must be a .dsp file"; return false; }
    if (!entry.empty() && type == "jsfx" && !ends_with(entry, ".jsfx")) { err = path + ": JSFX plugin entry must be a .jsfx file"; return false; }
    za_plugin_info pi;                                   // and the library must render this leaf as the same kind of plugin
    if (za_gpu_plugin_info(slug.c_str(), &pi) == ZA_OK && ((pi.kind == 1) != (type == "faust"))) {
        err = path + ": pluginType '" + type + "' does not match the library's " + (pi.kind == 1 ? "Faust" : "DSP-JSFX") + " leaf " + slug;
        return false;
    }
    return true;
}

bool parse_devices(const std::string& spec, std::vector<int>& out)     // "0,1,2" or "0-7" or a mix
{
    size_t i = 0;
    while (i < spec.size()) {
        char* e = nullptr;
        const long a = std::strtol(spec.c_str() + i, &e, 10);
        if (e == spec.c_str() + i || a < 0) return false;
        long b = a;
        i = (size_t)(e - spec.c_str());
        if (i < spec.size() && spec[i] == '-') { b = std::strtol(spec.c_str() + i + 1, &e, 10); if (b < a) return false; i = (size_t)(e - spec.c_str()); }
        for (long d = a; d <= b; ++d) out.push_back((int)d);
        if (i < spec.size()) { if (spec[i] != ',') return false; ++i; }
    }
    return !out.empty();
}

int list_plugins()
{
    for (int i = 0; i < za_gpu_plugin_count(); ++i) {
        const char* key = za_gpu_plugin_key(i);
        za_plugin_info pi;
        if (za_gpu_plugin_info(key, &pi) != ZA_OK) return fail(za_gpu_last_error());
        std::printf("%s: %d in, %d out, %d sliders, latency %d\n", key, pi.n_in, pi.n_out, pi.n_sliders, pi.latency);
        for (int c = 0; c < pi.n_sliders; ++c) {
            za_slider_info si;
            if (za_gpu_slider_info(key, c, &si) != ZA_OK) return fail(za_gpu_last_error());
            std::printf("    slider%d%s%s = %g  <%g, %g, %g>\n", si.index0 + 1, si.alias[0] ? ":" : "", si.alias, si.def, si.min, si.max, si.step);
        }
    }
    return 0;
}

int wav_info(const std::string& path)
{
    za_wav::Audio a;
    const std::string e = za_wav::read(path, a);
    if (!e.empty()) return fail(e);
    uint64_t h = 1469598103934665603ull;                   // FNV-1a over the decoded floats, channel-major
    for (const auto& c : a.ch)
        for (float v : c) {
            uint32_t u;
            std::memcpy(&u, &v, 4);
            for (int k = 0; k < 4; ++k) { h ^= (u >> (8 * k)) & 0xffu; h *= 1099511628211ull; }
        }
    std::printf("{\"channels\": %d, \"sample_rate\": %.3f, \"frames\": %lld, \"bits\": %d, \"float\": %s, \"fnv1a\": \"%016llx\"}\n", a.channels,
                a.sample_rate, (long long)a.frames, a.bits, a.is_float ? "true" : "false", (unsigned long long)h);
    return 0;
}

}  // namespace

int main(int argc, char** argv)
{
    std::string plugin, out_dir = "bounce_out", precision = "f64";
    int block = 512, device = 0, oversampling = 1;
    std::vector<int> devices{0};
    std::string state_in, state_out;
    double chunk_seconds = 10.0, tail_seconds = 0.0;
    std::map<int, float> sets;
    std::vector<std::string> inputs;
    std::string save_preset, err;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto need = [&](const char* name) -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "za_bounce: %s needs a value\n", name); std::exit(2); }
            return argv[++i];
        };
        if (a == "--list") return list_plugins();
        else if (a == "--wav-info") return wav_info(need("--wav-info"));
        else if (a == "--plugin") plugin = need("--plugin");
        else if (a == "--plugin-json") { if (!read_plugin_json(need("--plugin-json"), plugin, err)) return fail(err); }
        else if (a == "--devices") { devices.clear(); if (!parse_devices(need("--devices"), devices)) return usage(argv[0]); }
        else if (a == "--oversampling") oversampling = std::atoi(need("--oversampling"));
        else if (a == "--state-in") state_in = need("--state-in");
        else if (a == "--state-out") state_out = need("--state-out");
        else if (a == "--preset") { if (!read_preset(need("--preset"), sets, err)) return fail(err); }
        else if (a == "--save-preset") save_preset = need("--save-preset");
        else if (a == "--precision") precision = need("--precision");
        else if (a == "--block") block = std::atoi(need("--block"));
        else if (a == "--device") { device = std::atoi(need("--device")); devices.assign(1, device); }
        else if (a == "--chunk-seconds") chunk_seconds = std::atof(need("--chunk-seconds"));
        else if (a == "--tail-seconds") tail_seconds = std::atof(need("--tail-seconds"));
        else if (a == "--out-dir") out_dir = need("--out-dir");
        else if (a == "--set") {
            const std::string kv = need("--set");
            const size_t eq = kv.find('=');
            if (eq == std::string::npos) return usage(argv[0]);
            sets[std::atoi(kv.substr(0, eq).c_str())] = (float)std::atof(kv.substr(eq + 1).c_str());
        } else if (a == "-h" || a == "--help") { usage(argv[0]); return 0; }
        else if (a.size() > 2 && a[0] == '-' && a[1] == '-') return usage(argv[0]);
        else inputs.push_back(a);
    }
    if (plugin.empty() || inputs.empty() || block <= 0 || chunk_seconds <= 0.0 || (precision != "f32" && precision != "f64")) return usage(argv[0]);

    za_plugin_info pi;
    if (za_gpu_plugin_info(plugin.c_str(), &pi) != ZA_OK) return fail(za_gpu_last_error());
    const int ns = (int)inputs.size();

    // ---- read every job; one sample rate per batch (one handle = one srate, like one prepareToPlay) -------------------
    std::vector<za_wav::Audio> wav((size_t)ns);
    int64_t longest = 0;
    for (int s = 0; s < ns; ++s) {
        const std::string e = za_wav::read(inputs[(size_t)s], wav[(size_t)s]);
        if (!e.empty()) return fail(e);
        if (wav[(size_t)s].sample_rate != wav[0].sample_rate)
            return fail(inputs[(size_t)s] + ": sample rate differs from " + inputs[0] + " (render such files in separate batches)");
        longest = std::max(longest, wav[(size_t)s].frames);
    }
    const double sr = wav[0].sample_rate;
    const int64_t tail = (int64_t)std::llround(tail_seconds * sr);
    const int64_t total = longest + tail;
    int64_t chunk = (int64_t)std::llround(chunk_seconds * sr);
    chunk = std::max<int64_t>(block, chunk / block * block);            // chunks end on host-block boundaries

    // ---- parameters: declared defaults, --set overrides by slider number, same row for every stream ---------------------
    std::vector<float> row((size_t)pi.n_sliders);
    for (int c = 0; c < pi.n_sliders; ++c) {
        za_slider_info si;
        if (za_gpu_slider_info(plugin.c_str(), c, &si) != ZA_OK) return fail(za_gpu_last_error());
        row[(size_t)c] = si.def;
        const auto it = sets.find(si.index0 + 1);
        if (it != sets.end()) { row[(size_t)c] = it->second; sets.erase(it); }
    }
    if (!save_preset.empty()) {                      // the values this bounce used, in the form --preset reads
        FILE* f = std::fopen(save_preset.c_str(), "w");
        if (!f) return fail("cannot create " + save_preset);
        std::fprintf(f, "<?xml version=\"1.0\" encoding=\"UTF-8\"?>\n<PARAMS plugin=\"%s\">\n", json_escape(plugin).c_str());
        for (int c = 0; c < pi.n_sliders; ++c) {
            za_slider_info si;
            za_gpu_slider_info(plugin.c_str(), c, &si);
            std::fprintf(f, "  <PARAM id=\"slider%d\" value=\"%.9g\"/>\n", si.index0 + 1, (double)row[(size_t)c]);
        }
        std::fprintf(f, "</PARAMS>\n");
        std::fclose(f);
    }
    if (!sets.empty()) return fail("--set: plugin " + plugin + " declares no slider" + std::to_string(sets.begin()->first));
    std::vector<float> params((size_t)ns * (size_t)std::max(pi.n_sliders, 1));
    for (int s = 0; s < ns; ++s) std::copy(row.begin(), row.end(), params.begin() + (size_t)s * (size_t)pi.n_sliders);

    // one batch over the chosen devices: streams shard as contiguous ranges, one host thread per device inside za_batch_process
    if ((int)devices.size() > ns) devices.resize((size_t)ns);
    za_batch_handle h = nullptr;
    if (za_batch_create(plugin.c_str(), ns, sr, block, precision == "f64" ? ZA_F64 : ZA_F32, devices.data(), (int)devices.size(), oversampling, &h) != ZA_OK)
        return fail(za_gpu_last_error());
    if (za_batch_prepare(h, params.data()) != ZA_OK) return fail(za_gpu_last_error());
    // checkpoint / resume: the per-stream DSP state of every shard (za_gpu_state_save blobs, one after the other)
    auto for_each_shard_state = [&](bool load, const std::string& path) -> std::string {
        FILE* f = std::fopen(path.c_str(), load ? "rb" : "wb");
        if (!f) return "cannot open " + path;
        for (int g = 0; g < (int)devices.size(); ++g) {
            za_gpu_handle sh = nullptr;
            size_t bytes = 0;
            if (za_batch_shard(h, g, &sh) != ZA_OK || za_gpu_state_size(sh, &bytes) != ZA_OK) { std::fclose(f); return za_gpu_last_error(); }
            std::vector<char> blob(bytes);
            if (load) {
                if (std::fread(blob.data(), 1, bytes, f) != bytes) { std::fclose(f); return path + ": truncated or made for another batch shape"; }
                if (za_gpu_state_load(sh, blob.data(), bytes) != ZA_OK) { std::fclose(f); return za_gpu_last_error(); }
            } else {
                if (za_gpu_state_save(sh, blob.data(), bytes) != ZA_OK) { std::fclose(f); return za_gpu_last_error(); }
                if (std::fwrite(blob.data(), 1, bytes, f) != bytes) { std::fclose(f); return "write error on " + path; }
            }
        }
        std::fclose(f);
        return "";
    };
    if (!state_in.empty()) { const std::string e = for_each_shard_state(true, state_in); if (!e.empty()) return fail(e); }

    // ---- outputs: header now (lengths are known), samples appended chunk by chunk -----------------------------------------
    ::mkdir(out_dir.c_str(), 0777);
    std::vector<std::string> out_path((size_t)ns);
    std::vector<double> peak((size_t)ns, 0.0);
    for (int s = 0; s < ns; ++s) {
        out_path[(size_t)s] = out_dir + "/" + base_name(inputs[(size_t)s]) + (ns > 1 ? "." + std::to_string(s) : "") + ".wav";
        FILE* f = std::fopen(out_path[(size_t)s].c_str(), "wb");
        if (!f || !za_wav::write24_header(f, pi.n_out, wav[(size_t)s].frames + tail, sr)) return fail("cannot create " + out_path[(size_t)s]);
        std::fclose(f);
    }

    const int n_in = std::max(pi.n_in, 0), n_out = pi.n_out;
    std::vector<float> in((size_t)ns * (size_t)std::max(n_in, 1) * (size_t)chunk), out((size_t)ns * (size_t)n_out * (size_t)chunk);
    const auto t0 = std::chrono::steady_clock::now();
    for (int64_t f0 = 0; f0 < total; f0 += chunk) {
        const int64_t n = std::min(chunk, total - f0);
        for (int s = 0; s < ns; ++s) {
            const za_wav::Audio& w = wav[(size_t)s];
            for (int c = 0; c < n_in; ++c) {                     // channels the file does not have, and frames past its end, are silence
                float* dst = in.data() + ((size_t)s * (size_t)n_in + (size_t)c) * (size_t)chunk;
                const int64_t have = c < w.channels ? std::max<int64_t>(0, std::min(n, w.frames - f0)) : 0;
                if (have > 0) std::memcpy(dst, w.ch[(size_t)c].data() + f0, (size_t)have * sizeof(float));
                std::fill(dst + have, dst + n, 0.0f);
            }
        }
        if (za_batch_process(h, nullptr, 0, n_in ? in.data() : nullptr, out.data(), n, chunk) != ZA_OK) return fail(za_gpu_last_error());
        for (int s = 0; s < ns; ++s) {
            const int64_t keep = std::max<int64_t>(0, std::min(n, wav[(size_t)s].frames + tail - f0));
            if (keep == 0) continue;
            const float* src = out.data() + (size_t)s * (size_t)n_out * (size_t)chunk;
            for (int c = 0; c < n_out; ++c)
                for (int64_t i = 0; i < keep; ++i) peak[(size_t)s] = std::max(peak[(size_t)s], (double)std::fabs(src[(size_t)c * (size_t)chunk + (size_t)i]));
            FILE* f = std::fopen(out_path[(size_t)s].c_str(), "ab");
            if (!f || !za_wav::append24(f, src, n_out, keep, chunk)) return fail("write error on " + out_path[(size_t)s]);
            std::fclose(f);
        }
    }
    const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (!state_out.empty()) { const std::string e = for_each_shard_state(false, state_out); if (!e.empty()) return fail(e); }
    int64_t launches = 0;
    for (int g = 0; g < (int)devices.size(); ++g) {
        za_gpu_handle sh = nullptr;
        int64_t l = 0;
        if (za_batch_shard(h, g, &sh) == ZA_OK && za_gpu_launch_count(sh, &l) == ZA_OK) launches += l;
    }
    za_batch_destroy(h);

    // ---- report.json (the reference's exporter writes one beside its WAVs, JSFXCorrectnessCheck.h:1230-1255) ---------------
    const std::string rp = out_dir + "/report.json";
    FILE* r = std::fopen(rp.c_str(), "w");
    if (!r) return fail("cannot create " + rp);
    int64_t sum_frames = 0;
    for (int s = 0; s < ns; ++s) sum_frames += wav[(size_t)s].frames + tail;
    std::fprintf(r, "{\n  \"plugin\": \"%s\",\n  \"sampleRate\": %.3f,\n  \"frames\": %lld,\n  \"streams\": %d,\n  \"precision\": \"%s\",\n  \"block\": %d,\n"
                    "  \"chunkFrames\": %lld,\n  \"tailFrames\": %lld,\n  \"devices\": %d,\n  \"oversampling\": %d,\n  \"kernelLaunches\": %lld,\n  \"wallSeconds\": %.6f,\n  \"xRealtime\": %.3f,\n  \"files\": [\n",
                 json_escape(plugin).c_str(), sr, (long long)total, ns, precision.c_str(), block, (long long)chunk, (long long)tail, (int)devices.size(), oversampling, (long long)launches, wall,
                 wall > 0 ? (double)sum_frames / sr / wall : 0.0);
    for (int s = 0; s < ns; ++s)
        std::fprintf(r, "    {\"in\": \"%s\", \"out\": \"%s\", \"channelsIn\": %d, \"channelsOut\": %d, \"frames\": %lld, \"peakOut\": %.9g}%s\n",
                     json_escape(inputs[(size_t)s]).c_str(), json_escape(out_path[(size_t)s]).c_str(), wav[(size_t)s].channels, n_out,
                     (long long)(wav[(size_t)s].frames + tail), peak[(size_t)s], s + 1 < ns ? "," : "");
    std::fprintf(r, "  ]\n}\n");
    std::fclose(r);
    std::printf("za_bounce: %d stream(s) x %.3f s of %s in %.3f s (%.0f x realtime), %lld kernel launches -> %s\n", ns, (double)total / sr, plugin.c_str(), wall,
                wall > 0 ? (double)sum_frames / sr / wall : 0.0, (long long)launches, out_dir.c_str());
    return 0;
}
