// za_wav.h -- minimal RIFF/WAVE reader and 24-bit writer for the offline bounce driver (za_bounce.cpp).
// the step either side of the hot path.  The reference's nearest equivalent is the
// correctness exporter, which writes 24-bit PCM WAV through JUCE's WavAudioFormat
// (src/JSFXCorrectnessCheck.h:1177-1203); this header has no JUCE and reads what a DAW export contains:
// PCM 8/16/24/32 bit, IEEE float 32/64 bit, plain or WAVE_FORMAT_EXTENSIBLE headers.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace za_wav {

struct Audio {
    int channels = 0;
    double sample_rate = 0.0;
    int64_t frames = 0;
    int bits = 0;
    bool is_float = false;
    std::vector<std::vector<float>> ch;      // planar, [channel][frame]
};

inline uint32_t rd32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
inline uint16_t rd16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

// Returns "" on success, else what is wrong with the file.
inline std::string read(const std::string& path, Audio& a)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return "cannot open " + path;
    std::vector<uint8_t> buf;
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (sz < 12) { std::fclose(f); return path + ": not a RIFF file"; }
    buf.resize((size_t)sz);
    const size_t got = std::fread(buf.data(), 1, buf.size(), f);
    std::fclose(f);
    if (got != buf.size()) return path + ": short read";
    if (std::memcmp(buf.data(), "RIFF", 4) != 0 || std::memcmp(buf.data() + 8, "WAVE", 4) != 0) return path + ": not a RIFF/WAVE file";
    int fmt_tag = 0, block_align = 0;
    const uint8_t* data = nullptr;
    size_t data_len = 0;
    size_t pos = 12;
    while (pos + 8 <= buf.size()) {
        const uint8_t* ck = buf.data() + pos;
        size_t len = rd32(ck + 4);
        const size_t body = pos + 8;
        if (body + len > buf.size()) len = buf.size() - body;       // truncated / streamed files: take what is there
        if (std::memcmp(ck, "fmt ", 4) == 0 && len >= 16) {
            fmt_tag = rd16(ck + 8);
            a.channels = rd16(ck + 10);
            a.sample_rate = (double)rd32(ck + 12);
            block_align = rd16(ck + 20);
            a.bits = rd16(ck + 22);
            if (fmt_tag == 0xFFFE && len >= 40) fmt_tag = rd16(ck + 8 + 24);   // WAVE_FORMAT_EXTENSIBLE: first word of the sub-format GUID
        } else if (std::memcmp(ck, "data", 4) == 0) {
            data = buf.data() + body;
            data_len = len;
        }
        pos = body + len + (len & 1);
    }
    if (!data || a.channels <= 0 || a.channels > 64) return path + ": no fmt/data chunk or bad channel count";
    if (fmt_tag != 1 && fmt_tag != 3) return path + ": unsupported WAVE format tag " + std::to_string(fmt_tag);
    a.is_float = fmt_tag == 3;
    const int bytes = a.bits / 8;
    if ((a.is_float && bytes != 4 && bytes != 8) || (!a.is_float && (bytes < 1 || bytes > 4))) return path + ": unsupported sample size";
    if (block_align < bytes * a.channels) block_align = bytes * a.channels;
    a.frames = (int64_t)(data_len / (size_t)block_align);
    a.ch.assign((size_t)a.channels, std::vector<float>((size_t)a.frames));
    for (int64_t i = 0; i < a.frames; ++i) {
        const uint8_t* fr = data + (size_t)i * block_align;
        for (int c = 0; c < a.channels; ++c) {
            const uint8_t* s = fr + (size_t)c * bytes;
            float v;
            if (a.is_float) {
                if (bytes == 4) { std::memcpy(&v, s, 4); }
                else { double d; std::memcpy(&d, s, 8); v = (float)d; }
            } else if (bytes == 1) {
                v = ((int)s[0] - 128) * (1.0f / 128.0f);
            } else if (bytes == 2) {
                v = (int16_t)rd16(s) * (1.0f / 32768.0f);
            } else if (bytes == 3) {
                const int32_t x = (int32_t)(((uint32_t)s[0] << 8) | ((uint32_t)s[1] << 16) | ((uint32_t)s[2] << 24)) >> 8;
                v = (float)x * (1.0f / 8388608.0f);
            } else {
                v = (float)((double)(int32_t)rd32(s) * (1.0 / 2147483648.0));
            }
            a.ch[(size_t)c][(size_t)i] = v;
        }
    }
    return "";
}

// float -> 24-bit PCM the way JUCE's writer does it for the reference's exporter: clamp to [-1, 1], scale by 0x7fffff,
// round to nearest.
inline int32_t to_i24(float v)
{
    double x = (double)v;
    if (!(x == x)) x = 0.0;
    if (x > 1.0) x = 1.0;
    if (x < -1.0) x = -1.0;
    return (int32_t)std::lrint(x * 8388607.0);
}

// 44-byte header of a 24-bit PCM file whose length is known up front
inline bool write24_header(FILE* f, int channels, int64_t frames, double sample_rate)
{
    const uint32_t data_len = (uint32_t)((uint64_t)frames * (uint64_t)channels * 3u);
    uint8_t h[44];
    auto w32 = [&](int o, uint32_t v) { h[o] = (uint8_t)v; h[o + 1] = (uint8_t)(v >> 8); h[o + 2] = (uint8_t)(v >> 16); h[o + 3] = (uint8_t)(v >> 24); };
    auto w16 = [&](int o, uint32_t v) { h[o] = (uint8_t)v; h[o + 1] = (uint8_t)(v >> 8); };
    std::memcpy(h, "RIFF", 4); w32(4, 36 + data_len + (data_len & 1)); std::memcpy(h + 8, "WAVEfmt ", 8); w32(16, 16); w16(20, 1); w16(22, (uint32_t)channels);
    w32(24, (uint32_t)std::lrint(sample_rate)); w32(28, (uint32_t)std::lrint(sample_rate) * (uint32_t)channels * 3u); w16(32, (uint32_t)channels * 3u); w16(34, 24);
    std::memcpy(h + 36, "data", 4); w32(40, data_len);
    return std::fwrite(h, 1, 44, f) == 44;
}

// planar [channel][frame] (channel stride `stride` floats), n frames -> interleaved 24-bit samples appended to f
inline bool append24(FILE* f, const float* planar, int channels, int64_t n, int64_t stride)
{
    std::vector<uint8_t> row((size_t)channels * 3u * 4096u);
    for (int64_t i0 = 0; i0 < n; i0 += 4096) {
        const int64_t m = n - i0 < 4096 ? n - i0 : 4096;
        uint8_t* o = row.data();
        for (int64_t i = 0; i < m; ++i)
            for (int c = 0; c < channels; ++c) {
                const int32_t q = to_i24(planar[(size_t)c * (size_t)stride + (size_t)(i0 + i)]);
                *o++ = (uint8_t)q; *o++ = (uint8_t)(q >> 8); *o++ = (uint8_t)(q >> 16);
            }
        if (std::fwrite(row.data(), 1, (size_t)(o - row.data()), f) != (size_t)(o - row.data())) return false;
    }
    return true;
}

}  // namespace za_wav
