// hj_hdr.h — Radiance RGBE (.hdr) reader for the sky image: the reference loads it with stbi_loadf and copies RGB into
// a float4 image with w = 0 (/root/reference/include/renderer/texture.h:67-101). stb_image converts an RGBE texel as
// rgb * ldexp(1, e - (128 + 8)) and applies no gamma to float loads of .hdr files; the same is done here.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "hj_png.h"

namespace hj {

// returns float4 pixels (w = 0), row 0 first
inline std::vector<float> hdr_read(const std::string& path, int& W, int& H) {
    const std::vector<uint8_t> f = read_file(path);
    size_t p = 0;
    auto line = [&]() { std::string s; while (p < f.size() && f[p] != '\n') s += (char)f[p++]; if (p < f.size()) p++; return s; };
    const std::string magic = line();
    if (magic.compare(0, 2, "#?") != 0) throw std::runtime_error(path + ": not a Radiance HDR file");
    bool fmt = false;
    while (true) {
        const std::string l = line();
        if (l.empty()) break;
        if (l == "FORMAT=32-bit_rle_rgbe") fmt = true;
        if (p >= f.size()) break;
    }
    if (!fmt) throw std::runtime_error(path + ": unsupported HDR format");
    const std::string res = line();
    if (sscanf(res.c_str(), "-Y %d +X %d", &H, &W) != 2) throw std::runtime_error(path + ": unsupported HDR orientation");
    std::vector<uint8_t> rgbe((size_t)W * H * 4);
    for (int y = 0; y < H; y++) {
        uint8_t* row = &rgbe[(size_t)y * W * 4];
        if (W >= 8 && W < 32768 && p + 4 <= f.size() && f[p] == 2 && f[p + 1] == 2 && !(f[p + 2] & 0x80) && ((f[p + 2] << 8) | f[p + 3]) == W) {
            p += 4;
            for (int c = 0; c < 4; c++) {           // each channel run-length coded on its own
                int x = 0;
                while (x < W) {
                    if (p >= f.size()) throw std::runtime_error(path + ": truncated");
                    int n = f[p++];
                    if (n > 128) { n -= 128; if (x + n > W || p >= f.size()) throw std::runtime_error(path + ": bad run"); const uint8_t v = f[p++]; for (int k = 0; k < n; k++) row[4 * (x++) + c] = v; }
                    else { if (n == 0 || x + n > W || p + (size_t)n > f.size()) throw std::runtime_error(path + ": bad run"); for (int k = 0; k < n; k++) row[4 * (x++) + c] = f[p++]; }
                }
            }
        } else {                                    // flat scanline
            if (p + (size_t)W * 4 > f.size()) throw std::runtime_error(path + ": truncated");
            memcpy(row, &f[p], (size_t)W * 4); p += (size_t)W * 4;
        }
    }
    std::vector<float> out((size_t)W * H * 4);
    for (size_t i = 0; i < (size_t)W * H; i++) {
        const uint8_t* q = &rgbe[4 * i];
        if (q[3]) { const float s = (float)ldexp(1.0f, (int)q[3] - (128 + 8)); out[4 * i] = q[0] * s; out[4 * i + 1] = q[1] * s; out[4 * i + 2] = q[2] * s; }
        else out[4 * i] = out[4 * i + 1] = out[4 * i + 2] = 0.0f;
        out[4 * i + 3] = 0.0f;
    }
    return out;
}

}  // namespace hj
