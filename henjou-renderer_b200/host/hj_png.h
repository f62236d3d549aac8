// hj_png.h — PNG in and out for the host loop, on zlib only.
//   read : the glTF textures and the thin-film LUT, decoded to 8-bit RGBA the way the reference's Texture does through
//          stb_image with 4 requested channels (/root/reference/include/renderer/texture.h:16-39);
//   write: the frame, `sutil::saveImage(name, buffer, false)` of an UNSIGNED_BYTE4 buffer
//          (/root/reference/include/renderer/renderer.h:1285-1302), which stores rows bottom-up (row 0 of the launch is
//          the bottom of the picture) — so the writer flips vertically.
#pragma once
#include <zlib.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace hj {

struct Image8 { int width = 0, height = 0; std::vector<uint8_t> rgba; };

inline std::vector<uint8_t> read_file(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    std::vector<uint8_t> d;
    uint8_t buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) d.insert(d.end(), buf, buf + n);
    fclose(f);
    return d;
}

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// 8/16-bit, colour types 0 2 3 4 6, non-interlaced (what Blender's glTF exporter and the LUT generator write)
inline Image8 png_decode(const std::vector<uint8_t>& file, const std::string& name = "png") {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (file.size() < 8 || memcmp(file.data(), sig, 8) != 0) throw std::runtime_error(name + ": not a PNG file");
    uint32_t W = 0, H = 0; int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    size_t p = 8;
    while (p + 12 <= file.size()) {
        const uint32_t len = be32(&file[p]);
        const char* tag = (const char*)&file[p + 4];
        if (p + 12 + (size_t)len > file.size()) throw std::runtime_error(name + ": truncated chunk");
        const uint8_t* d = &file[p + 8];
        if (!memcmp(tag, "IHDR", 4)) { W = be32(d); H = be32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12]; }
        else if (!memcmp(tag, "PLTE", 4)) plte.assign(d, d + len);
        else if (!memcmp(tag, "tRNS", 4)) trns.assign(d, d + len);
        else if (!memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), d, d + len);
        else if (!memcmp(tag, "IEND", 4)) break;
        p += 12 + (size_t)len;
    }
    if (W == 0 || H == 0) throw std::runtime_error(name + ": no IHDR");
    if (interlace) throw std::runtime_error(name + ": interlaced PNG is not supported");
    if (depth != 8 && depth != 16) throw std::runtime_error(name + ": only 8- and 16-bit PNG is supported");
    const int nch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!nch || (ctype == 3 && depth != 8)) throw std::runtime_error(name + ": unsupported colour type");
    const size_t bpp = (size_t)nch * depth / 8, stride = (size_t)W * bpp;
    std::vector<uint8_t> raw((stride + 1) * H);
    uLongf out_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size())
        throw std::runtime_error(name + ": zlib stream is corrupt");
    std::vector<uint8_t> pix(stride * H);
    for (uint32_t y = 0; y < H; y++) {          // undo the scanline filters
        const uint8_t ft = raw[(stride + 1) * y];
        const uint8_t* in = &raw[(stride + 1) * y + 1];
        uint8_t* cur = &pix[stride * y];
        const uint8_t* up = y ? &pix[stride * (y - 1)] : nullptr;
        for (size_t x = 0; x < stride; x++) {
            const int a = x >= bpp ? cur[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
            int v = in[x];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: { const int pa = abs(b - c), pb = abs(a - c), pc = abs(a + b - 2 * c); v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
                default: throw std::runtime_error(name + ": bad filter type");
            }
            cur[x] = (uint8_t)v;
        }
    }
    Image8 im; im.width = (int)W; im.height = (int)H; im.rgba.resize((size_t)W * H * 4);
    const size_t step = depth / 8;               // 16-bit: keep the high byte (stb_image's 16 -> 8 conversion)
    for (size_t i = 0; i < (size_t)W * H; i++) {
        const uint8_t* s = &pix[i * bpp];
        uint8_t* o = &im.rgba[i * 4];
        switch (ctype) {
            case 0: o[0] = o[1] = o[2] = s[0]; o[3] = 255; break;
            case 2: o[0] = s[0]; o[1] = s[step]; o[2] = s[2 * step]; o[3] = 255; break;
            case 3: {
                const size_t k = s[0];
                if (3 * k + 2 >= plte.size()) throw std::runtime_error(name + ": palette index out of range");
                o[0] = plte[3 * k]; o[1] = plte[3 * k + 1]; o[2] = plte[3 * k + 2]; o[3] = k < trns.size() ? trns[k] : 255; break;
            }
            case 4: o[0] = o[1] = o[2] = s[0]; o[3] = s[step]; break;
            default: o[0] = s[0]; o[1] = s[step]; o[2] = s[2 * step]; o[3] = s[3 * step]; break;
        }
    }
    return im;
}
inline Image8 png_read(const std::string& path) { return png_decode(read_file(path), path); }

// RGBA8, filter 0, one IDAT; `flip`: write the rows bottom-up
inline void png_write(const std::string& path, const uint8_t* rgba, int W, int H, bool flip) {
    std::vector<uint8_t> raw((size_t)(W * 4 + 1) * H);
    for (int y = 0; y < H; y++) {
        uint8_t* row = &raw[(size_t)(W * 4 + 1) * y];
        row[0] = 0;
        memcpy(row + 1, rgba + (size_t)W * 4 * (flip ? H - 1 - y : y), (size_t)W * 4);
    }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) throw std::runtime_error("png_write: zlib failed");
    comp.resize(clen);
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    auto chunk = [&](const char* tag, const uint8_t* d, uint32_t n) {
        uint8_t hdr[8] = {(uint8_t)(n >> 24), (uint8_t)(n >> 16), (uint8_t)(n >> 8), (uint8_t)n, (uint8_t)tag[0], (uint8_t)tag[1], (uint8_t)tag[2], (uint8_t)tag[3]};
        fwrite(hdr, 1, 8, f);
        if (n) fwrite(d, 1, n, f);
        uLong c = crc32(0L, hdr + 4, 4);
        if (n) c = crc32(c, d, n);
        const uint8_t tail[4] = {(uint8_t)(c >> 24), (uint8_t)(c >> 16), (uint8_t)(c >> 8), (uint8_t)c};
        fwrite(tail, 1, 4, f);
    };
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    fwrite(sig, 1, 8, f);
    uint8_t ihdr[13] = {(uint8_t)(W >> 24), (uint8_t)(W >> 16), (uint8_t)(W >> 8), (uint8_t)W, (uint8_t)(H >> 24), (uint8_t)(H >> 16), (uint8_t)(H >> 8), (uint8_t)H, 8, 6, 0, 0, 0};
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", comp.data(), (uint32_t)comp.size());
    chunk("IEND", nullptr, 0);
    fclose(f);
}

}  // namespace hj
