// Minimal PNG reader/writer (zlib) for the host side: image textures in
// (texture/image.rs:29 uses the `image` crate) and the CLI's PNG output
// (src/bin/raysnail.rs:431-445).  8-bit, non-interlaced, colour types 0/2/4/6.
#include <zlib.h>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>

#include "scene.hpp"

namespace raysnail {

static uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

std::shared_ptr<Image> load_png(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("cannot open image " + path);
    std::vector<uint8_t> file((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (file.size() < 8 || std::memcmp(file.data(), sig, 8) != 0) throw std::runtime_error("not a PNG/PPM image: " + path);
    size_t pos = 8;
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat;
    while (pos + 12 <= file.size()) {
        uint32_t len = be32(&file[pos]);
        std::string type((const char*)&file[pos + 4], 4);
        const uint8_t* data = &file[pos + 8];
        if (pos + 12 + len > file.size()) throw std::runtime_error("truncated PNG: " + path);
        if (type == "IHDR") {
            w = be32(data); h = be32(data + 4); depth = data[8]; ctype = data[9]; interlace = data[12];
        } else if (type == "IDAT") {
            idat.insert(idat.end(), data, data + len);
        } else if (type == "IEND") {
            break;
        }
        pos += 12 + len;
    }
    if (depth != 8 || interlace != 0 || (ctype != 0 && ctype != 2 && ctype != 4 && ctype != 6))
        throw std::runtime_error("unsupported PNG format (need 8-bit non-interlaced grey/RGB/RGBA): " + path);
    const int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 4 ? 2 : 4;
    const size_t stride = (size_t)w * ch;
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size())
        throw std::runtime_error("PNG inflate failed: " + path);
    std::vector<uint8_t> img(stride * h);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t* src = &raw[(stride + 1) * y];
        uint8_t* dst = &img[stride * y];
        const uint8_t* up = y ? &img[stride * (y - 1)] : nullptr;
        int filter = src[0];
        for (size_t x = 0; x < stride; ++x) {
            int a = x >= (size_t)ch ? dst[x - ch] : 0;
            int b = up ? up[x] : 0;
            int c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
            int v = src[1 + x];
            switch (filter) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) / 2; break;
                case 4: {
                    int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
                    v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                    break;
                }
                default: throw std::runtime_error("bad PNG filter: " + path);
            }
            dst[x] = (uint8_t)v;
        }
    }
    std::vector<uint8_t> rgba((size_t)w * h * 4);
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        const uint8_t* p = &img[i * ch];
        uint8_t r, g, b, a = 255;
        if (ch == 1) { r = g = b = p[0]; }
        else if (ch == 2) { r = g = b = p[0]; a = p[1]; }
        else { r = p[0]; g = p[1]; b = p[2]; if (ch == 4) a = p[3]; }
        rgba[4 * i] = r; rgba[4 * i + 1] = g; rgba[4 * i + 2] = b; rgba[4 * i + 3] = a;
    }
    return std::make_shared<Image>((int)w, (int)h, std::move(rgba));
}

static void put_chunk(std::vector<uint8_t>& out, const char* type, const uint8_t* data, size_t len) {
    auto be = [&](uint32_t v) { out.push_back(v >> 24); out.push_back(v >> 16); out.push_back(v >> 8); out.push_back(v); };
    be((uint32_t)len);
    size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    if (len) out.insert(out.end(), data, data + len);
    be((uint32_t)crc32(0, &out[start], (uInt)(out.size() - start)));
}

bool write_png_rgb8(const std::string& path, const uint8_t* rgb, int w, int h) {
    std::vector<uint8_t> raw((size_t)(w * 3 + 1) * h);
    for (int y = 0; y < h; ++y) {
        raw[(size_t)(w * 3 + 1) * y] = 0;
        std::memcpy(&raw[(size_t)(w * 3 + 1) * y + 1], rgb + (size_t)w * 3 * y, (size_t)w * 3);
    }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) return false;
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    uint8_t ihdr[13] = {(uint8_t)(w >> 24), (uint8_t)(w >> 16), (uint8_t)(w >> 8), (uint8_t)w,
                        (uint8_t)(h >> 24), (uint8_t)(h >> 16), (uint8_t)(h >> 8), (uint8_t)h, 8, 2, 0, 0, 0};
    put_chunk(out, "IHDR", ihdr, 13);
    put_chunk(out, "IDAT", comp.data(), clen);
    put_chunk(out, "IEND", nullptr, 0);
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    std::fclose(f);
    return ok;
}

}  // namespace raysnail
