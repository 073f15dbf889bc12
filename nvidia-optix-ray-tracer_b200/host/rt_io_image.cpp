// rt_io_image.cpp — headless image outputs replacing the reference's presentation layer
// (src/GraphicsAPI/*: pure image copies into a GL texture / Vulkan image / D3D surface).
//   PFM  : "PF" colour, little-endian float32 RGB, rows bottom-to-top (the format's convention)
//   PPM  : "P6" 8-bit RGB from an RGBA8 buffer
//   EXR  : OpenEXR 2 single-part scanline file, uncompressed, channels A B G R as float32
//          (no OpenEXR library in the image: the writer follows the published file layout)
#include <cstdint>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "rt_io_internal.hpp"

extern "C" {

int rt_io_write_pfm(const char *path, uint32_t w, uint32_t h, const float *rgba, int flip_y) {
    if (!path || !rgba || !w || !h) return fail(RT_IO_ERR_ARG, "bad argument");
    std::ofstream f(path, std::ios::out | std::ios::binary);
    if (!f.is_open()) return fail(RT_IO_ERR_OPEN, "Failed to open file: %s!", path);
    const std::string hdr = "PF\n" + std::to_string(w) + " " + std::to_string(h) + "\n-1.0\n";
    f.write(hdr.data(), (std::streamsize)hdr.size());
    std::vector<float> row(3 * (size_t)w);
    for (uint32_t y = 0; y < h; y++) {
        // PFM stores the bottom row first.  flip_y == 0: buffer row 0 is the TOP of the image.
        const uint32_t src = flip_y ? y : (h - 1 - y);
        const float *s = rgba + 4 * (size_t)src * w;
        for (uint32_t x = 0; x < w; x++) {
            row[3 * x] = s[4 * x];
            row[3 * x + 1] = s[4 * x + 1];
            row[3 * x + 2] = s[4 * x + 2];
        }
        f.write(reinterpret_cast<const char *>(row.data()), (std::streamsize)(row.size() * 4));
    }
    return f ? RT_IO_OK : fail(RT_IO_ERR_OPEN, "write failed: %s", path);
}

int rt_io_read_pfm(const char *path, uint32_t *w, uint32_t *h, float *rgb_out, size_t capacity_floats) {
    if (!path || !w || !h) return fail(RT_IO_ERR_ARG, "bad argument");
    std::string buf;
    if (!read_file(path, buf)) return fail(RT_IO_ERR_OPEN, "Failed to open file: %s!", path);
    unsigned ww = 0, hh = 0;
    float scale = 0;
    int used = 0;
    if (sscanf(buf.c_str(), "PF %u %u %f%n", &ww, &hh, &scale, &used) != 3) return fail(RT_IO_ERR_PARSE, "not a colour PFM: %s", path);
    used += 1;  // the single whitespace after the scale
    if (scale >= 0) return fail(RT_IO_ERR_FORMAT, "big-endian PFM not supported: %s", path);
    *w = ww;
    *h = hh;
    const size_t need = 3 * (size_t)ww * hh;
    if (!rgb_out) return RT_IO_OK;
    if (capacity_floats < need || buf.size() < used + need * 4) return fail(RT_IO_ERR_FORMAT, "PFM buffer too small / file truncated");
    for (uint32_t y = 0; y < hh; y++)  // return rows top-down
        std::memcpy(rgb_out + 3 * (size_t)(hh - 1 - y) * ww, buf.data() + used + 12 * (size_t)y * ww, 12 * (size_t)ww);
    return RT_IO_OK;
}

int rt_io_write_ppm(const char *path, uint32_t w, uint32_t h, const uint8_t *rgba8, int flip_y) {
    if (!path || !rgba8 || !w || !h) return fail(RT_IO_ERR_ARG, "bad argument");
    std::ofstream f(path, std::ios::out | std::ios::binary);
    if (!f.is_open()) return fail(RT_IO_ERR_OPEN, "Failed to open file: %s!", path);
    const std::string hdr = "P6\n" + std::to_string(w) + " " + std::to_string(h) + "\n255\n";
    f.write(hdr.data(), (std::streamsize)hdr.size());
    std::vector<uint8_t> row(3 * (size_t)w);
    for (uint32_t y = 0; y < h; y++) {
        const uint8_t *s = rgba8 + 4 * (size_t)(flip_y ? h - 1 - y : y) * w;
        for (uint32_t x = 0; x < w; x++) {
            row[3 * x] = s[4 * x];
            row[3 * x + 1] = s[4 * x + 1];
            row[3 * x + 2] = s[4 * x + 2];
        }
        f.write(reinterpret_cast<const char *>(row.data()), (std::streamsize)row.size());
    }
    return f ? RT_IO_OK : fail(RT_IO_ERR_OPEN, "write failed: %s", path);
}

int rt_io_write_exr(const char *path, uint32_t w, uint32_t h, const float *rgba, int flip_y) {
    if (!path || !rgba || !w || !h) return fail(RT_IO_ERR_ARG, "bad argument");
    std::string hd;
    auto put32 = [](std::string &s, uint32_t v) { s.append(reinterpret_cast<const char *>(&v), 4); };
    auto putf = [](std::string &s, float v) { s.append(reinterpret_cast<const char *>(&v), 4); };
    auto attr = [&](const char *name, const char *type, const std::string &val) {
        hd.append(name);
        hd.push_back('\0');
        hd.append(type);
        hd.push_back('\0');
        put32(hd, (uint32_t)val.size());
        hd.append(val);
    };
    put32(hd, 20000630u);  // magic
    put32(hd, 2u);         // version 2, single-part scanline
    {
        std::string ch;
        for (const char *c : {"A", "B", "G", "R"}) {  // alphabetical
            ch.append(c);
            ch.push_back('\0');
            put32(ch, 2u);  // FLOAT
            ch.push_back('\0');  // pLinear
            ch.append(3, '\0');
            put32(ch, 1u);
            put32(ch, 1u);
        }
        ch.push_back('\0');
        attr("channels", "chlist", ch);
    }
    attr("compression", "compression", std::string(1, '\0'));
    {
        std::string b;
        put32(b, 0);
        put32(b, 0);
        put32(b, w - 1);
        put32(b, h - 1);
        attr("dataWindow", "box2i", b);
        attr("displayWindow", "box2i", b);
    }
    attr("lineOrder", "lineOrder", std::string(1, '\0'));
    {
        std::string v;
        putf(v, 1.0f);
        attr("pixelAspectRatio", "float", v);
    }
    {
        std::string v;
        putf(v, 0.0f);
        putf(v, 0.0f);
        attr("screenWindowCenter", "v2f", v);
    }
    {
        std::string v;
        putf(v, 1.0f);
        attr("screenWindowWidth", "float", v);
    }
    hd.push_back('\0');
    std::ofstream f(path, std::ios::out | std::ios::binary);
    if (!f.is_open()) return fail(RT_IO_ERR_OPEN, "Failed to open file: %s!", path);
    f.write(hd.data(), (std::streamsize)hd.size());
    const uint64_t line_bytes = 16ull * w, chunk = 8 + line_bytes;
    uint64_t off = hd.size() + 8ull * h;
    for (uint32_t y = 0; y < h; y++, off += chunk) f.write(reinterpret_cast<const char *>(&off), 8);
    std::vector<float> line(4 * (size_t)w);
    for (uint32_t y = 0; y < h; y++) {
        const float *s = rgba + 4 * (size_t)(flip_y ? h - 1 - y : y) * w;
        for (uint32_t x = 0; x < w; x++) {
            line[x] = s[4 * x + 3];
            line[w + x] = s[4 * x + 2];
            line[2 * (size_t)w + x] = s[4 * x + 1];
            line[3 * (size_t)w + x] = s[4 * x];
        }
        const int32_t yy = (int32_t)y;
        const uint32_t sz = (uint32_t)line_bytes;
        f.write(reinterpret_cast<const char *>(&yy), 4);
        f.write(reinterpret_cast<const char *>(&sz), 4);
        f.write(reinterpret_cast<const char *>(line.data()), (std::streamsize)line_bytes);
    }
    return f ? RT_IO_OK : fail(RT_IO_ERR_OPEN, "write failed: %s", path);
}

}  // extern "C"
