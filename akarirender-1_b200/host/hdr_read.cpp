// hdr_read.cpp — Radiance .hdr (RGBE) reader for image textures.
//
// The reference reads `.hdr` files with `stbi_loadf(path, &x, &y, &channel, 3)` (core/image.cpp:85-103); stb_image is a
// third-party header vendored by the reference (external/misc/stb_image.h, v2.x), absent from this repo. This restates
// its published HDR path (stbi__hdr_load / stbi__hdr_convert): header tokens up to the empty line, FORMAT=32-bit_rle_rgbe
// required, resolution line `-Y h +X w`, then flat RGBE pixels (width < 8 or >= 32768) or new-style per-channel RLE
// scanlines; a pixel converts as byte * ldexp(1, e - 136) per channel (exactly representable: no rounding), e = 0 -> 0.
// No gamma, no flip. Pinned by tests against the reference's own reader (tests/golden/make_golden_r2.py).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "scene_file.h"

namespace akb_host {
namespace {

struct Bytes {
    const std::vector<uint8_t> &b;
    size_t p = 0;
    int get8() { return p < b.size() ? b[p++] : 0; } // stbi__get8 returns 0 at the end of the data
    bool eof() const { return p >= b.size(); }
};

// stbi__hdr_gettoken: one line without its newline, at most 1023 characters (the rest of a longer line is skipped)
std::string token(Bytes &s) {
    std::string t;
    int c = s.get8();
    while (!s.eof() && c != '\n') {
        t.push_back((char)c);
        if (t.size() == 1023) {
            while (!s.eof() && s.get8() != '\n') {
            }
            break;
        }
        c = s.get8();
    }
    return t;
}

void convert(float *out, const uint8_t *rgbe) { // stbi__hdr_convert, req_comp = 3
    if (rgbe[3] != 0) {
        const float f1 = (float)ldexp(1.0f, (int)rgbe[3] - (128 + 8));
        out[0] = rgbe[0] * f1, out[1] = rgbe[1] * f1, out[2] = rgbe[2] * f1;
    } else {
        out[0] = out[1] = out[2] = 0.0f;
    }
}

} // namespace

void read_hdr_rgb(const std::string &path, int *width, int *height, std::vector<float> *rgb) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("error loading " + path + ": cannot open");
    std::vector<uint8_t> b;
    uint8_t tmp[65536];
    for (size_t n; (n = fread(tmp, 1, sizeof tmp, f)) > 0;) b.insert(b.end(), tmp, tmp + n);
    fclose(f);
    auto fail = [&](const std::string &why) { throw std::runtime_error("error loading " + path + ": " + why); };
    Bytes s{b};
    const std::string id = token(s);
    if (id != "#?RADIANCE" && id != "#?RGBE") fail("not a Radiance HDR file");
    bool valid = false;
    for (;;) {
        const std::string t = token(s);
        if (t.empty()) break;
        if (t == "FORMAT=32-bit_rle_rgbe") valid = true;
    }
    if (!valid) fail("unsupported HDR format (need FORMAT=32-bit_rle_rgbe)");
    const std::string res = token(s);
    if (res.compare(0, 3, "-Y ") != 0) fail("unsupported HDR data layout (need -Y h +X w)");
    char *end = nullptr;
    const long h = strtol(res.c_str() + 3, &end, 10);
    while (*end == ' ') ++end;
    if (strncmp(end, "+X ", 3) != 0) fail("unsupported HDR data layout (need -Y h +X w)");
    const long w = strtol(end + 3, nullptr, 10);
    if (w <= 0 || h <= 0 || w > 65536 || h > 65536 || (size_t)w * h > ((size_t)1 << 28)) fail("bad HDR resolution");
    rgb->assign((size_t)w * h * 3, 0.0f);
    float *out = rgb->data();
    auto flat_from = [&](size_t first_pixel) { // the rest of the file as flat RGBE pixels
        for (size_t i = first_pixel; i < (size_t)w * h; i++) {
            uint8_t px[4];
            for (int k = 0; k < 4; k++) px[k] = (uint8_t)s.get8();
            convert(out + 3 * i, px);
        }
    };
    if (w < 8 || w >= 32768) {
        flat_from(0);
    } else {
        std::vector<uint8_t> line((size_t)w * 4);
        for (long j = 0; j < h; j++) {
            const int c1 = s.get8(), c2 = s.get8();
            int len = s.get8();
            if (c1 != 2 || c2 != 2 || (len & 0x80)) {
                // not run-length encoded: stb takes these bytes as pixel 0 and reads everything after it as flat pixels
                const uint8_t px[4] = {(uint8_t)c1, (uint8_t)c2, (uint8_t)len, (uint8_t)s.get8()};
                convert(out, px);
                flat_from(1);
                break;
            }
            len = (len << 8) | s.get8();
            if (len != w) fail("corrupt HDR: invalid decoded scanline length");
            for (int k = 0; k < 4; k++) {
                long i = 0, nleft;
                while ((nleft = w - i) > 0) {
                    int count = s.get8();
                    if (count > 128) { // run
                        const int value = s.get8();
                        count -= 128;
                        if (count > nleft) fail("corrupt HDR: bad RLE data");
                        for (int z = 0; z < count; z++) line[(size_t)(i++) * 4 + k] = (uint8_t)value;
                    } else { // dump
                        if (count > nleft) fail("corrupt HDR: bad RLE data");
                        if (count == 0 && s.eof()) fail("corrupt HDR: truncated");
                        for (int z = 0; z < count; z++) line[(size_t)(i++) * 4 + k] = (uint8_t)s.get8();
                    }
                }
            }
            for (long i = 0; i < w; i++) convert(out + 3 * ((size_t)j * w + i), &line[(size_t)i * 4]);
        }
    }
    *width = (int)w, *height = (int)h;
}

} // namespace akb_host
