// png_read.cpp — minimal PNG reader for image textures (the reference reads them through stb_image,
// core/image.cpp:105-123: `stbi_load(path, &x, &y, &channel, 3)`, then byte / 255.0f per channel).
// Supports what texture files in practice are: 8 bits per channel, grey / grey+alpha / RGB / RGBA, non-interlaced.
// Own inflate (RFC 1951: stored, fixed and dynamic Huffman blocks) - no zlib dependency in the library.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <vector>

#include "scene_file.h"

namespace akb_host {
namespace {

struct BitReader {
    const uint8_t *p, *end;
    uint32_t buf = 0;
    int cnt = 0;
    uint32_t bits(int n) {
        while (cnt < n) {
            if (p >= end) throw std::runtime_error("truncated deflate stream");
            buf |= (uint32_t)*p++ << cnt;
            cnt += 8;
        }
        const uint32_t v = n ? (buf & ((1u << n) - 1u)) : 0;
        buf >>= n, cnt -= n;
        return v;
    }
    void align() { buf = 0, cnt = 0; }
};

// canonical Huffman decoding table (counts per length + symbols sorted by code), decoded bit by bit
struct Huffman {
    uint16_t count[16] = {0}, symbol[320] = {0};
    void build(const uint8_t *lengths, int n) {
        memset(count, 0, sizeof count);
        for (int i = 0; i < n; i++) count[lengths[i]]++;
        count[0] = 0;
        uint16_t offs[16];
        offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
        for (int i = 0; i < n; i++)
            if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
    int decode(BitReader &br) const {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= 15; l++) {
            code |= (int)br.bits(1);
            const int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c, first += c;
            first <<= 1, code <<= 1;
        }
        throw std::runtime_error("bad Huffman code");
    }
};

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

// `limit`: the decoded size the caller expects (from IHDR); a stream that inflates beyond it is rejected before the
// memory is spent (a small crafted file can expand 1000:1).
void inflate_block(BitReader &br, const Huffman &lit, const Huffman &dist, std::vector<uint8_t> &out, size_t limit) {
    for (;;) {
        if (out.size() > limit) throw std::runtime_error("image data longer than the header says");
        const int s = lit.decode(br);
        if (s < 256) {
            out.push_back((uint8_t)s);
        } else if (s == 256) {
            return;
        } else {
            if (s > 285) throw std::runtime_error("bad length symbol");
            const int len = kLenBase[s - 257] + (int)br.bits(kLenExtra[s - 257]);
            const int ds = dist.decode(br);
            if (ds > 29) throw std::runtime_error("bad distance symbol");
            const size_t d = kDistBase[ds] + br.bits(kDistExtra[ds]);
            if (d > out.size()) throw std::runtime_error("distance too far back");
            for (int i = 0; i < len; i++) out.push_back(out[out.size() - d]);
        }
    }
}

std::vector<uint8_t> zlib_inflate(const std::vector<uint8_t> &z, size_t limit) {
    if (z.size() < 6 || (z[0] & 0x0F) != 8 || ((z[0] << 8) | z[1]) % 31 != 0 || (z[1] & 0x20)) throw std::runtime_error("bad zlib header");
    BitReader br{z.data() + 2, z.data() + z.size()};
    std::vector<uint8_t> out;
    out.reserve(std::min<size_t>(limit, (size_t)64 << 20));
    Huffman fixed_lit, fixed_dist;
    {
        uint8_t l[288];
        for (int i = 0; i < 288; i++) l[i] = i < 144 ? 8 : (i < 256 ? 9 : (i < 280 ? 7 : 8));
        fixed_lit.build(l, 288);
        uint8_t d[30];
        memset(d, 5, sizeof d);
        fixed_dist.build(d, 30);
    }
    for (int last = 0; !last;) {
        last = (int)br.bits(1);
        const int type = (int)br.bits(2);
        if (type == 0) {
            br.align();
            if (br.end - br.p < 4) throw std::runtime_error("truncated stored block");
            const uint32_t len = br.p[0] | (br.p[1] << 8), nlen = br.p[2] | (br.p[3] << 8);
            br.p += 4;
            if ((len ^ 0xFFFFu) != nlen || (size_t)(br.end - br.p) < len) throw std::runtime_error("bad stored block");
            if (out.size() + len > limit) throw std::runtime_error("image data longer than the header says");
            out.insert(out.end(), br.p, br.p + len);
            br.p += len;
        } else if (type == 1) {
            inflate_block(br, fixed_lit, fixed_dist, out, limit);
        } else if (type == 2) {
            const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
            static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            uint8_t lengths[320];
            memset(lengths, 0, sizeof lengths);
            for (int i = 0; i < ncode; i++) lengths[order[i]] = (uint8_t)br.bits(3);
            Huffman cl;
            cl.build(lengths, 19);
            memset(lengths, 0, sizeof lengths);
            for (int i = 0; i < nlen + ndist;) {
                const int s = cl.decode(br);
                if (s < 16) {
                    lengths[i++] = (uint8_t)s;
                } else {
                    int rep, val = 0;
                    if (s == 16) {
                        if (i == 0) throw std::runtime_error("repeat without a previous length");
                        val = lengths[i - 1], rep = 3 + (int)br.bits(2);
                    } else if (s == 17) {
                        rep = 3 + (int)br.bits(3);
                    } else {
                        rep = 11 + (int)br.bits(7);
                    }
                    if (i + rep > nlen + ndist) throw std::runtime_error("too many code lengths");
                    while (rep--) lengths[i++] = (uint8_t)val;
                }
            }
            Huffman lit, dist;
            lit.build(lengths, nlen);
            dist.build(lengths + nlen, ndist);
            inflate_block(br, lit, dist, out, limit);
        } else {
            throw std::runtime_error("bad deflate block type");
        }
    }
    return out;
}

uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }

} // namespace

void read_png_rgb8(const std::string &path, int *width, int *height, std::vector<uint8_t> *rgb) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("error loading " + path + ": cannot open");
    std::vector<uint8_t> b;
    uint8_t tmp[65536];
    for (size_t n; (n = fread(tmp, 1, sizeof tmp, f)) > 0;) b.insert(b.end(), tmp, tmp + n);
    fclose(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    auto fail = [&](const std::string &why) { throw std::runtime_error("error loading " + path + ": " + why); };
    if (b.size() < 8 || memcmp(b.data(), sig, 8) != 0) fail("not a PNG file (only PNG image textures are supported)");
    int w = 0, h = 0, ctype = -1;
    std::vector<uint8_t> idat;
    for (size_t o = 8; o + 12 <= b.size();) {
        const uint32_t n = be32(&b[o]);
        if (o + 12 + n > b.size()) fail("truncated chunk");
        const char *tag = (const char *)&b[o + 4];
        if (!memcmp(tag, "IHDR", 4) && n >= 13) {
            w = (int)be32(&b[o + 8]), h = (int)be32(&b[o + 12]);
            ctype = b[o + 17];
            if (b[o + 16] != 8 || b[o + 20] != 0 || !(ctype == 0 || ctype == 2 || ctype == 4 || ctype == 6))
                fail("unsupported PNG variant (need 8 bits per channel, non-interlaced, no palette)");
        } else if (!memcmp(tag, "IDAT", 4)) {
            idat.insert(idat.end(), &b[o + 8], &b[o + 8] + n);
        }
        o += 12 + n;
    }
    if (w <= 0 || h <= 0 || ctype < 0) fail("no IHDR");
    if (w > 65536 || h > 65536 || (size_t)w * h > ((size_t)1 << 28)) fail("image too large (limit 65536 x 65536, 2^28 pixels)");
    const int ch = ctype == 0 ? 1 : (ctype == 2 ? 3 : (ctype == 4 ? 2 : 4));
    std::vector<uint8_t> raw;
    try {
        raw = zlib_inflate(idat, ((size_t)w * ch + 1) * (size_t)h);
    } catch (const std::exception &e) {
        fail(e.what());
    }
    const size_t stride = (size_t)w * ch;
    if (raw.size() < (stride + 1) * (size_t)h) fail("image data too short");
    std::vector<uint8_t> px(stride * h);
    for (int y = 0; y < h; y++) { // PNG filters 0..4 (RFC 2083 §6)
        const uint8_t *line = &raw[(stride + 1) * y + 1];
        const int ft = raw[(stride + 1) * y];
        uint8_t *cur = &px[stride * y];
        const uint8_t *prev = y ? &px[stride * (y - 1)] : nullptr;
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= (size_t)ch ? cur[i - ch] : 0, up = prev ? prev[i] : 0, c = (prev && i >= (size_t)ch) ? prev[i - ch] : 0;
            int pr = 0;
            if (ft == 1) pr = a;
            else if (ft == 2) pr = up;
            else if (ft == 3) pr = (a + up) >> 1;
            else if (ft == 4) {
                const int pa = abs(up - c), pb = abs(a - c), pc = abs(a + up - 2 * c);
                pr = (pa <= pb && pa <= pc) ? a : (pb <= pc ? up : c);
            } else if (ft != 0) fail("bad filter type");
            cur[i] = (uint8_t)(line[i] + pr);
        }
    }
    rgb->resize((size_t)w * h * 3);
    for (size_t i = 0; i < (size_t)w * h; i++) // stbi_load(..., 3): grey is replicated, alpha dropped
        for (int k = 0; k < 3; k++) (*rgb)[3 * i + k] = px[i * ch + (ch >= 3 ? k : 0)];
    *width = w, *height = h;
}

} // namespace akb_host
