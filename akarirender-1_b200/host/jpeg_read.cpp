// jpeg_read.cpp — JPEG texture files for the scene front-end.
//
// The reference reads every image that is not ".hdr" with `stbi_load(path, &x, &y, &channel, 3)` (core/image.cpp:105,
// stb_image v2.23 vendored under external/misc) and turns bytes into byte / 255.0f. A texel is a parity input (it
// scales a BSDF), so this reader has to return the SAME bytes as stb does, and a JPEG decoder is free in exactly
// three places, which are therefore restated from stb_image.h (its scalar paths; its SSE2 kernels are documented to
// produce identical results):
//   * the inverse DCT: stb's fixed-point "islow" transform, 12 fractional bits, two extra bits kept between the column
//     and the row pass, rounding constants 512 and 65536 + (128 << 17) (stb_image.h:2233-2322);
//   * chroma upsampling: the 3:1 "tent" filters, per row pair (stb_image.h:3240-3303) and the row stepping of
//     load_jpeg_image (stb_image.h:3669-3714); other factors replicate samples;
//   * YCbCr -> RGB in 20-bit fixed point with the constants rounded to 12 bits first (stb_image.h:3433-3460).
// Everything else (markers, Huffman decoding, spectral selection / successive approximation of progressive files,
// restart intervals) is ITU-T T.81 and written from it; any conforming decoder arrives at the same coefficients.
// Like stb: 8-bit precision, Huffman coding (SOF0 / SOF1 / SOF2), 1, 3 or 4 components, sampling factors 1..4; JFIF
// files are YCbCr, Adobe files say so themselves, three components named 'R','G','B' are RGB. Grey files are
// replicated to RGB (what stbi_load(..., 3) returns).
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <vector>

#include "scene_file.h"

namespace akb_host {
namespace {

[[noreturn]] void fail(const std::string &path, const char *what) { throw std::runtime_error(path + ": " + what); }

constexpr uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                 41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// Canonical Huffman table (T.81 annex C, decoding per F.2.2.3): codes of length l occupy [first[l], first[l] + count[l]).
struct Huffman {
    bool defined = false;
    int count[17] = {};
    int first_code[17] = {}, first_index[17] = {};
    uint8_t symbol[256] = {};

    bool build(const int counts[16]) {
        int n = 0, code = 0;
        for (int l = 1; l <= 16; l++) {
            count[l] = counts[l - 1];
            first_code[l] = code, first_index[l] = n;
            n += count[l], code += count[l];
            if (code > (1 << l)) return false; // more codes of this length than fit
            code <<= 1;
        }
        defined = n <= 256;
        return defined;
    }
};

// The entropy-coded segment as a bit stream: 0xFF00 is a data byte 0xFF, 0xFF followed by anything else ends the data
// (the marker is remembered and zero bits follow), fill bytes 0xFF 0xFF ... are skipped.
struct BitReader {
    const uint8_t *p, *end;
    uint32_t acc = 0;
    int n_bits = 0;
    int marker = -1;

    void reset_bits() { acc = 0, n_bits = 0, marker = -1; }
    void fill() {
        while (n_bits <= 24) {
            uint32_t b = 0;
            if (marker < 0 && p < end) {
                b = *p++;
                if (b == 0xFF) {
                    int c = p < end ? *p++ : 0;
                    while (c == 0xFF) c = p < end ? *p++ : 0;
                    if (c != 0) {
                        marker = c;
                        b = 0;
                    }
                }
            }
            acc |= b << (24 - n_bits);
            n_bits += 8;
        }
    }
    int bit() {
        if (n_bits < 1) fill();
        const int v = (int)(acc >> 31);
        acc <<= 1, n_bits--;
        return v;
    }
    int bits(int n) { // n in [0, 16]
        if (n == 0) return 0;
        if (n_bits < n) fill();
        const int v = (int)(acc >> (32 - n));
        acc <<= n, n_bits -= n;
        return v;
    }
    // T.81 F.2.2.1 EXTEND: an n-bit magnitude category value
    int receive_extend(int n) {
        if (n == 0) return 0;
        const int v = bits(n);
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }
    int decode(const Huffman &h) {
        if (n_bits < 16) fill();
        int code = 0;
        for (int l = 1; l <= 16; l++) {
            code = (code << 1) | (int)((acc >> (32 - l)) & 1u);
            if (code - h.first_code[l] < h.count[l] && code >= h.first_code[l]) {
                acc <<= l, n_bits -= l;
                return h.symbol[h.first_index[l] + code - h.first_code[l]];
            }
        }
        return -1;
    }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
    int px_w = 0, px_h = 0; // samples that carry image (T.81 A.1.1)
    int bw = 0, bh = 0;     // blocks per row / column, padded to whole MCUs
    int dc_pred = 0;
    std::vector<int16_t> coef; // progressive: bw * bh * 64, natural order
    std::vector<uint8_t> plane; // (bw * 8) x (bh * 8) decoded samples
};

// stb's inverse DCT of one dequantised block (natural order) into 8 rows of `stride` bytes. Integer arithmetic wraps
// like stb's `int` does (unsigned here to keep it defined).
inline int32_t fx(double x) { return (int32_t)(x * 4096 + 0.5); }
struct Idct1D {
    int32_t x0, x1, x2, x3, t0, t1, t2, t3;
    Idct1D(int32_t s0, int32_t s1, int32_t s2, int32_t s3, int32_t s4, int32_t s5, int32_t s6, int32_t s7) {
        static const int32_t c0 = fx(0.5411961f), c1 = fx(-1.847759065f), c2 = fx(0.765366865f), c3 = fx(1.175875602f), c4 = fx(0.298631336f),
                             c5 = fx(2.053119869f), c6 = fx(3.072711026f), c7 = fx(1.501321110f), c8 = fx(-0.899976223f), c9 = fx(-2.562915447f),
                             c10 = fx(-1.961570560f), c11 = fx(-0.390180644f);
        auto mul = [](int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); };
        auto add = [](int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); };
        auto sub = [](int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); };
        // even part
        int32_t p1 = mul(add(s2, s6), c0);
        const int32_t e2 = add(p1, mul(s6, c1)), e3 = add(p1, mul(s2, c2));
        const int32_t e0 = mul(add(s0, s4), 4096), e1 = mul(sub(s0, s4), 4096);
        x0 = add(e0, e3), x3 = sub(e0, e3), x1 = add(e1, e2), x2 = sub(e1, e2);
        // odd part
        const int32_t p3 = add(s7, s3), p4 = add(s5, s1), q1 = add(s7, s1), q2 = add(s5, s3);
        const int32_t p5 = mul(add(p3, p4), c3);
        const int32_t r1 = add(p5, mul(q1, c8)), r2 = add(p5, mul(q2, c9)), r3 = mul(p3, c10), r4 = mul(p4, c11);
        t3 = add(mul(s1, c7), add(r1, r4));
        t2 = add(mul(s3, c6), add(r2, r3));
        t1 = add(mul(s5, c5), add(r2, r4));
        t0 = add(mul(s7, c4), add(r1, r3));
    }
};
inline uint8_t clamp8(int32_t x) { return (uint8_t)(x < 0 ? 0 : (x > 255 ? 255 : x)); }

void idct_block(uint8_t *out, int stride, const int16_t d[64]) {
    int32_t v[64];
    auto add = [](int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); };
    auto sub = [](int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); };
    for (int c = 0; c < 8; c++) { // columns; two extra bits of precision are kept
        if (!(d[c + 8] | d[c + 16] | d[c + 24] | d[c + 32] | d[c + 40] | d[c + 48] | d[c + 56])) {
            const int32_t dc = d[c] * 4;
            for (int r = 0; r < 8; r++) v[c + 8 * r] = dc;
            continue;
        }
        Idct1D k(d[c], d[c + 8], d[c + 16], d[c + 24], d[c + 32], d[c + 40], d[c + 48], d[c + 56]);
        const int32_t x0 = add(k.x0, 512), x1 = add(k.x1, 512), x2 = add(k.x2, 512), x3 = add(k.x3, 512);
        v[c] = add(x0, k.t3) >> 10, v[c + 56] = sub(x0, k.t3) >> 10;
        v[c + 8] = add(x1, k.t2) >> 10, v[c + 48] = sub(x1, k.t2) >> 10;
        v[c + 16] = add(x2, k.t1) >> 10, v[c + 40] = sub(x2, k.t1) >> 10;
        v[c + 24] = add(x3, k.t0) >> 10, v[c + 32] = sub(x3, k.t0) >> 10;
    }
    for (int r = 0; r < 8; r++) { // rows: 12 + 2 + 3 bits to remove, rounded, level shift 128 folded in
        const int32_t *s = v + 8 * r;
        uint8_t *o = out + (size_t)r * stride;
        Idct1D k(s[0], s[1], s[2], s[3], s[4], s[5], s[6], s[7]);
        const int32_t bias = 65536 + (128 << 17);
        const int32_t x0 = add(k.x0, bias), x1 = add(k.x1, bias), x2 = add(k.x2, bias), x3 = add(k.x3, bias);
        o[0] = clamp8(add(x0, k.t3) >> 17), o[7] = clamp8(sub(x0, k.t3) >> 17);
        o[1] = clamp8(add(x1, k.t2) >> 17), o[6] = clamp8(sub(x1, k.t2) >> 17);
        o[2] = clamp8(add(x2, k.t1) >> 17), o[5] = clamp8(sub(x2, k.t1) >> 17);
        o[3] = clamp8(add(x3, k.t0) >> 17), o[4] = clamp8(sub(x3, k.t0) >> 17);
    }
}

struct Decoder {
    const std::string &path;
    const uint8_t *p, *end;
    int width = 0, height = 0, n_comp = 0;
    bool progressive = false, jfif = false;
    int adobe_transform = -1, n_rgb_ids = 0;
    int h_max = 1, v_max = 1, mcu_x = 0, mcu_y = 0;
    int restart_interval = 0;
    uint16_t quant[4][64] = {};
    Huffman dc_tab[4], ac_tab[4];
    Component comp[4];
    // current scan
    int scan_n = 0, order[4] = {}, ss = 0, se = 63, ah = 0, al = 0, eob_run = 0;

    Decoder(const std::string &path_, const uint8_t *b, const uint8_t *e) : path(path_), p(b), end(e) {}

    int get8() { return p < end ? *p++ : 0; }
    int get16() {
        const int a = get8();
        return (a << 8) | get8();
    }
    void skip(int n) { p = (end - p < n) ? end : p + n; }

    // a marker code, or -1 if the next two bytes are not one (fill bytes 0xFF are skipped)
    int next_marker() {
        if (get8() != 0xFF) return -1;
        int c = get8();
        while (c == 0xFF) c = get8();
        return c;
    }

    void tables_and_misc(int m) {
        if (m == 0xDD) { // DRI
            if (get16() != 4) fail(path, "bad DRI segment");
            restart_interval = get16();
        } else if (m == 0xDB) { // DQT: 8- or 16-bit entries in zigzag order
            int len = get16() - 2;
            while (len > 0) {
                const int q = get8(), wide = q >> 4, t = q & 15;
                if (wide > 1 || t > 3) fail(path, "bad quantisation table header");
                for (int i = 0; i < 64; i++) quant[t][kZigzag[i]] = (uint16_t)(wide ? get16() : get8());
                len -= wide ? 129 : 65;
            }
            if (len != 0) fail(path, "bad DQT segment");
        } else if (m == 0xC4) { // DHT
            int len = get16() - 2;
            while (len > 0) {
                const int q = get8(), cls = q >> 4, t = q & 15;
                if (cls > 1 || t > 3) fail(path, "bad Huffman table header");
                int counts[16], n = 0;
                for (int i = 0; i < 16; i++) n += counts[i] = get8();
                Huffman &h = cls ? ac_tab[t] : dc_tab[t];
                if (!h.build(counts)) fail(path, "bad Huffman code lengths");
                for (int i = 0; i < n; i++) h.symbol[i] = (uint8_t)get8();
                len -= 17 + n;
            }
            if (len != 0) fail(path, "bad DHT segment");
        } else if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE) { // APPn / COM
            int len = get16();
            if (len < 2) fail(path, "bad APP / COM segment");
            len -= 2;
            if (m == 0xE0 && len >= 5) {
                static const char tag[5] = {'J', 'F', 'I', 'F', 0};
                bool ok = true;
                for (int i = 0; i < 5; i++) ok &= get8() == (uint8_t)tag[i];
                len -= 5;
                if (ok) jfif = true;
            } else if (m == 0xEE && len >= 12) {
                static const char tag[6] = {'A', 'd', 'o', 'b', 'e', 0};
                bool ok = true;
                for (int i = 0; i < 6; i++) ok &= get8() == (uint8_t)tag[i];
                len -= 6;
                if (ok) {
                    skip(5); // version, flags0, flags1
                    adobe_transform = get8();
                    len -= 6;
                }
            }
            skip(len);
        } else {
            fail(path, m < 0 ? "marker expected" : "unsupported JPEG marker (arithmetic coding, lossless and hierarchical files are not read)");
        }
    }

    void frame_header() {
        const int len = get16();
        if (len < 11) fail(path, "bad SOF segment");
        if (get8() != 8) fail(path, "only 8-bit JPEG files are read");
        height = get16(), width = get16();
        if (height == 0 || width == 0) fail(path, "JPEG without a size in its frame header");
        n_comp = get8();
        if (n_comp != 1 && n_comp != 3 && n_comp != 4) fail(path, "bad JPEG component count");
        if (len != 8 + 3 * n_comp) fail(path, "bad SOF segment");
        if ((uint64_t)width * height > ((uint64_t)1 << 26)) fail(path, "image too large");
        static const char rgb[3] = {'R', 'G', 'B'};
        for (int i = 0; i < n_comp; i++) {
            Component &c = comp[i];
            c.id = get8();
            if (n_comp == 3 && c.id == rgb[i]) n_rgb_ids++;
            const int q = get8();
            c.h = q >> 4, c.v = q & 15, c.tq = get8();
            if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) fail(path, "bad JPEG component description");
            h_max = std::max(h_max, c.h), v_max = std::max(v_max, c.v);
        }
        mcu_x = (width + 8 * h_max - 1) / (8 * h_max), mcu_y = (height + 8 * v_max - 1) / (8 * v_max);
        for (int i = 0; i < n_comp; i++) {
            Component &c = comp[i];
            c.px_w = (width * c.h + h_max - 1) / h_max, c.px_h = (height * c.v + v_max - 1) / v_max;
            c.bw = mcu_x * c.h, c.bh = mcu_y * c.v;
            c.plane.assign((size_t)c.bw * 8 * c.bh * 8 + (size_t)width + 16, 0); // (+ slack: sampling factors that do not divide h_max read past a row, as in stb)
            if (progressive) c.coef.assign((size_t)c.bw * c.bh * 64, 0);
        }
    }

    void scan_header() {
        const int len = get16();
        scan_n = get8();
        if (scan_n < 1 || scan_n > 4 || scan_n > n_comp || len != 6 + 2 * scan_n) fail(path, "bad SOS segment");
        for (int i = 0; i < scan_n; i++) {
            const int id = get8(), q = get8();
            int which = 0;
            while (which < n_comp && comp[which].id != id) which++;
            if (which == n_comp || (q >> 4) > 3 || (q & 15) > 3) fail(path, "bad SOS component");
            comp[which].td = q >> 4, comp[which].ta = q & 15;
            order[i] = which;
        }
        ss = get8(), se = get8();
        const int a = get8();
        ah = a >> 4, al = a & 15;
        if (progressive) {
            if (ss > 63 || se > 63 || ss > se || ah > 13 || al > 13) fail(path, "bad progressive scan parameters");
        } else {
            if (ss != 0 || ah != 0 || al != 0) fail(path, "bad sequential scan parameters");
            se = 63;
        }
    }

    // ---- block decoding (T.81 F.2.2 sequential; G.1.2 progressive) ----
    void block_sequential(BitReader &br, Component &c, int16_t d[64]) {
        memset(d, 0, 64 * sizeof(int16_t));
        const uint16_t *q = quant[c.tq];
        const int t = br.decode(dc_tab[c.td]);
        if (t < 0 || t > 15) fail(path, "bad Huffman code");
        c.dc_pred = (int)((uint32_t)c.dc_pred + (uint32_t)br.receive_extend(t)); // (wraps on corrupt data instead of overflowing)
        d[0] = (int16_t)((uint32_t)c.dc_pred * q[0]);
        for (int k = 1; k < 64;) {
            const int rs = br.decode(ac_tab[c.ta]);
            if (rs < 0) fail(path, "bad Huffman code");
            const int r = rs >> 4, s = rs & 15;
            if (s == 0) {
                if (rs != 0xF0) break; // end of block
                k += 16;
            } else {
                k += r;
                const int z = kZigzag[std::min(k, 63)]; // (a corrupt run past the block: stb's table continues with 63s)
                k++;
                d[z] = (int16_t)(br.receive_extend(s) * q[z]);
            }
        }
    }
    void block_dc_progressive(BitReader &br, Component &c, int16_t *d) {
        if (se != 0) fail(path, "progressive scan mixes DC and AC coefficients");
        if (ah == 0) {
            memset(d, 0, 64 * sizeof(int16_t));
            const int t = br.decode(dc_tab[c.td]);
            if (t < 0 || t > 15) fail(path, "bad Huffman code");
            c.dc_pred = (int)((uint32_t)c.dc_pred + (uint32_t)br.receive_extend(t));
            d[0] = (int16_t)((uint32_t)c.dc_pred << al);
        } else if (br.bit()) {
            d[0] = (int16_t)(d[0] + (int16_t)(1 << al));
        }
    }
    void refine(BitReader &br, int16_t &coef, int16_t bit) { // correction bit of a coefficient with history (G.1.2.3)
        if (br.bit() && (coef & bit) == 0) coef = (int16_t)(coef > 0 ? coef + bit : coef - bit);
    }
    void block_ac_progressive(BitReader &br, Component &c, int16_t *d) {
        if (ss == 0) fail(path, "progressive scan mixes DC and AC coefficients");
        const Huffman &h = ac_tab[c.ta];
        if (ah == 0) { // first pass over this band
            if (eob_run) {
                eob_run--;
                return;
            }
            for (int k = ss; k <= se;) {
                const int rs = br.decode(h);
                if (rs < 0) fail(path, "bad Huffman code");
                const int r = rs >> 4, s = rs & 15;
                if (s == 0) {
                    if (r < 15) { // EOBn: this block and eob_run more end here
                        eob_run = (1 << r) - 1;
                        if (r) eob_run += br.bits(r);
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    d[kZigzag[std::min(k, 63)]] = (int16_t)(br.receive_extend(s) * (1 << al));
                    k++;
                }
            }
            return;
        }
        const int16_t bit = (int16_t)(1 << al); // refinement pass
        if (eob_run) {
            eob_run--;
            for (int k = ss; k <= se; k++) {
                int16_t &v = d[kZigzag[k]];
                if (v != 0) refine(br, v, bit);
            }
            return;
        }
        for (int k = ss; k <= se;) {
            const int rs = br.decode(h);
            if (rs < 0) fail(path, "bad Huffman code");
            int r = rs >> 4;
            const int s = rs & 15;
            int16_t value = 0;
            if (s == 0) {
                if (r < 15) {
                    eob_run = (1 << r) - 1;
                    if (r) eob_run += br.bits(r);
                    r = 64; // the rest of the band only receives correction bits
                }
            } else {
                if (s != 1) fail(path, "bad Huffman code");
                value = br.bit() ? bit : (int16_t)-bit;
            }
            while (k <= se) { // skip r zero-history coefficients, correcting the others on the way, then place the new one
                int16_t &v = d[kZigzag[k++]]; // (se <= 63)
                if (v != 0) {
                    refine(br, v, bit);
                } else {
                    if (r == 0) {
                        v = value;
                        break;
                    }
                    r--;
                }
            }
        }
    }

    // One scan. Returns the marker that ended it (-1: none seen).
    int scan() {
        BitReader br{p, end};
        int todo = restart_interval ? restart_interval : 0x7fffffff;
        auto restart = [&]() { // after every MCU; false: the data goes on without the restart marker it should have
            if (--todo > 0) return true;
            if (br.n_bits < 24) br.fill();
            if (br.marker < 0xD0 || br.marker > 0xD7) return false;
            br.reset_bits();
            for (Component &c : comp) c.dc_pred = 0;
            eob_run = 0;
            todo = restart_interval ? restart_interval : 0x7fffffff;
            return true;
        };
        for (Component &c : comp) c.dc_pred = 0;
        eob_run = 0;
        int16_t block[64];
        bool go_on = true;
        if (scan_n == 1) { // one component: its blocks in raster order, only those that carry image
            Component &c = comp[order[0]];
            const int w = (c.px_w + 7) >> 3, h = (c.px_h + 7) >> 3;
            for (int j = 0; j < h && go_on; j++)
                for (int i = 0; i < w && go_on; i++) {
                    if (!progressive) {
                        block_sequential(br, c, block);
                        idct_block(c.plane.data() + ((size_t)j * 8 * c.bw + i) * 8, c.bw * 8, block);
                    } else if (ss == 0) {
                        block_dc_progressive(br, c, c.coef.data() + 64 * ((size_t)j * c.bw + i));
                    } else {
                        block_ac_progressive(br, c, c.coef.data() + 64 * ((size_t)j * c.bw + i));
                    }
                    go_on = restart();
                }
        } else { // interleaved: h x v blocks of every component per MCU
            for (int j = 0; j < mcu_y && go_on; j++)
                for (int i = 0; i < mcu_x && go_on; i++) {
                    for (int k = 0; k < scan_n; k++) {
                        Component &c = comp[order[k]];
                        for (int y = 0; y < c.v; y++)
                            for (int x = 0; x < c.h; x++) {
                                const int bx = i * c.h + x, by = j * c.v + y;
                                if (!progressive) {
                                    block_sequential(br, c, block);
                                    idct_block(c.plane.data() + ((size_t)by * 8 * c.bw + bx) * 8, c.bw * 8, block);
                                } else {
                                    block_dc_progressive(br, c, c.coef.data() + 64 * ((size_t)by * c.bw + bx));
                                }
                            }
                    }
                    go_on = restart();
                }
        }
        p = br.p;
        return br.marker;
    }

    void finish_progressive() {
        for (int n = 0; n < n_comp; n++) {
            Component &c = comp[n];
            const int w = (c.px_w + 7) >> 3, h = (c.px_h + 7) >> 3;
            const uint16_t *q = quant[c.tq];
            for (int j = 0; j < h; j++)
                for (int i = 0; i < w; i++) {
                    int16_t *d = c.coef.data() + 64 * ((size_t)j * c.bw + i);
                    for (int k = 0; k < 64; k++) d[k] = (int16_t)(d[k] * q[k]);
                    idct_block(c.plane.data() + ((size_t)j * 8 * c.bw + i) * 8, c.bw * 8, d);
                }
        }
    }

    void decode() {
        if (next_marker() != 0xD8) fail(path, "not a JPEG file");
        int m = next_marker();
        while (m != 0xC0 && m != 0xC1 && m != 0xC2) {
            tables_and_misc(m);
            m = next_marker();
            while (m < 0) { // padding between segments
                if (p >= end) fail(path, "JPEG without a frame header");
                m = next_marker();
            }
        }
        progressive = m == 0xC2;
        frame_header();
        m = next_marker();
        while (m != 0xD9) {
            if (m == 0xDA) {
                scan_header();
                m = scan();
                if (m < 0) { // zero padding after the entropy-coded data
                    while (p < end)
                        if (get8() == 0xFF) {
                            m = get8();
                            break;
                        }
                    if (m < 0) fail(path, "JPEG data ends without an end marker");
                }
                continue;
            }
            if (m == 0xDC) { // DNL
                if (get16() != 4 || get16() != height) fail(path, "bad DNL segment");
            } else {
                tables_and_misc(m);
            }
            m = next_marker();
            if (m < 0 && p >= end) fail(path, "JPEG data ends without an end marker");
        }
        if (progressive) finish_progressive();
    }

    // ---- planes -> RGB rows ----
    static uint8_t blinn(uint8_t x, uint8_t y) {
        const unsigned t = x * y + 128;
        return (uint8_t)((t + (t >> 8)) >> 8);
    }
    static const uint8_t *upsample(uint8_t *out, const uint8_t *near, const uint8_t *far, int w, int hs, int vs) {
        if (hs == 1 && vs == 1) return near;
        if (hs == 1 && vs == 2) {
            for (int i = 0; i < w; i++) out[i] = (uint8_t)((3 * near[i] + far[i] + 2) >> 2);
        } else if (hs == 2 && vs == 1) {
            if (w == 1) {
                out[0] = out[1] = near[0];
                return out;
            }
            out[0] = near[0];
            out[1] = (uint8_t)((near[0] * 3 + near[1] + 2) >> 2);
            for (int i = 1; i < w - 1; i++) {
                const int n = 3 * near[i] + 2;
                out[2 * i] = (uint8_t)((n + near[i - 1]) >> 2);
                out[2 * i + 1] = (uint8_t)((n + near[i + 1]) >> 2);
            }
            out[2 * (w - 1)] = (uint8_t)((near[w - 2] * 3 + near[w - 1] + 2) >> 2);
            out[2 * (w - 1) + 1] = near[w - 1];
        } else if (hs == 2 && vs == 2) {
            if (w == 1) {
                out[0] = out[1] = (uint8_t)((3 * near[0] + far[0] + 2) >> 2);
                return out;
            }
            int t1 = 3 * near[0] + far[0];
            out[0] = (uint8_t)((t1 + 2) >> 2);
            for (int i = 1; i < w; i++) {
                const int t0 = t1;
                t1 = 3 * near[i] + far[i];
                out[2 * i - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
                out[2 * i] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
            }
            out[2 * w - 1] = (uint8_t)((t1 + 2) >> 2);
        } else {
            for (int i = 0; i < w; i++)
                for (int j = 0; j < hs; j++) out[i * hs + j] = near[i];
        }
        return out;
    }
    static void ycc_to_rgb(uint8_t *out, uint8_t y, uint8_t cb_, uint8_t cr_) {
        static const int32_t k_r = fx(1.40200f) << 8, k_g_cr = fx(0.71414f) << 8, k_g_cb = fx(0.34414f) << 8, k_b = fx(1.77200f) << 8;
        const int32_t yf = (y << 20) + (1 << 19), cr = cr_ - 128, cb = cb_ - 128;
        const int32_t r = (yf + cr * k_r) >> 20;
        const int32_t g = (int32_t)(yf + cr * -k_g_cr + (int32_t)((uint32_t)(cb * -k_g_cb) & 0xffff0000u)) >> 20;
        const int32_t b = (yf + cb * k_b) >> 20;
        out[0] = clamp8(r), out[1] = clamp8(g), out[2] = clamp8(b);
    }

    void to_rgb(std::vector<uint8_t> *rgb) {
        rgb->assign((size_t)width * height * 3, 0);
        const bool is_rgb = n_comp == 3 && (n_rgb_ids == 3 || (adobe_transform == 0 && !jfif));
        struct Row {
            int hs, vs, ystep, w_lores, ypos;
            const uint8_t *line0, *line1;
            std::vector<uint8_t> buf;
        } row[4];
        for (int k = 0; k < n_comp; k++) {
            Row &r = row[k];
            r.hs = h_max / comp[k].h, r.vs = v_max / comp[k].v;
            r.ystep = r.vs >> 1, r.w_lores = (width + r.hs - 1) / r.hs, r.ypos = 0;
            r.line0 = r.line1 = comp[k].plane.data();
            r.buf.assign((size_t)width + 3 + 8, 0);
        }
        const uint8_t *line[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int j = 0; j < height; j++) {
            for (int k = 0; k < n_comp; k++) {
                Row &r = row[k];
                const bool bottom = r.ystep >= (r.vs >> 1); // the nearer source row is the lower one
                line[k] = upsample(r.buf.data(), bottom ? r.line1 : r.line0, bottom ? r.line0 : r.line1, r.w_lores, r.hs, r.vs);
                if (++r.ystep >= r.vs) {
                    r.ystep = 0;
                    r.line0 = r.line1;
                    if (++r.ypos < comp[k].px_h) r.line1 += comp[k].bw * 8;
                }
            }
            uint8_t *out = rgb->data() + (size_t)j * width * 3;
            for (int i = 0; i < width; i++, out += 3) {
                if (n_comp == 1) {
                    out[0] = out[1] = out[2] = line[0][i];
                } else if (n_comp == 3 && is_rgb) {
                    out[0] = line[0][i], out[1] = line[1][i], out[2] = line[2][i];
                } else if (n_comp == 3) {
                    ycc_to_rgb(out, line[0][i], line[1][i], line[2][i]);
                } else if (adobe_transform == 0) { // CMYK
                    const uint8_t m = line[3][i];
                    out[0] = blinn(line[0][i], m), out[1] = blinn(line[1][i], m), out[2] = blinn(line[2][i], m);
                } else if (adobe_transform == 2) { // YCCK
                    const uint8_t m = line[3][i];
                    ycc_to_rgb(out, line[0][i], line[1][i], line[2][i]);
                    out[0] = blinn((uint8_t)(255 - out[0]), m), out[1] = blinn((uint8_t)(255 - out[1]), m), out[2] = blinn((uint8_t)(255 - out[2]), m);
                } else { // four components without a statement about them: the first three as YCbCr
                    ycc_to_rgb(out, line[0][i], line[1][i], line[2][i]);
                }
            }
        }
    }
};

} // namespace

bool looks_like_jpeg(const uint8_t *bytes, size_t n) { return n >= 3 && bytes[0] == 0xFF && bytes[1] == 0xD8 && bytes[2] == 0xFF; }

void decode_jpeg_rgb8(const std::string &name, const uint8_t *bytes, size_t n, int *width, int *height, std::vector<uint8_t> *rgb) {
    Decoder d(name, bytes, bytes + n);
    d.decode();
    d.to_rgb(rgb);
    *width = d.width, *height = d.height;
}

void read_image_rgb8(const std::string &path, int *width, int *height, std::vector<uint8_t> *rgb) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error(path + ": cannot open image file");
    uint8_t head[3] = {0, 0, 0};
    const size_t got = fread(head, 1, 3, f);
    if (!looks_like_jpeg(head, got)) {
        fclose(f);
        read_png_rgb8(path, width, height, rgb);
        return;
    }
    std::vector<uint8_t> bytes(head, head + got);
    uint8_t buf[1 << 16];
    for (size_t n; (n = fread(buf, 1, sizeof buf, f)) > 0;) {
        bytes.insert(bytes.end(), buf, buf + n);
        if (bytes.size() > ((size_t)1 << 30)) {
            fclose(f);
            throw std::runtime_error(path + ": image file too large");
        }
    }
    fclose(f);
    decode_jpeg_rgb8(path, bytes.data(), bytes.size(), width, height, rgb);
}

} // namespace akb_host
