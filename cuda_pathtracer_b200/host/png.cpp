// png.cpp -- PNG decoder for TEXTURE_FILE images (zlib inflate + the five scan-line filters + Adam7).
//
// The reference loads textures with stbi_load(filename, &w, &h, &components, 0) (sceneStructs.h:170-175;
// the vendored stb_image.h is v2.06).  PNG is lossless, so "the same bytes as stb_image" is a matter of
// its OUTPUT conventions, restated here from that file:
//   * 1/2/4/8 bits per channel only -- 16-bit PNGs are a load error there (:4328) and here;
//   * components reported and stored: grey 1, grey+alpha 2, RGB 3, RGBA 4; a palette image becomes RGB, or
//     RGBA when it has a tRNS chunk (:4370-4374, :4422-4427); a grey / RGB image with a tRNS colour key gains
//     an alpha channel that is 0 where the pixel equals the key and 255 elsewhere (:4179-4201, :4414-4418);
//   * 1/2/4-bit grey samples are scaled to 0..255 (x 0xff / 0x55 / 0x11), palette indices are not (:4069),
//     and the tRNS key of such an image is scaled the same way before the comparison (:4378);
//   * ancillary chunks are skipped, an unknown CRITICAL chunk is an error (:4441-4453); chunk CRCs and the
//     zlib Adler-32 are not checked (stb does not either).
// Only Texture::get_color's 3-component case ever reaches a pixel (sceneStructs.h:180-187), but the loader dump
// of the tests compares every byte of every texture with what the reference's loader produced.
#include "mesh.hpp"

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200host {
namespace {

struct PngError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// ---- inflate (RFC 1951) ------------------------------------------------------------------------------------
struct Huff {
    uint16_t count[16] = {0};
    uint16_t symbol[288] = {0};

    void build(const uint8_t* lengths, int n)
    {
        std::memset(count, 0, sizeof(count));
        for (int i = 0; i < n; ++i) count[lengths[i]]++;
        count[0] = 0;
        int left = 1;
        for (int l = 1; l < 16; ++l) {
            left = (left << 1) - count[l];
            if (left < 0) throw PngError("bad code lengths");
        }
        uint16_t offs[16];
        offs[1] = 0;
        for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
        for (int i = 0; i < n; ++i)
            if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
};

struct Inflater {
    const uint8_t* in;
    size_t n, at = 0;
    uint32_t acc = 0;
    int bits = 0;
    std::vector<uint8_t> out;
    size_t cap = ~(size_t)0;          // the image needs no more output than this: a small file must not expand without bound

    void grew() { if (out.size() > cap) throw PngError("too much pixel data"); }
    int bit()
    {
        if (bits == 0) {
            if (at >= n) throw PngError("zlib data truncated");
            acc = in[at++];
            bits = 8;
        }
        const int b = (int)(acc & 1u);
        acc >>= 1;
        --bits;
        return b;
    }
    int take(int k)
    {
        int v = 0;
        for (int i = 0; i < k; ++i) v |= bit() << i;
        return v;
    }
    int decode(const Huff& h)
    {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l < 16; ++l) {
            code |= bit();
            const int c = h.count[l];
            if (code - c < first) return h.symbol[index + (code - first)];
            index += c;
            first += c;
            first <<= 1;
            code <<= 1;
        }
        throw PngError("bad huffman code");
    }
    void block(const Huff& lit, const Huff& dist)
    {
        static const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
        static const uint8_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
        static const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
        static const uint8_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
        for (;;) {
            const int s = decode(lit);
            if (s < 256) { out.push_back((uint8_t)s); grew(); continue; }
            if (s == 256) return;
            if (s > 285) throw PngError("bad length code");
            const int len = lbase[s - 257] + take(lext[s - 257]);
            const int ds = decode(dist);
            if (ds > 29) throw PngError("bad distance code");
            const size_t d = (size_t)dbase[ds] + (size_t)take(dext[ds]);
            if (d > out.size()) throw PngError("bad distance");
            const size_t from = out.size() - d;
            for (int i = 0; i < len; ++i) out.push_back(out[from + (size_t)i]);
            grew();
        }
    }
    void run()
    {
        if (n < 2) throw PngError("zlib header missing");
        const int cmf = in[0], flg = in[1];
        if ((cmf * 256 + flg) % 31 != 0 || (cmf & 15) != 8 || (flg & 32)) throw PngError("bad zlib header");
        at = 2;
        int last;
        do {
            last = bit();
            const int type = take(2);
            if (type == 0) {
                bits = 0;
                if (at + 4 > n) throw PngError("zlib data truncated");
                const unsigned len = in[at] | (in[at + 1] << 8), nlen = in[at + 2] | (in[at + 3] << 8);
                at += 4;
                if ((len ^ 0xffffu) != nlen || at + len > n) throw PngError("bad stored block");
                out.insert(out.end(), in + at, in + at + len);
                grew();
                at += len;
            } else if (type == 1) {
                uint8_t l[288];
                for (int i = 0; i < 288; ++i) l[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
                uint8_t dl[30];
                std::memset(dl, 5, sizeof(dl));
                Huff lit, dist;
                lit.build(l, 288);
                dist.build(dl, 30);
                block(lit, dist);
            } else if (type == 2) {
                static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                const int hlit = take(5) + 257, hdist = take(5) + 1, hclen = take(4) + 4;
                if (hlit > 286 || hdist > 30) throw PngError("bad code counts");
                uint8_t cl[19] = {0};
                for (int i = 0; i < hclen; ++i) cl[order[i]] = (uint8_t)take(3);
                Huff clh;
                clh.build(cl, 19);
                uint8_t l[320] = {0};
                int i = 0;
                while (i < hlit + hdist) {
                    const int s = decode(clh);
                    if (s < 16) { l[i++] = (uint8_t)s; continue; }
                    int rep, val = 0;
                    if (s == 16) { if (i == 0) throw PngError("bad repeat"); val = l[i - 1]; rep = 3 + take(2); }
                    else if (s == 17) rep = 3 + take(3);
                    else rep = 11 + take(7);
                    if (i + rep > hlit + hdist) throw PngError("bad repeat");
                    while (rep--) l[i++] = (uint8_t)val;
                }
                Huff lit, dist;
                lit.build(l, hlit);
                dist.build(l + hlit, hdist);
                block(lit, dist);
            } else {
                throw PngError("bad block type");
            }
        } while (!last);
    }
};

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline int paeth(int a, int b, int c)
{
    const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// one (sub-)image of w x h pixels out of `raw` (filter byte + packed samples per row) -> unpacked samples,
// `chan` per pixel, grey samples of less than 8 bits scaled when `scale_grey`.  Returns the bytes consumed.
size_t unfilter_pass(const uint8_t* raw, size_t raw_len, int w, int h, int chan, int depth, bool scale_grey, std::vector<uint8_t>& px)
{
    const size_t row_bytes = ((size_t)w * chan * depth + 7) / 8;
    const int bpp = depth < 8 ? 1 : chan;            // the filters work on whole bytes
    if ((row_bytes + 1) * (size_t)h > raw_len) throw PngError("not enough pixels");
    std::vector<uint8_t> cur(row_bytes), prev(row_bytes, 0);
    px.assign((size_t)w * h * chan, 0);
    const int scale = !scale_grey ? 1 : depth == 1 ? 0xff : depth == 2 ? 0x55 : depth == 4 ? 0x11 : 1;
    for (int y = 0; y < h; ++y) {
        const uint8_t* line = raw + (row_bytes + 1) * (size_t)y;
        const int filter = line[0];
        if (filter > 4) throw PngError("invalid filter");
        for (size_t i = 0; i < row_bytes; ++i) {
            const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= (size_t)bpp ? prev[i - bpp] : 0;
            const int x = line[1 + i];
            int v = x;
            if (filter == 1) v = x + a;
            else if (filter == 2) v = x + b;
            else if (filter == 3) v = x + ((a + b) >> 1);
            else if (filter == 4) v = x + paeth(a, b, c);
            cur[i] = (uint8_t)v;
        }
        uint8_t* o = px.data() + (size_t)y * w * chan;
        const int n = w * chan;
        if (depth == 8) {
            std::memcpy(o, cur.data(), (size_t)n);
        } else {
            const int per = 8 / depth, mask = (1 << depth) - 1;
            for (int k = 0; k < n; ++k) o[k] = (uint8_t)(scale * ((cur[(size_t)(k / per)] >> (8 - depth * (k % per + 1))) & mask));
        }
        prev.swap(cur);
    }
    return (row_bytes + 1) * (size_t)h;
}

}  // namespace

bool looks_like_png(const unsigned char* d, size_t n)
{
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    return n >= 8 && std::memcmp(d, sig, 8) == 0;
}

TextureImage decode_png(const unsigned char* data, size_t size, const std::string& name)
{
    try {
        size_t at = 8;
        bool have_ihdr = false, have_trans = false, done = false;
        int w = 0, h = 0, depth = 0, color = 0, interlace = 0, img_n = 0, pal_n = 0, pal_len = 0;
        uint8_t palette[1024], tc[3] = {0, 0, 0};
        std::vector<uint8_t> idat;
        while (!done) {
            if (at + 8 > size) throw PngError("truncated");
            const uint32_t len = be32(data + at);
            const uint8_t* type = data + at + 4;
            const uint8_t* body = data + at + 8;
            if (len > size || at + 12 + (size_t)len > size) throw PngError("truncated chunk");
            auto is = [&](const char* t) { return std::memcmp(type, t, 4) == 0; };
            if (is("IHDR")) {
                if (have_ihdr || len != 13) throw PngError("bad IHDR");
                have_ihdr = true;
                const uint32_t uw = be32(body), uh = be32(body + 4);
                if (uw == 0 || uh == 0 || uw > (1u << 24) || uh > (1u << 24)) throw PngError("bad size");
                w = (int)uw; h = (int)uh;
                depth = body[8]; color = body[9];
                if (depth != 1 && depth != 2 && depth != 4 && depth != 8) throw PngError("1/2/4/8-bit only");
                if (color > 6 || (color != 3 && (color & 1))) throw PngError("bad colour type");
                if (body[10] || body[11] || body[12] > 1) throw PngError("bad compression / filter / interlace method");
                interlace = body[12];
                if (color == 3) { pal_n = 3; img_n = 1; }
                else img_n = ((color & 2) ? 3 : 1) + ((color & 4) ? 1 : 0);
                if ((1u << 30) / uw / 4 < uh) throw PngError("too large");
            } else if (!have_ihdr) {
                throw PngError("first chunk is not IHDR");
            } else if (is("PLTE")) {
                if (len > 768 || len % 3) throw PngError("invalid PLTE");
                pal_len = (int)len / 3;
                for (int i = 0; i < pal_len; ++i) {
                    palette[4 * i] = body[3 * i]; palette[4 * i + 1] = body[3 * i + 1]; palette[4 * i + 2] = body[3 * i + 2];
                    palette[4 * i + 3] = 255;
                }
            } else if (is("tRNS")) {
                if (!idat.empty()) throw PngError("tRNS after IDAT");
                if (pal_n) {
                    if (pal_len == 0 || (int)len > pal_len) throw PngError("bad tRNS");
                    pal_n = 4;
                    for (uint32_t i = 0; i < len; ++i) palette[4 * i + 3] = body[i];
                } else {
                    if (!(img_n & 1) || len != (uint32_t)img_n * 2) throw PngError("bad tRNS");
                    have_trans = true;
                    const int scale = depth == 1 ? 0xff : depth == 2 ? 0x55 : depth == 4 ? 0x11 : 1;
                    for (int k = 0; k < img_n; ++k) tc[k] = (uint8_t)((uint8_t)(body[2 * k + 1]) * scale);
                }
            } else if (is("IDAT")) {
                if (pal_n && !pal_len) throw PngError("no PLTE");
                idat.insert(idat.end(), body, body + len);
            } else if (is("IEND")) {
                done = true;
            } else if (!(type[0] & 32)) {
                throw PngError("unknown critical chunk");
            }
            at += 12 + (size_t)len;
        }
        if (idat.empty()) throw PngError("no IDAT");
        Inflater z;
        z.in = idat.data();
        z.n = idat.size();
        z.out.reserve(((size_t)w * img_n * depth / 8 + 2) * (size_t)h);
        // upper bound of what any layout of this image inflates to (Adam7: at most 7 extra filter bytes per row of
        // the 8-row pattern and 7 partial rows): more is a corrupt or hostile file
        z.cap = (((size_t)w * img_n * depth + 7) / 8 + 8) * ((size_t)h + 8) + 64;
        z.run();

        // samples of the file: img_n per pixel, unpacked (grey scaled, palette indices not)
        std::vector<uint8_t> samples((size_t)w * h * img_n, 0);
        if (!interlace) {
            const size_t expect = (((size_t)w * img_n * depth + 7) / 8 + 1) * (size_t)h;
            if (z.out.size() != expect) throw PngError("not enough pixels");
            unfilter_pass(z.out.data(), z.out.size(), w, h, img_n, depth, color == 0, samples);
        } else {
            static const int xo[7] = {0, 4, 0, 2, 0, 1, 0}, yo[7] = {0, 0, 4, 0, 2, 0, 1};
            static const int xs[7] = {8, 8, 4, 4, 2, 2, 1}, ys[7] = {8, 8, 8, 4, 4, 2, 2};
            size_t used = 0;
            std::vector<uint8_t> sub;
            for (int p = 0; p < 7; ++p) {
                const int pw = (w - xo[p] + xs[p] - 1) / xs[p], ph = (h - yo[p] + ys[p] - 1) / ys[p];
                if (pw <= 0 || ph <= 0) continue;
                used += unfilter_pass(z.out.data() + used, z.out.size() - used, pw, ph, img_n, depth, color == 0, sub);
                for (int y = 0; y < ph; ++y)
                    for (int x = 0; x < pw; ++x)
                        std::memcpy(&samples[((size_t)(y * ys[p] + yo[p]) * w + (size_t)(x * xs[p] + xo[p])) * img_n],
                                    &sub[((size_t)y * pw + x) * img_n], (size_t)img_n);
            }
        }

        TextureImage t;
        t.width = w;
        t.height = h;
        const size_t n_px = (size_t)w * h;
        if (pal_n) {
            t.components = pal_n;
            t.pixels.resize(n_px * pal_n);
            for (size_t i = 0; i < n_px; ++i) {
                const int idx = samples[i];        // (an index beyond the palette reads the zero-initialised / stale entry in stb; 0 here)
                const uint8_t* e = idx < 256 ? palette + 4 * idx : palette;
                for (int k = 0; k < pal_n; ++k) t.pixels[i * pal_n + k] = idx < pal_len ? e[k] : (k == 3 ? 255 : 0);
            }
        } else if (have_trans) {
            const int out_n = img_n + 1;
            t.components = out_n;
            t.pixels.resize(n_px * out_n);
            for (size_t i = 0; i < n_px; ++i) {
                bool key = true;
                for (int k = 0; k < img_n; ++k) {
                    t.pixels[i * out_n + k] = samples[i * img_n + k];
                    key = key && samples[i * img_n + k] == tc[k];
                }
                t.pixels[i * out_n + img_n] = key ? 0 : 255;
            }
        } else {
            t.components = img_n;
            t.pixels = std::move(samples);
        }
        return t;
    } catch (const PngError& e) {
        throw std::runtime_error("Texture load error! (" + name + ": PNG: " + e.what() + ")");
    }
}

}  // namespace b200host
