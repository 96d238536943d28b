// jpeg.cpp -- baseline (SOF0/SOF1, Huffman, 8-bit) JPEG decoder for TEXTURE_FILE images.
//
// The reference loads textures with stbi_load(filename, &w, &h, &components, 0)
// (/root/reference/path_tracer/src/sceneStructs.h:170-175) and its own scene, scenes/room.json, names
// two JPEGs (Textures/chair.jpg, Textures/wallpaper.jpg: baseline, 4:2:0).  A texel that differs by
// one level changes the colour of every path that touches it, so the decoder must return the same
// BYTES as stb_image 2.x does, not merely a visually equal picture.  JPEG decoding is only specified
// up to an accuracy bound; the choices that pin the bytes are restated here from the vendored
// external/include/stb_image.h (each cited below):
//   * the integer inverse DCT: jidctint "ISLOW" butterfly with 12-bit constants, column pass keeping
//     two extra bits (+512 >> 10), row pass removing 17 bits with the +128 level shift folded into
//     the rounding term (stb_image.h:1906-2001);
//   * de-quantised coefficients are truncated to 16 bits before the transform (:1712, :1731, :1743);
//   * chroma upsampling: "fancy" triangle filters -- 3:1 vertically, 3:1 horizontally, 9:3:3:1 for
//     2x2 -- with the edge samples replicated and the row pair chosen by the ystep rule of
//     load_jpeg_image (:2849-2911, :3327-3368); other ratios replicate samples (:2996-3006);
//   * YCbCr -> RGB in 20.12 fixed point with the constants rounded to 12 bits first and the Cb term
//     of green masked to its upper 16 bits (the "same result in SIMD and scalar" variant, :3071-3094).
// The SSE2 kernels stb selects on x86-64 are bit-identical to these scalar forms by design (:2004-2006,
// :3069-3070).  tests/test_host_cli.py checks the decoded texels of both reference textures byte for
// byte against what the reference's own loader (stb_image, via oracle/ref_scene_dump) produces.
//
// Progressive files (SOF2: spectral selection + successive approximation) are decoded as well: the scans fill
// a coefficient store per component (stb_image.h:1744-1890, restated below: first and refinement passes of the
// DC and the AC band, end-of-band runs), which is de-quantised -- with the same truncation to 16 bits -- and
// transformed when the last scan is in (:2527-2545).  Arithmetic-coded and 12-bit JPEGs are a load error, as in stb.
#include "mesh.hpp"

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200host {
namespace {

struct JpegError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// zig-zag scan position -> natural (row-major) coefficient index
const unsigned char kNatural[64] = {
    0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct HuffTable {
    // canonical code: for each length l, codes [first[l], first[l] + count[l]) map to symbols[offset[l] ...]
    int count[17] = {0};
    int first[17] = {0};
    int offset[17] = {0};
    unsigned char symbols[256] = {0};
    bool present = false;

    void build(const unsigned char counts[16], const unsigned char* syms, int n_syms)
    {
        int code = 0, at = 0;
        for (int l = 1; l <= 16; ++l) {
            count[l] = counts[l - 1];
            first[l] = code;
            offset[l] = at;
            code += count[l];
            if (code > (1 << l)) throw JpegError("bad code lengths");
            at += count[l];
            code <<= 1;
        }
        if (at != n_syms) throw JpegError("bad huffman table");
        std::memcpy(symbols, syms, (size_t)n_syms);
        present = true;
    }
};

// Entropy-coded segment reader: MSB-first bits, 0xFF00 -> 0xFF, stops at any other marker and then
// supplies zero bits (as stb does once it has seen a marker, stb_image.h:1567-1585).
struct BitReader {
    const unsigned char* p;
    const unsigned char* end;
    uint32_t acc = 0;
    int bits = 0;
    int marker = -1;          // marker met inside the entropy-coded data (without the 0xFF), or -1

    BitReader(const unsigned char* b, const unsigned char* e) : p(b), end(e) {}

    int next_byte()
    {
        if (marker >= 0 || p >= end) return 0;
        int b = *p++;
        if (b == 0xFF) {
            int c = p < end ? *p++ : 0;
            while (c == 0xFF) c = p < end ? *p++ : 0;      // fill bytes
            if (c != 0) { marker = c; return 0; }
        }
        return b;
    }
    int bit()
    {
        if (bits == 0) { acc = (uint32_t)next_byte(); bits = 8; }
        --bits;
        return (int)((acc >> bits) & 1u);
    }
    int receive(int n)
    {
        int v = 0;
        for (int i = 0; i < n; ++i) v = (v << 1) | bit();
        return v;
    }
    // JPEG "EXTEND": an n-bit magnitude whose leading bit 0 means negative
    int receive_extend(int n)
    {
        if (n == 0) return 0;
        const int v = receive(n);
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }
    int decode(const HuffTable& h)
    {
        int code = 0;
        for (int l = 1; l <= 16; ++l) {
            code = (code << 1) | bit();
            if (code - h.first[l] < h.count[l]) return h.symbols[h.offset[l] + code - h.first[l]];
        }
        throw JpegError("bad huffman code");
    }
    void reset() { acc = 0; bits = 0; }
    // after a restart interval: consume the RSTn marker (already latched or the next two bytes)
    void take_restart()
    {
        reset();
        if (marker < 0 && p + 1 < end && p[0] == 0xFF && p[1] >= 0xD0 && p[1] <= 0xD7) { p += 2; return; }
        if (marker >= 0xD0 && marker <= 0xD7) marker = -1;
    }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
    int dc_pred = 0;
    int x = 0, y = 0;         // samples actually covered by the image
    int w2 = 0, h2 = 0;       // allocated (whole MCUs)
    std::vector<unsigned char> data;
    std::vector<short> coef;  // progressive: 64 coefficients per block, (w2 / 8) blocks per row
};

inline unsigned char clamp_u8(int x) { return (unsigned char)(x < 0 ? 0 : (x > 255 ? 255 : x)); }

constexpr int fix12(double x) { return (int)(x * 4096 + 0.5); }

// 32-bit integer with two's-complement wrap-around: what the transform's arithmetic does on every machine this runs
// on, spelled out so that a corrupt file (coefficients far outside what a quantiser can produce) wraps instead of
// running into signed-overflow undefined behaviour.  Valid files never wrap.
struct W {
    int v;
};
inline W operator+(W a, W b) { return W{(int)((unsigned)a.v + (unsigned)b.v)}; }
inline W operator-(W a, W b) { return W{(int)((unsigned)a.v - (unsigned)b.v)}; }
inline W operator*(W a, W b) { return W{(int)((unsigned)a.v * (unsigned)b.v)}; }
inline W K(double x) { return W{fix12(x)}; }

// One 8-point pass of the ISLOW butterfly.  Returns the even part in e[0..3] and the odd part in
// o[0..3] (both scaled by 2^12) so that out[k] = e[k] + o[3-k], out[7-k] = e[k] - o[3-k].
inline void idct_1d(W s0, W s1, W s2, W s3, W s4, W s5, W s6, W s7, W e[4], W o[4])
{
    // even part
    const W z = (s2 + s6) * K(0.5411961f);
    const W a2 = z + s6 * K(-1.847759065f);
    const W a3 = z + s2 * K(0.765366865f);
    const W a0 = (s0 + s4) * W{4096};
    const W a1 = (s0 - s4) * W{4096};
    e[0] = a0 + a3;
    e[3] = a0 - a3;
    e[1] = a1 + a2;
    e[2] = a1 - a2;
    // odd part
    W b0 = s7, b1 = s5, b2 = s3, b3 = s1;
    W p3 = b0 + b2, p4 = b1 + b3, p1 = b0 + b3, p2 = b1 + b2;
    const W p5 = (p3 + p4) * K(1.175875602f);
    b0 = b0 * K(0.298631336f);
    b1 = b1 * K(2.053119869f);
    b2 = b2 * K(3.072711026f);
    b3 = b3 * K(1.501321110f);
    p1 = p5 + p1 * K(-0.899976223f);
    p2 = p5 + p2 * K(-2.562915447f);
    p3 = p3 * K(-1.961570560f);
    p4 = p4 * K(-0.390180644f);
    o[3] = b3 + p1 + p4;
    o[2] = b2 + p2 + p3;
    o[1] = b1 + p2 + p4;
    o[0] = b0 + p1 + p3;
}

void idct_block(unsigned char* out, int stride, const short d[64])
{
    W v[64], e[4], o[4];
    for (int c = 0; c < 8; ++c) {
        if (!d[c + 8] && !d[c + 16] && !d[c + 24] && !d[c + 32] && !d[c + 40] && !d[c + 48] && !d[c + 56]) {
            const W dc = W{d[c]} * W{4};             // DC-only column: same scale as the general case
            for (int r = 0; r < 8; ++r) v[c + 8 * r] = dc;
            continue;
        }
        idct_1d(W{d[c]}, W{d[c + 8]}, W{d[c + 16]}, W{d[c + 24]}, W{d[c + 32]}, W{d[c + 40]}, W{d[c + 48]}, W{d[c + 56]}, e, o);
        for (int k = 0; k < 4; ++k) {
            const W x = e[k] + W{512};               // keep two extra bits: >> 10 instead of >> 12
            v[c + 8 * k] = W{(x + o[3 - k]).v >> 10};
            v[c + 8 * (7 - k)] = W{(x - o[3 - k]).v >> 10};
        }
    }
    for (int r = 0; r < 8; ++r) {
        const W* s = v + 8 * r;
        idct_1d(s[0], s[1], s[2], s[3], s[4], s[5], s[6], s[7], e, o);
        unsigned char* q = out + (size_t)r * stride;
        for (int k = 0; k < 4; ++k) {
            const W x = e[k] + W{65536 + (128 << 17)};   // rounding + level shift, removed by >> 17
            q[k] = clamp_u8((x + o[3 - k]).v >> 17);
            q[7 - k] = clamp_u8((x - o[3 - k]).v >> 17);
        }
    }
}

// ---- chroma upsampling (one output row from the nearer and the farther input row) ----
inline unsigned char div4(int x) { return (unsigned char)(x >> 2); }
inline unsigned char div16(int x) { return (unsigned char)(x >> 4); }

const unsigned char* upsample_row(unsigned char* out, const unsigned char* near_row, const unsigned char* far_row,
                                  int w, int hs, int vs)
{
    if (hs == 1 && vs == 1) return near_row;
    if (hs == 1 && vs == 2) {
        for (int i = 0; i < w; ++i) out[i] = div4(3 * near_row[i] + far_row[i] + 2);
        return out;
    }
    if (hs == 2 && vs == 1) {
        const unsigned char* in = near_row;
        if (w == 1) { out[0] = out[1] = in[0]; return out; }
        out[0] = in[0];
        out[1] = div4(in[0] * 3 + in[1] + 2);
        for (int i = 1; i < w - 1; ++i) {
            const int n = 3 * in[i] + 2;
            out[2 * i] = div4(n + in[i - 1]);
            out[2 * i + 1] = div4(n + in[i + 1]);
        }
        out[2 * (w - 1)] = div4(in[w - 2] * 3 + in[w - 1] + 2);
        out[2 * (w - 1) + 1] = in[w - 1];
        return out;
    }
    if (hs == 2 && vs == 2) {
        if (w == 1) { out[0] = out[1] = div4(3 * near_row[0] + far_row[0] + 2); return out; }
        int cur = 3 * near_row[0] + far_row[0];
        out[0] = div4(cur + 2);
        for (int i = 1; i < w; ++i) {
            const int prev = cur;
            cur = 3 * near_row[i] + far_row[i];
            out[2 * i - 1] = div16(3 * prev + cur + 8);
            out[2 * i] = div16(3 * cur + prev + 8);
        }
        out[2 * w - 1] = div4(cur + 2);
        return out;
    }
    for (int i = 0; i < w; ++i)                      // any other ratio: nearest neighbour
        for (int j = 0; j < hs; ++j) out[i * hs + j] = near_row[i];
    return out;
}

constexpr int fix20(float x) { return ((int)(x * 4096.0f + 0.5f)) << 8; }

inline void ycc_to_rgb(unsigned char* out, int y, int cb, int cr)
{
    const int yf = (y << 20) + (1 << 19);
    cr -= 128;
    cb -= 128;
    int r = yf + cr * fix20(1.40200f);
    int g = yf + (cr * -fix20(0.71414f)) + (int)((unsigned)(cb * -fix20(0.34414f)) & 0xffff0000u);
    int b = yf + cb * fix20(1.77200f);
    out[0] = clamp_u8(r >> 20);
    out[1] = clamp_u8(g >> 20);
    out[2] = clamp_u8(b >> 20);
}

struct Decoder {
    const unsigned char* base;
    size_t size;
    size_t at = 0;
    int width = 0, height = 0, n_comp = 0;
    int h_max = 1, v_max = 1, mcus_x = 0, mcus_y = 0;
    int restart_interval = 0;
    bool progressive = false;
    int spec_start = 0, spec_end = 0, succ_high = 0, succ_low = 0, eob_run = 0;
    unsigned char quant[4][64] = {{0}};
    HuffTable dc_tab[4], ac_tab[4];
    Component comp[4];

    int u8()
    {
        if (at >= size) throw JpegError("truncated");
        return base[at++];
    }
    int u16()
    {
        const int hi = u8();
        return hi * 256 + u8();
    }

    void read_dqt(int len)
    {
        while (len > 0) {
            const int q = u8(), precision = q >> 4, t = q & 15;
            if (precision != 0) throw JpegError("16-bit quantisation tables are not supported");
            if (t > 3) throw JpegError("bad DQT table");
            for (int i = 0; i < 64; ++i) quant[t][kNatural[i]] = (unsigned char)u8();
            len -= 65;
        }
        if (len != 0) throw JpegError("bad DQT len");
    }
    void read_dht(int len)
    {
        while (len > 0) {
            const int q = u8(), tc = q >> 4, th = q & 15;
            if (tc > 1 || th > 3) throw JpegError("bad DHT header");
            unsigned char counts[16], syms[256];
            int n = 0;
            for (int i = 0; i < 16; ++i) { counts[i] = (unsigned char)u8(); n += counts[i]; }
            if (n > 256) throw JpegError("bad DHT header");
            for (int i = 0; i < n; ++i) syms[i] = (unsigned char)u8();
            (tc == 0 ? dc_tab : ac_tab)[th].build(counts, syms, n);
            len -= 17 + n;
        }
        if (len != 0) throw JpegError("bad DHT len");
    }
    void read_sof(int len)
    {
        if (len < 11) throw JpegError("bad SOF len");
        if (u8() != 8) throw JpegError("only 8-bit JPEG is supported");
        height = u16();
        width = u16();
        if (height == 0 || width == 0) throw JpegError("0-sized image");
        n_comp = u8();
        if (n_comp != 1 && n_comp != 3) throw JpegError("bad component count");      // stb: 1 or 3 only
        if (len != 8 + 3 * n_comp) throw JpegError("bad SOF len");
        for (int i = 0; i < n_comp; ++i) {
            Component& c = comp[i];
            c.id = u8();
            const int q = u8();
            c.h = q >> 4;
            c.v = q & 15;
            c.tq = u8();
            if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) throw JpegError("bad component header");
            if (c.h > h_max) h_max = c.h;
            if (c.v > v_max) v_max = c.v;
        }
        mcus_x = (width + 8 * h_max - 1) / (8 * h_max);
        mcus_y = (height + 8 * v_max - 1) / (8 * v_max);
        for (int i = 0; i < n_comp; ++i) {
            Component& c = comp[i];
            c.x = (width * c.h + h_max - 1) / h_max;
            c.y = (height * c.v + v_max - 1) / v_max;
            c.w2 = mcus_x * c.h * 8;
            c.h2 = mcus_y * c.v * 8;
            c.data.assign((size_t)c.w2 * c.h2, 0);
            if (progressive) c.coef.assign((size_t)(c.w2 / 8) * (c.h2 / 8) * 64, 0);
        }
    }

    // progressive, DC band (stb_image.h:1747-1770): first pass = prediction + point transform, later passes add one bit
    void prog_dc(BitReader& br, Component& c, short* d)
    {
        if (spec_end != 0) throw JpegError("can't merge dc and ac");
        if (succ_high == 0) {
            const HuffTable& hd = dc_tab[c.td];
            if (!hd.present) throw JpegError("missing huffman table");
            std::memset(d, 0, 64 * sizeof(short));
            const int t = br.decode(hd);
            if (t > 15) throw JpegError("bad DC code");
            c.dc_pred += br.receive_extend(t);
            d[0] = (short)(c.dc_pred * (1 << succ_low));
        } else if (br.bit()) {
            d[0] = (short)(d[0] + (short)(1 << succ_low));
        }
    }

    // progressive, AC band [spec_start, spec_end] of one block (stb_image.h:1774-1890)
    void prog_ac(BitReader& br, Component& c, short* d)
    {
        if (spec_start == 0) throw JpegError("can't merge dc and ac");
        const HuffTable& ha = ac_tab[c.ta];
        if (!ha.present) throw JpegError("missing huffman table");
        if (succ_high == 0) {
            if (eob_run) { --eob_run; return; }
            int k = spec_start;
            do {
                const int rs = br.decode(ha), sbits = rs & 15, r = rs >> 4;
                if (sbits == 0) {
                    if (r < 15) {
                        eob_run = 1 << r;
                        if (r) eob_run += br.receive(r);
                        --eob_run;
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    if (k > 63) throw JpegError("bad AC run");
                    d[kNatural[k++]] = (short)(br.receive_extend(sbits) * (1 << succ_low));
                }
            } while (k <= spec_end);
            return;
        }
        const short bit = (short)(1 << succ_low);
        auto refine = [&](short* p) {
            if (br.bit() && (*p & bit) == 0) *p = (short)(*p > 0 ? *p + bit : *p - bit);
        };
        if (eob_run) {
            --eob_run;
            for (int k = spec_start; k <= spec_end; ++k) {
                short* p = &d[kNatural[k]];
                if (*p != 0) refine(p);
            }
            return;
        }
        int k = spec_start;
        do {
            const int rs = br.decode(ha);
            int sv = rs & 15, r = rs >> 4;
            if (sv == 0) {
                if (r < 15) {
                    eob_run = (1 << r) - 1;
                    if (r) eob_run += br.receive(r);
                    r = 64;                                   // force the end of the block
                }                                             // r == 15: sixteen zeros, nothing to place
            } else {
                if (sv != 1) throw JpegError("bad huffman code");
                sv = br.bit() ? bit : -bit;
            }
            while (k <= spec_end) {                           // skip r zero coefficients, refining the others on the way
                short* p = &d[kNatural[k++]];
                if (*p != 0) {
                    refine(p);
                } else {
                    if (r == 0) { *p = (short)sv; break; }
                    --r;
                }
            }
        } while (k <= spec_end);
    }

    // progressive: every scan is in -- de-quantise and transform the blocks that cover the image (stb_image.h:2527-2545)
    void finish_progressive()
    {
        for (int n = 0; n < n_comp; ++n) {
            Component& c = comp[n];
            const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3, stride = c.w2 / 8;
            const unsigned char* q = quant[c.tq];
            for (int j = 0; j < bh; ++j)
                for (int i = 0; i < bw; ++i) {
                    short* d = c.coef.data() + 64 * ((size_t)i + (size_t)j * stride);
                    for (int k = 0; k < 64; ++k) d[k] = (short)(d[k] * q[k]);
                    idct_block(c.data.data() + (size_t)c.w2 * j * 8 + i * 8, c.w2, d);
                }
        }
    }

    void decode_block(BitReader& br, Component& c, short d[64])
    {
        const HuffTable& hd = dc_tab[c.td];
        const HuffTable& ha = ac_tab[c.ta];
        if (!hd.present || !ha.present) throw JpegError("missing huffman table");
        const unsigned char* q = quant[c.tq];
        std::memset(d, 0, 64 * sizeof(short));
        const int t = br.decode(hd);
        if (t > 15) throw JpegError("bad DC code");
        c.dc_pred += br.receive_extend(t);
        d[0] = (short)(c.dc_pred * q[0]);
        for (int k = 1; k < 64;) {
            const int rs = br.decode(ha), s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (rs != 0xF0) break;               // end of block
                k += 16;
                continue;
            }
            k += r;
            if (k > 63) throw JpegError("bad AC run");
            const int nat = kNatural[k++];
            d[nat] = (short)(br.receive_extend(s) * q[nat]);
        }
    }

    // one scan (baseline: all 64 coefficients of every block it covers); returns the marker that ended it
    int read_scan()
    {
        const int len = u16();
        const int n = u8();
        if (n < 1 || n > n_comp || len != 6 + 2 * n) throw JpegError("bad SOS");
        int order[4];
        for (int i = 0; i < n; ++i) {
            const int id = u8(), q = u8();
            int which = 0;
            while (which < n_comp && comp[which].id != id) ++which;
            if (which == n_comp) throw JpegError("bad SOS component");
            comp[which].td = q >> 4;
            comp[which].ta = q & 15;
            if (comp[which].td > 3 || comp[which].ta > 3) throw JpegError("bad SOS table");
            order[i] = which;
        }
        const int ss = u8(), se = u8(), ahl = u8();
        if (progressive) {
            spec_start = ss; spec_end = se; succ_high = ahl >> 4; succ_low = ahl & 15;
            if (ss > 63 || se > 63 || ss > se || succ_high > 13 || succ_low > 13) throw JpegError("bad SOS");
            if (ss != 0 && n != 1) throw JpegError("bad SOS");    // an AC band is never interleaved
            eob_run = 0;
        } else if (ss != 0 || ahl != 0) {
            throw JpegError("bad SOS");                           // baseline (stb ignores `se` here too)
        }

        BitReader br(base + at, base + size);
        for (int i = 0; i < n_comp; ++i) comp[i].dc_pred = 0;
        int todo = restart_interval ? restart_interval : 0x7fffffff;
        short d[64];
        auto restart_if_due = [&]() {
            if (--todo > 0) return true;
            // interval finished: the next thing must be RSTn, else the scan is over
            br.reset();
            if (br.marker < 0 && br.p + 1 < br.end && br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7) {
                br.p += 2;
            } else if (br.marker >= 0xD0 && br.marker <= 0xD7) {
                br.marker = -1;
            } else {
                return false;
            }
            for (int i = 0; i < n_comp; ++i) comp[i].dc_pred = 0;
            eob_run = 0;
            todo = restart_interval ? restart_interval : 0x7fffffff;
            return true;
        };
        // one block of this scan: straight to samples (baseline) or into the coefficient store (progressive)
        auto block = [&](Component& c, int bx, int by) {
            if (!progressive) {
                decode_block(br, c, d);
                idct_block(c.data.data() + (size_t)c.w2 * by * 8 + bx * 8, c.w2, d);
                return;
            }
            short* cd = c.coef.data() + 64 * ((size_t)bx + (size_t)by * (c.w2 / 8));
            if (spec_start == 0) prog_dc(br, c, cd);
            else prog_ac(br, c, cd);
        };
        bool go = true;
        if (n == 1) {
            // non-interleaved: the blocks that cover the component's own samples, row by row
            Component& c = comp[order[0]];
            const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
            for (int j = 0; j < bh && go; ++j)
                for (int i = 0; i < bw && go; ++i) {
                    block(c, i, j);
                    go = restart_if_due();
                }
        } else {
            for (int j = 0; j < mcus_y && go; ++j)
                for (int i = 0; i < mcus_x && go; ++i) {
                    for (int k = 0; k < n; ++k) {
                        Component& c = comp[order[k]];
                        for (int y = 0; y < c.v; ++y)
                            for (int x = 0; x < c.h; ++x) {
                                block(c, i * c.h + x, j * c.v + y);
                            }
                    }
                    go = restart_if_due();
                }
        }
        // find the marker that follows the entropy-coded data
        at = (size_t)(br.p - base);
        if (br.marker >= 0) return br.marker;
        while (at + 1 < size) {
            if (base[at] == 0xFF && base[at + 1] != 0 && base[at + 1] != 0xFF) {
                const int m = base[at + 1];
                at += 2;
                return m;
            }
            ++at;
        }
        return 0xD9;
    }

    TextureImage run()
    {
        if (u8() != 0xFF || u8() != 0xD8) throw JpegError("no SOI");
        bool have_sof = false, have_scan = false;
        int pending = -1;
        for (;;) {
            int m = pending;
            pending = -1;
            if (m < 0) {
                int b = u8();
                if (b != 0xFF) throw JpegError("expected marker");
                do { m = u8(); } while (m == 0xFF);
            }
            if (m == 0xD9) break;
            if (m == 0xC0 || m == 0xC1) {
                if (have_sof) throw JpegError("two SOF markers");
                read_sof(u16());
                have_sof = true;
            } else if (m == 0xC2) {
                if (have_sof) throw JpegError("two SOF markers");
                progressive = true;
                read_sof(u16());
                have_sof = true;
            } else if (m == 0xC3 || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC)) {
                throw JpegError("unsupported JPEG coding process");
            } else if (m == 0xC4) {
                read_dht(u16() - 2);
            } else if (m == 0xDB) {
                read_dqt(u16() - 2);
            } else if (m == 0xDD) {
                if (u16() != 4) throw JpegError("bad DRI len");
                restart_interval = u16();
            } else if (m == 0xDA) {
                if (!have_sof) throw JpegError("SOS before SOF");
                pending = read_scan();
                have_scan = true;
            } else if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE || m == 0xCC || m == 0xDC) {
                const int len = u16();
                if (len < 2 || at + (size_t)(len - 2) > size) throw JpegError("bad segment length");
                at += (size_t)(len - 2);
            } else if (m >= 0xD0 && m <= 0xD7) {
                // stray restart marker between segments: ignored
            } else {
                throw JpegError("unknown marker");
            }
        }
        if (!have_scan) throw JpegError("no image data");
        if (progressive) finish_progressive();

        TextureImage t;
        t.width = width;
        t.height = height;
        t.components = n_comp;
        t.pixels.assign((size_t)width * height * n_comp, 0);

        // row-pair bookkeeping of load_jpeg_image (stb_image.h:3316-3368): line0/line1 are the two
        // input rows that bracket the output row; which one is "near" alternates with ystep.
        struct Up {
            int hs, vs, ystep, w_lores, ypos;
            const unsigned char *line0, *line1;
            std::vector<unsigned char> buf;
        } up[4];
        for (int k = 0; k < n_comp; ++k) {
            Up& u = up[k];
            u.hs = h_max / comp[k].h;
            u.vs = v_max / comp[k].v;
            u.ystep = u.vs >> 1;
            u.w_lores = (width + u.hs - 1) / u.hs;
            u.ypos = 0;
            u.line0 = u.line1 = comp[k].data.data();
            u.buf.assign((size_t)width + 8, 0);
        }
        const unsigned char* rows[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int j = 0; j < height; ++j) {
            for (int k = 0; k < n_comp; ++k) {
                Up& u = up[k];
                const bool bottom = u.ystep >= (u.vs >> 1);
                rows[k] = upsample_row(u.buf.data(), bottom ? u.line1 : u.line0, bottom ? u.line0 : u.line1,
                                       u.w_lores, u.hs, u.vs);
                if (++u.ystep >= u.vs) {
                    u.ystep = 0;
                    u.line0 = u.line1;
                    if (++u.ypos < comp[k].y) u.line1 += comp[k].w2;
                }
            }
            unsigned char* out = t.pixels.data() + (size_t)j * width * n_comp;
            if (n_comp == 3)
                for (int i = 0; i < width; ++i) ycc_to_rgb(out + 3 * i, rows[0][i], rows[1][i], rows[2][i]);
            else
                std::memcpy(out, rows[0], (size_t)width);
        }
        return t;
    }
};

}  // namespace

bool looks_like_jpeg(const unsigned char* data, size_t size)
{
    return size >= 3 && data[0] == 0xFF && data[1] == 0xD8 && data[2] == 0xFF;
}

TextureImage decode_jpeg(const unsigned char* data, size_t size, const std::string& name)
{
    try {
        Decoder d{data, size};
        return d.run();
    } catch (const JpegError& e) {
        throw std::runtime_error("Texture load error! (" + name + ": JPEG: " + e.what() + ")");
    }
}

}  // namespace b200host
