// image_io.cpp -- see image_io.hpp.
#include "image_io.hpp"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <ctime>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace b200host {

namespace {

uint32_t crc_table[256];
bool crc_ready = false;

uint32_t crc32(const unsigned char* p, size_t n, uint32_t crc = 0)
{
    if (!crc_ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            crc_table[i] = c;
        }
        crc_ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = crc_table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return ~crc;
}

void put_be32(std::vector<unsigned char>& v, uint32_t x)
{
    v.push_back((unsigned char)(x >> 24));
    v.push_back((unsigned char)(x >> 16));
    v.push_back((unsigned char)(x >> 8));
    v.push_back((unsigned char)x);
}

void write_chunk(FILE* f, const char type[4], const std::vector<unsigned char>& data)
{
    std::vector<unsigned char> head;
    put_be32(head, (uint32_t)data.size());
    std::fwrite(head.data(), 1, 4, f);
    std::vector<unsigned char> body(type, type + 4);
    body.insert(body.end(), data.begin(), data.end());
    std::fwrite(body.data(), 1, body.size(), f);
    std::vector<unsigned char> tail;
    put_be32(tail, crc32(body.data(), body.size()));
    std::fwrite(tail.data(), 1, 4, f);
}

// zlib stream made of "stored" deflate blocks
std::vector<unsigned char> zlib_stored(const std::vector<unsigned char>& raw)
{
    std::vector<unsigned char> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78);
    z.push_back(0x01);
    size_t at = 0;
    do {
        const size_t n = raw.size() - at < 65535 ? raw.size() - at : 65535;
        z.push_back(at + n == raw.size() ? 1 : 0);      // BFINAL, BTYPE = 00
        z.push_back((unsigned char)(n & 0xFF));
        z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 0xFF));
        z.push_back((unsigned char)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + (long)at, raw.begin() + (long)(at + n));
        at += n;
    } while (at < raw.size());
    uint32_t a = 1, b = 0;                               // Adler-32
    for (unsigned char c : raw) { a = (a + c) % 65521u; b = (b + a) % 65521u; }
    put_be32(z, (b << 16) | a);
    return z;
}

}  // namespace

Image::Image(int x, int y) : xSize(x), ySize(y), pixels(3 * (size_t)x * y, 0.0f)
{
    if (x <= 0 || y <= 0) throw std::runtime_error("Image: bad size");
}

void Image::setPixel(int x, int y, float r, float g, float b)
{
    if (x < 0 || y < 0 || x >= xSize || y >= ySize) throw std::runtime_error("Image::setPixel out of range");
    float* p = &pixels[3 * ((size_t)y * xSize + x)];
    p[0] = r; p[1] = g; p[2] = b;
}

std::vector<unsigned char> Image::bytes() const
{
    std::vector<unsigned char> out(pixels.size());
    for (size_t i = 0; i < pixels.size(); ++i) {
        // glm::clamp(pix, 0, 1) * 255.f, then (unsigned char): min(max(x, 0), 1); a NaN stays NaN in
        // GLM's clamp and its conversion is undefined -- written as 0 here
        float v = pixels[i];
        v = v < 0.0f ? 0.0f : v;
        v = v > 1.0f ? 1.0f : v;
        v = v * 255.f;
        out[i] = std::isnan(v) ? 0 : (unsigned char)v;
    }
    return out;
}

void Image::savePNG(const std::string& baseFilename) const
{
    const std::string filename = baseFilename + ".png";
    const std::vector<unsigned char> rgb = bytes();
    std::vector<unsigned char> raw;
    raw.reserve(((size_t)xSize * 3 + 1) * ySize);
    for (int y = 0; y < ySize; ++y) {
        raw.push_back(0);                                // filter type: none
        raw.insert(raw.end(), rgb.begin() + (long)((size_t)y * xSize * 3), rgb.begin() + (long)((size_t)(y + 1) * xSize * 3));
    }
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + filename);
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    std::fwrite(sig, 1, 8, f);
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, (uint32_t)xSize);
    put_be32(ihdr, (uint32_t)ySize);
    ihdr.push_back(8);    // bit depth
    ihdr.push_back(2);    // colour type: RGB
    ihdr.push_back(0);
    ihdr.push_back(0);
    ihdr.push_back(0);
    write_chunk(f, "IHDR", ihdr);
    write_chunk(f, "IDAT", zlib_stored(raw));
    write_chunk(f, "IEND", {});
    std::fclose(f);
    std::cout << "Saved " << filename << "." << std::endl;
}

void Image::saveHDR(const std::string& baseFilename) const
{
    const std::string filename = baseFilename + ".hdr";
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + filename);
    std::fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", ySize, xSize);
    std::vector<unsigned char> row((size_t)xSize * 4);
    for (int y = 0; y < ySize; ++y) {
        for (int x = 0; x < xSize; ++x) {
            const float* p = &pixels[3 * ((size_t)y * xSize + x)];
            const float mx = std::fmax(p[0], std::fmax(p[1], p[2]));
            unsigned char* o = &row[(size_t)x * 4];
            if (!(mx > 1e-32f)) { o[0] = o[1] = o[2] = o[3] = 0; continue; }
            int e;
            const float norm = std::frexp(mx, &e) * 256.0f / mx;
            o[0] = (unsigned char)(p[0] * norm);
            o[1] = (unsigned char)(p[1] * norm);
            o[2] = (unsigned char)(p[2] * norm);
            o[3] = (unsigned char)(e + 128);
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    std::cout << "Saved " + filename + "." << std::endl;
}

Image image_from_sum(const float* sum_rgb, int width, int height, float samples)
{
    Image img(width, height);
    for (int x = 0; x < width; ++x)
        for (int y = 0; y < height; ++y) {
            const float* pix = sum_rgb + 3 * ((size_t)x + (size_t)y * width);
            img.setPixel(width - 1 - x, y, pix[0] / samples, pix[1] / samples, pix[2] / samples);
        }
    return img;
}

std::string output_basename(const std::string& imageName, const std::string& startTimeString, float samples)
{
    std::ostringstream ss;
    ss << imageName << "." << startTimeString << "." << samples << "samp";
    return ss.str();
}

std::string currentTimeString()
{
    time_t now;
    time(&now);
    char buf[sizeof "0000-00-00_00-00-00z"];
    strftime(buf, sizeof buf, "%Y-%m-%d_%H-%M-%Sz", gmtime(&now));
    return std::string(buf);
}

}  // namespace b200host
