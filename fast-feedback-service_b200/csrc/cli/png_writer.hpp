// png_writer.hpp -- smallest possible RGB8 PNG encoder for the --writeout diagnostics
// (the reference links lodepng for the same files, spotfinder.cc:641-645,987-991).
// Pixel data goes into "stored" deflate blocks: no compression, no dependency.
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace pngw {

inline uint32_t crc32(const uint8_t* p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}

inline void be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x);
}

inline bool write_chunk(FILE* f, const char* type, const std::vector<uint8_t>& data) {
    std::vector<uint8_t> head;
    be32(head, (uint32_t)data.size());
    head.insert(head.end(), type, type + 4);
    uint32_t crc = crc32(head.data() + 4, 4);
    crc = crc32(data.data(), data.size(), crc);
    std::vector<uint8_t> tail;
    be32(tail, crc);
    return fwrite(head.data(), 1, head.size(), f) == head.size() &&
           (data.empty() || fwrite(data.data(), 1, data.size(), f) == data.size()) &&
           fwrite(tail.data(), 1, 4, f) == 4;
}

// rgb: width*height*3 bytes, rows top to bottom
inline bool write_rgb(const std::string& path, const uint8_t* rgb, uint32_t width, uint32_t height) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    bool ok = fwrite(sig, 1, 8, f) == 8;
    std::vector<uint8_t> ihdr;
    be32(ihdr, width);
    be32(ihdr, height);
    ihdr.insert(ihdr.end(), {8, 2, 0, 0, 0});  // 8 bit, RGB, deflate, no filter method, no interlace
    ok = ok && write_chunk(f, "IHDR", ihdr);
    // raw scanlines: filter byte 0 + row
    const size_t row = (size_t)width * 3, raw_n = (row + 1) * height;
    std::vector<uint8_t> z;
    z.reserve(raw_n + raw_n / 65535 * 5 + 16);
    z.push_back(0x78);
    z.push_back(0x01);
    uint32_t a = 1, b = 0;  // adler32 of the raw stream
    std::vector<uint8_t> raw(raw_n);
    for (uint32_t y = 0; y < height; ++y) {
        raw[(row + 1) * y] = 0;
        std::copy(rgb + row * y, rgb + row * (y + 1), raw.begin() + (row + 1) * y + 1);
    }
    for (size_t pos = 0; pos < raw_n || pos == 0;) {
        const size_t n = std::min<size_t>(65535, raw_n - pos);
        z.push_back(pos + n >= raw_n ? 1 : 0);  // BFINAL, BTYPE = 00 (stored)
        z.push_back(n & 0xff); z.push_back(n >> 8);
        z.push_back(~n & 0xff); z.push_back((~n >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        pos += n;
        if (n == 0) break;
    }
    for (size_t i = 0; i < raw_n;) {  // adler32 in blocks that cannot overflow
        const size_t n = std::min<size_t>(5552, raw_n - i);
        for (size_t k = 0; k < n; ++k) {
            a += raw[i + k];
            b += a;
        }
        a %= 65521;
        b %= 65521;
        i += n;
    }
    be32(z, (b << 16) | a);
    ok = ok && write_chunk(f, "IDAT", z);
    ok = ok && write_chunk(f, "IEND", {});
    return fclose(f) == 0 && ok;
}

}  // namespace pngw
