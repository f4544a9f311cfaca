// bslz4.hpp -- bitshuffle + LZ4 chunk decoder (host side of frame ingestion).
//
// The reference links the external bitshuffle 0.5.2 library and calls
//   bshuf_decompress_lz4(chunk + 12, out, n_pixels, sizeof(pixel_t), 0)
// (spotfinder/spotfinder.cc:830-834; h5read/src/read_chunks.cc:23-24).  That library is not
// available offline, so its published stream format is restated here:
//   HDF5 filter header : 8-byte BE uncompressed byte count, 4-byte BE block size in bytes
//   body               : for every block of `block` elements (default 8192 / elem_size
//                        rounded down to a multiple of 8; the last partial block is rounded
//                        down to a multiple of 8 elements): 4-byte BE compressed length +
//                        one raw LZ4 block that inflates to block*elem_size bytes holding the
//                        BIT-TRANSPOSED elements: row (byte j, bit k) of elem_size*8 rows,
//                        block/8 bytes per row, element i in bit i%8 of byte i/8;
//                        finally the trailing (n % 8) elements are stored raw.
// Parity: self-pinned by round trip with the encoder in tests/bslz4_ref.py (no golden data
// offline).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

namespace bslz4 {

inline uint32_t be32(const uint8_t* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }
inline uint64_t be64(const uint8_t* p) { return (uint64_t)be32(p) << 32 | be32(p + 4); }

// Raw LZ4 block decode.  Returns bytes written, or -1 on malformed input.
inline long lz4_block_decode(const uint8_t* src, size_t src_len, uint8_t* dst, size_t dst_cap) {
    const uint8_t* ip = src;
    const uint8_t* const iend = src + src_len;
    uint8_t* op = dst;
    uint8_t* const oend = dst + dst_cap;
    while (ip < iend) {
        const unsigned token = *ip++;
        size_t lit = token >> 4;
        if (lit == 15) {
            unsigned b;
            do {
                if (ip >= iend) return -1;
                b = *ip++;
                lit += b;
            } while (b == 255);
        }
        if ((size_t)(iend - ip) < lit || (size_t)(oend - op) < lit) return -1;
        memcpy(op, ip, lit);
        ip += lit;
        op += lit;
        if (ip >= iend) break;  // last sequence has no match
        if (iend - ip < 2) return -1;
        const size_t offset = ip[0] | (size_t)ip[1] << 8;
        ip += 2;
        if (offset == 0 || offset > (size_t)(op - dst)) return -1;
        size_t mlen = token & 15;
        if (mlen == 15) {
            unsigned b;
            do {
                if (ip >= iend) return -1;
                b = *ip++;
                mlen += b;
            } while (b == 255);
        }
        mlen += 4;
        if ((size_t)(oend - op) < mlen) return -1;
        const uint8_t* match = op - offset;
        if (offset >= mlen) {
            memcpy(op, match, mlen);
            op += mlen;
        } else {
            for (size_t k = 0; k < mlen; ++k) *op++ = *match++;  // overlapping copy
        }
    }
    return (long)(op - dst);
}

// 8x8 bit-matrix transpose of a little-endian uint64 (byte r = row r).
inline uint64_t transpose8(uint64_t x) {
    uint64_t t;
    t = (x ^ (x >> 7)) & 0x00AA00AA00AA00AAull;
    x = x ^ t ^ (t << 7);
    t = (x ^ (x >> 14)) & 0x0000CCCC0000CCCCull;
    x = x ^ t ^ (t << 14);
    t = (x ^ (x >> 28)) & 0x00000000F0F0F0F0ull;
    x = x ^ t ^ (t << 28);
    return x;
}

// in: elem_size*8 bit rows of n/8 bytes;  out: n elements.  n is a multiple of 8.
inline void bit_untranspose(const uint8_t* in, uint8_t* out, size_t n, size_t elem_size) {
    const size_t row_bytes = n / 8;
    for (size_t j = 0; j < elem_size; ++j) {
        const uint8_t* rows = in + j * 8 * row_bytes;
        for (size_t c = 0; c < row_bytes; ++c) {
            uint64_t x = 0;
            for (int k = 0; k < 8; ++k) x |= (uint64_t)rows[k * row_bytes + c] << (8 * k);
            // x: byte k = bit-row k (bit i = element 8c+i)  ->  byte i = byte j of element 8c+i
            x = transpose8(x);
            uint8_t* o = out + (8 * c) * elem_size + j;
            for (int i = 0; i < 8; ++i) o[i * elem_size] = (uint8_t)(x >> (8 * i));
        }
    }
}

inline size_t default_block(size_t elem_size) {
    size_t b = (8192 / elem_size) / 8 * 8;
    return b < 128 ? 128 : b;
}

// Decode a bitshuffle-LZ4 body (no 12-byte header).  block == 0 selects the default.
// Returns 0 on success, negative on error.
inline int decompress(const uint8_t* in, size_t in_len, void* out_v, size_t n, size_t elem_size, size_t block,
                      std::vector<uint8_t>& scratch) {
    uint8_t* out = static_cast<uint8_t*>(out_v);
    if (block == 0) block = default_block(elem_size);
    if (block % 8) return -2;
    scratch.resize(block * elem_size);
    const uint8_t* ip = in;
    const uint8_t* const iend = in + in_len;
    size_t done = 0;
    auto one = [&](size_t cnt) -> int {
        if (iend - ip < 4) return -3;
        const uint32_t clen = be32(ip);
        ip += 4;
        if ((size_t)(iend - ip) < clen) return -3;
        if (lz4_block_decode(ip, clen, scratch.data(), cnt * elem_size) != (long)(cnt * elem_size)) return -4;
        ip += clen;
        bit_untranspose(scratch.data(), out + done * elem_size, cnt, elem_size);
        done += cnt;
        return 0;
    };
    for (size_t b = 0; b < n / block; ++b)
        if (int rc = one(block)) return rc;
    size_t last = n % block;
    last -= last % 8;
    if (last)
        if (int rc = one(last)) return rc;
    const size_t leftover = (n % 8) * elem_size;
    if ((size_t)(iend - ip) < leftover) return -3;
    memcpy(out + done * elem_size, ip, leftover);
    return 0;
}

// HDF5-filter chunk: 12-byte header + body, as written to image_%06d_2 / raw HDF5 chunks.
inline int decompress_chunk(const uint8_t* chunk, size_t len, void* out, size_t n, size_t elem_size,
                            std::vector<uint8_t>& scratch) {
    if (len < 12) return -1;
    if (be64(chunk) != (uint64_t)n * elem_size) return -5;
    size_t block = be32(chunk + 8) / elem_size;
    if (block == 0 || block % 8) block = 0;
    return decompress(chunk + 12, len - 12, out, n, elem_size, block, scratch);
}

}  // namespace bslz4
