/*
 * bslz4_oracle.c -- CPU ORACLE (test / bench infrastructure only, see ffs_oracle.h) for the
 * chunk decompression step of the reference's frame loop:
 *
 *     bshuf_decompress_lz4(buffer.data() + 12, host_image, width * height, sizeof(pixel_t), 0)
 *                                                   (spotfinder/spotfinder.cc:830-834)
 *
 * bitshuffle (0.5.2, environment.yml:6) and LZ4 are third-party libraries that are not in the
 * reference tree nor in this image; this is a scalar restatement of their PUBLISHED stream
 * formats, parity-pinned by (a) round trips through the real liblz4 (pyarrow "lz4_raw", the
 * encoder the tests and the bench use) and (b) equality with the original pixels.
 *
 *   chunk  = 8-byte BE uncompressed size | 4-byte BE block size in bytes | body      (HDF5 filter)
 *   body   = for every block of `block` elements (block * elem_size = block size; the last block
 *            holds the remaining multiple of 8 elements): 4-byte BE compressed length + one raw
 *            LZ4 block that decodes to the block's bit-transposed bytes; the final (n mod 8)
 *            elements follow uncompressed.
 *   bit transposition of a block of B elements of S bytes: S*8 bit planes of B/8 bytes, plane
 *   8*j + k = bit k of byte j of every element, element i at bit (i mod 8) of plane byte i/8.
 *
 * It is the CPU arm's decoder in bench.py's e2e_from_chunks leg ("port": the real library is
 * SIMD-vectorised and faster; the figure is reported, not optimised).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "ffs_oracle.h"

static uint32_t be32(const uint8_t *p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}
static uint64_t be64(const uint8_t *p) { return ((uint64_t)be32(p) << 32) | be32(p + 4); }

/* One raw LZ4 block (LZ4 block format: token = literal length << 4 | match length - 4, 255-run
 * length extensions, 2-byte LE offset; the last sequence has literals only). */
static int lz4_block(const uint8_t *src, size_t n_src, uint8_t *dst, size_t n_dst) {
    size_t ip = 0, op = 0;
    while (ip < n_src) {
        const unsigned token = src[ip++];
        size_t lit = token >> 4;
        if (lit == 15) {
            unsigned b;
            do {
                if (ip >= n_src) return -1;
                b = src[ip++];
                lit += b;
            } while (b == 255);
        }
        if (ip + lit > n_src || op + lit > n_dst) return -1;
        memcpy(dst + op, src + ip, lit);
        ip += lit;
        op += lit;
        if (ip >= n_src) break; /* last sequence */
        if (ip + 2 > n_src) return -1;
        const size_t offset = (size_t)src[ip] | ((size_t)src[ip + 1] << 8);
        ip += 2;
        if (offset == 0 || offset > op) return -1;
        size_t mlen = (token & 15u);
        if (mlen == 15) {
            unsigned b;
            do {
                if (ip >= n_src) return -1;
                b = src[ip++];
                mlen += b;
            } while (b == 255);
        }
        mlen += 4;
        if (op + mlen > n_dst) return -1;
        for (size_t k = 0; k < mlen; ++k, ++op) dst[op] = dst[op - offset]; /* overlapping copies repeat */
    }
    return op == n_dst ? 0 : -1;
}

/* 8x8 bit-matrix transpose of the 8 bytes of x (row r = byte r, little-endian). */
static uint64_t transpose8(uint64_t x) {
    uint64_t t;
    t = (x ^ (x >> 7)) & 0x00AA00AA00AA00AAull;
    x ^= t ^ (t << 7);
    t = (x ^ (x >> 14)) & 0x0000CCCC0000CCCCull;
    x ^= t ^ (t << 14);
    t = (x ^ (x >> 28)) & 0x00000000F0F0F0F0ull;
    x ^= t ^ (t << 28);
    return x;
}

/* planes (elem_size*8 rows of n/8 bytes) -> n elements */
static void unshuffle_block(const uint8_t *planes, uint8_t *out, size_t n, size_t elem_size) {
    const size_t row = n / 8;
    for (size_t c = 0; c < row; ++c)
        for (size_t j = 0; j < elem_size; ++j) {
            uint64_t x = 0;
            for (int k = 0; k < 8; ++k) x |= (uint64_t)planes[(8 * j + k) * row + c] << (8 * k);
            /* row k = bit k of byte j of elements 8c..8c+7 (element i at bit i): after the
             * transpose byte i holds byte j of element 8c + i */
            x = transpose8(x);
            for (int i = 0; i < 8; ++i) out[(8 * c + i) * elem_size + j] = (uint8_t)(x >> (8 * i));
        }
}

int ffs_oracle_bslz4_decode(const void *chunk, size_t chunk_bytes, void *out, size_t n_elements,
                            size_t elem_size) {
    const uint8_t *p = (const uint8_t *)chunk;
    if (!chunk || !out || chunk_bytes < 12 || (elem_size != 1 && elem_size != 2 && elem_size != 4 && elem_size != 8))
        return -1;
    if (be64(p) != (uint64_t)n_elements * elem_size) return -2;
    size_t block = be32(p + 8) / elem_size;
    if (block == 0) { /* bitshuffle's default: 8192 bytes, a multiple of 8 elements, at least 128 */
        block = (8192 / elem_size) / 8 * 8;
        if (block < 128) block = 128;
    }
    if (block % 8) return -3;
    size_t ip = 12, done = 0;
    uint8_t *dst = (uint8_t *)out;
    uint8_t *planes = (uint8_t *)malloc(block * elem_size);
    if (!planes) return -10;
    int rc = 0;
    while (n_elements - done >= 8) {
        size_t n = n_elements - done;
        if (n > block) n = block;
        n = n / 8 * 8;
        if (ip + 4 > chunk_bytes) { rc = -4; break; }
        const size_t clen = be32(p + ip);
        ip += 4;
        if (ip + clen > chunk_bytes || lz4_block(p + ip, clen, planes, n * elem_size)) { rc = -5; break; }
        ip += clen;
        unshuffle_block(planes, dst + done * elem_size, n, elem_size);
        done += n;
        if (n < block) break;
    }
    free(planes);
    if (rc) return rc;
    const size_t tail = (n_elements - done) * elem_size;
    if (ip + tail > chunk_bytes) return -6;
    memcpy(dst + done * elem_size, p + ip, tail);
    return 0;
}
