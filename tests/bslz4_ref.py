"""Test-side bitshuffle-LZ4 ENCODER (numpy), restating the published bitshuffle 0.5.x stream
format, used to round-trip the product's decoder (csrc/cli/bslz4.hpp).  LZ4 blocks are
emitted with simple run/repeat matching so the match path of the decoder is exercised."""
import struct

import numpy as np


def _lz4_block(data: bytes) -> bytes:
    """A valid (not optimal) LZ4 block: greedy matches against offset-1 runs and a 4-byte hash."""
    n = len(data)
    out = bytearray()
    i = anchor = 0
    table = {}
    end_guard = n - 12  # LZ4 rule: last 5 bytes literal, matches must start 12 bytes before end

    def emit(lit: bytes, mlen: int, offset: int):
        token_l = min(len(lit), 15)
        token_m = min(mlen - 4, 15) if mlen else 0
        out.append(token_l << 4 | token_m)
        if len(lit) >= 15:
            r = len(lit) - 15
            while r >= 255:
                out.append(255)
                r -= 255
            out.append(r)
        out.extend(lit)
        if mlen:
            out.extend(struct.pack("<H", offset))
            if mlen - 4 >= 15:
                r = mlen - 4 - 15
                while r >= 255:
                    out.append(255)
                    r -= 255
                out.append(r)

    while i < end_guard:
        key = data[i:i + 4]
        cand = table.get(key)
        table[key] = i
        if cand is not None and i - cand < 65535:
            m = 4
            while i + m < n - 5 and data[cand + m] == data[i + m]:
                m += 1
            emit(data[anchor:i], m, i - cand)
            i += m
            anchor = i
        else:
            i += 1
    emit(data[anchor:], 0, 0)
    return bytes(out)


def _bit_transpose(block: np.ndarray) -> bytes:
    """elements (n multiple of 8) -> elem_size*8 bit rows of n/8 bytes, element i in bit i%8."""
    n = block.size
    s = block.dtype.itemsize
    b = block.view(np.uint8).reshape(n, s)                       # [element, byte j]
    bits = np.unpackbits(b[:, :, None], axis=2, bitorder="little")  # [element, j, bit k]
    rows = bits.transpose(1, 2, 0).reshape(s * 8, n)             # row (j, k) over elements
    return np.packbits(rows, axis=1, bitorder="little").tobytes()


def compress_chunk(arr: np.ndarray, block: int | None = None) -> bytes:
    """HDF5-filter chunk: 12-byte header + bitshuffle-LZ4 body."""
    flat = np.ascontiguousarray(arr).reshape(-1)
    s = flat.dtype.itemsize
    n = flat.size
    if block is None:
        block = max((8192 // s) // 8 * 8, 128)
    out = bytearray(struct.pack(">QI", n * s, block * s))
    pos = 0
    for _ in range(n // block):
        c = _lz4_block(_bit_transpose(flat[pos:pos + block]))
        out += struct.pack(">I", len(c)) + c
        pos += block
    last = (n % block) // 8 * 8
    if last:
        c = _lz4_block(_bit_transpose(flat[pos:pos + last]))
        out += struct.pack(">I", len(c)) + c
        pos += last
    out += flat[pos:].tobytes()
    return bytes(out)
