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


def bit_untranspose(raw: bytes, dtype) -> np.ndarray:
    """Inverse of _bit_transpose: elem_size*8 bit rows of n/8 bytes -> n elements."""
    s = np.dtype(dtype).itemsize
    n = len(raw) * 8 // (s * 8)
    rows = np.unpackbits(np.frombuffer(raw, np.uint8).reshape(s * 8, n // 8), axis=1, bitorder="little")
    bits = rows.reshape(s, 8, n).transpose(2, 0, 1)                  # [element, byte j, bit k]
    return np.packbits(bits, axis=2, bitorder="little").reshape(n, s).copy().view(dtype).reshape(n)


# every literal / match shape the decoders distinguish: literal runs inside and beyond the 15-value
# token field (and beyond one 32-byte parse window), match lengths around the token limit, one
# extension byte, an extension byte of 255, the 128-byte short-copy limit; offsets 1/2/4 (runs),
# other powers of two, odd short periods, periods just above a 32-byte round, both sides of 132
_LITS = (0, 0, 0, 1, 2, 5, 13, 14, 15, 16, 31, 32, 33, 40, 269, 270, 300)
_MLENS = (4, 5, 7, 18, 19, 20, 33, 64, 100, 127, 128, 129, 130, 200, 273, 274, 275, 529, 600)
_OFFSETS = (1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 31, 32, 33, 47, 64, 100, 127, 128, 131, 132, 133, 200, 512, 513)


def crafted_lz4_block(rng, out_len: int):
    """(lz4 block, the bytes it decodes to): random sequences drawn from the lists above."""
    out = bytearray()
    dec = bytearray()

    def put_len(r):
        while r >= 255:
            out.append(255)
            r -= 255
        out.append(r)

    def seq(lit: bytes, mlen: int, offset: int):
        out.append(min(len(lit), 15) << 4 | (min(mlen - 4, 15) if mlen else 0))
        if len(lit) >= 15:
            put_len(len(lit) - 15)
        out.extend(lit)
        dec.extend(lit)
        if mlen:
            out.extend(struct.pack("<H", offset))
            if mlen - 4 >= 15:
                put_len(mlen - 4 - 15)
            for _ in range(mlen):
                dec.append(dec[-offset])

    first = True
    while True:
        room = out_len - len(dec)
        if room < 16:
            break
        lit_n = 37 if first else int(rng.choice(_LITS))
        first = False
        mlen = int(rng.choice(_MLENS))
        if lit_n + mlen > room - 5:      # leave room for the closing literal-only sequence
            lit_n = min(lit_n, max(room - 5 - 4, 0))
            mlen = room - 5 - lit_n
            if mlen < 4:
                break
        offs = [o for o in _OFFSETS if o <= len(dec) + lit_n]
        offset = int(rng.choice(offs)) if offs else 1
        if len(dec) + lit_n == 0:
            lit_n = 1
        seq(bytes(rng.integers(0, 256, lit_n, dtype=np.uint8)), mlen, offset)
    seq(bytes(rng.integers(0, 256, out_len - len(dec), dtype=np.uint8)), 0, 0)
    assert len(dec) == out_len
    return bytes(out), bytes(dec)


def crafted_chunk(rng, dtype, n_blocks: int, block_bytes: int = 8192):
    """HDF5-filter chunk of n_blocks crafted LZ4 blocks and the elements it decodes to."""
    out = bytearray(struct.pack(">QI", n_blocks * block_bytes, block_bytes))
    elems = []
    for _ in range(n_blocks):
        c, raw = crafted_lz4_block(rng, block_bytes)
        out += struct.pack(">I", len(c)) + c
        elems.append(bit_untranspose(raw, dtype))
    return bytes(out), np.concatenate(elems)
