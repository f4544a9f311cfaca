"""bitshuffle-LZ4 chunk ENCODER for workload generation (tests, bench): the HDF5 filter
stream the detector writes (12-byte header, per block a 4-byte BE length + a raw LZ4 block of
the bit-transposed elements, see csrc/cli/bslz4.hpp).  The LZ4 blocks come from the real
liblz4 through pyarrow's "lz4_raw" codec, so the decoders are exercised on genuine LZ4
output.  numpy only; not part of the hot path."""
from __future__ import annotations

import struct

import numpy as np


def default_block(elem_size: int) -> int:
    return max((8192 // elem_size) // 8 * 8, 128)


def bit_transpose_blocks(a: np.ndarray) -> np.ndarray:
    """a: (n_blocks, block) elements, block % 8 == 0 -> (n_blocks, elem_size*8*block/8) bytes:
    plane (byte j, bit k) = row 8j+k of block/8 bytes, element i in bit i%8 of byte i/8."""
    nb, blk = a.shape
    s = a.dtype.itemsize
    b = np.ascontiguousarray(a).view(np.uint8).reshape(nb, blk // 8, 8, s)      # [block, c, i, j]
    bits = np.unpackbits(b[..., None], axis=4, bitorder="little")               # [block, c, i, j, k]
    bits = bits.transpose(0, 3, 4, 1, 2)                                        # [block, j, k, c, i]
    return np.packbits(bits, axis=4, bitorder="little").reshape(nb, s * 8 * (blk // 8))


def _codec():
    import pyarrow as pa
    return pa.Codec("lz4_raw")


def compress_chunk(arr: np.ndarray, block: int | None = None, group: int = 256) -> bytes:
    flat = np.ascontiguousarray(arr).reshape(-1)
    s = flat.dtype.itemsize
    n = flat.size
    if block is None:
        block = default_block(s)
    codec = _codec()
    out = [struct.pack(">QI", n * s, block * s)]
    n_full = n // block
    for g0 in range(0, n_full, group):  # bounded temporaries
        g1 = min(n_full, g0 + group)
        planes = bit_transpose_blocks(flat[g0 * block:g1 * block].reshape(g1 - g0, block))
        for row in planes:
            c = codec.compress(row.tobytes(), asbytes=True)
            out.append(struct.pack(">I", len(c)))
            out.append(c)
    pos = n_full * block
    last = (n % block) // 8 * 8
    if last:
        planes = bit_transpose_blocks(flat[pos:pos + last].reshape(1, last))
        c = codec.compress(planes[0].tobytes(), asbytes=True)
        out.append(struct.pack(">I", len(c)))
        out.append(c)
        pos += last
    out.append(flat[pos:].tobytes())
    return b"".join(out)
