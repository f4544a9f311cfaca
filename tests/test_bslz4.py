"""CPU-only: the bitshuffle-LZ4 chunk decoder (host ingestion, csrc/cli/bslz4.hpp) round-trips
the test-side encoder, including ragged tails and corrupt input."""
import ctypes as C

import numpy as np
import pytest

from bslz4_ref import compress_chunk
from ffs_b200 import lib


def decode(chunk: bytes, n: int, dtype):
    out = np.zeros(n, dtype)
    rc = lib.load().ffs_bslz4_decode_chunk(chunk, len(chunk), out.ctypes.data, n, out.dtype.itemsize)
    return rc, out


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("n", [8, 9, 4096, 4097, 5000, 8191, 12345, 70 * 90])
def test_roundtrip(n, dtype):
    rng = np.random.default_rng(n)
    a = rng.poisson(3.0, n).astype(dtype)
    a[rng.integers(0, n, max(1, n // 50))] = np.iinfo(dtype).max  # gap pixels
    rc, out = decode(compress_chunk(a), n, dtype)
    assert rc == 0 and np.array_equal(out, a)


def test_incompressible_and_constant():
    rng = np.random.default_rng(1)
    a = rng.integers(0, 65536, 10000).astype(np.uint16)
    rc, out = decode(compress_chunk(a), a.size, np.uint16)
    assert rc == 0 and np.array_equal(out, a)
    z = np.zeros(20000, np.uint32)
    chunk = compress_chunk(z)
    assert len(chunk) < z.nbytes // 20
    rc, out = decode(chunk, z.size, np.uint32)
    assert rc == 0 and not out.any()


def test_corrupt_input_is_rejected():
    a = np.arange(5000, dtype=np.uint16)
    chunk = compress_chunk(a)
    assert decode(chunk[:-7], a.size, np.uint16)[0] != 0       # truncated
    assert decode(chunk, a.size + 8, np.uint16)[0] != 0        # wrong element count
    bad = bytearray(chunk)
    bad[12:16] = b"\xff\xff\xff\xff"                           # absurd block length
    assert decode(bytes(bad), a.size, np.uint16)[0] != 0


def test_lz4_blocks_pinned_against_liblz4():
    """The LZ4 block format is pinned against the real liblz4 (pyarrow's lz4_raw codec): it
    decodes what the test encoder emits, and the product's host decoder decodes what liblz4
    emits (chunks built by ffs_b200.bslz4_encode)."""
    pa = pytest.importorskip("pyarrow")
    from bslz4_ref import _lz4_block
    from ffs_b200 import bslz4_encode
    codec = pa.Codec("lz4_raw")
    rng = np.random.default_rng(5)
    for data in (bytes(5000), rng.integers(0, 4, 3000).astype(np.uint8).tobytes(),
                 rng.integers(0, 256, 2000).astype(np.uint8).tobytes(), b"abcabcabc" * 300):
        assert codec.decompress(_lz4_block(data), decompressed_size=len(data), asbytes=True) == data
    for dtype in (np.uint16, np.uint32):
        for n in (8, 4100, 9001, 70 * 90):
            a = rng.poisson(3.0, n).astype(dtype)
            chunk = bslz4_encode.compress_chunk(a)
            assert chunk[:12] == compress_chunk(a)[:12]
            rc, out = decode(chunk, n, dtype)
            assert rc == 0 and np.array_equal(out, a)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_oracle_decoder_matches_original_and_product_host_decoder(dtype):
    """oracle/bslz4_oracle.c (the CPU arm's decoder in bench.py) on chunks from the real liblz4
    (bslz4_encode) and from the odd-offset test encoder; corrupt input rejected."""
    from ffs_b200 import bslz4_encode
    from oracle import oracle
    rng = np.random.default_rng(9)
    for n in (8, 9, 4096, 4097, 12345, 70 * 90):
        a = rng.poisson(3.0, n).astype(dtype)
        a[rng.integers(0, n, max(1, n // 50))] = np.iinfo(dtype).max
        for enc in (bslz4_encode.compress_chunk, compress_chunk):
            chunk = enc(a)
            assert np.array_equal(oracle.bslz4_decode(chunk, dtype, n), a)
            rc, out = decode(chunk, n, dtype)
            assert rc == 0 and np.array_equal(out, a)
    chunk = bslz4_encode.compress_chunk(np.arange(5000, dtype=dtype))
    with pytest.raises(ValueError):
        oracle.bslz4_decode(chunk[:-9], dtype, 5000)
    with pytest.raises(ValueError):
        oracle.bslz4_decode(chunk, dtype, 5008)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_crafted_sequences_host_decoder_and_oracle(dtype):
    """LZ4 blocks written sequence by sequence (bslz4_ref.crafted_lz4_block: every literal length,
    match length and offset class the decoders distinguish), expected elements from the
    generator's own byte-wise expansion and an independent numpy bit un-transpose."""
    from bslz4_ref import crafted_chunk
    from oracle import oracle
    rng = np.random.default_rng(2026)
    for _ in range(4):
        chunk, want = crafted_chunk(rng, dtype, 5)
        assert np.array_equal(oracle.bslz4_decode(chunk, dtype, want.size).reshape(-1), want)
        rc, out = decode(chunk, want.size, dtype)
        assert rc == 0 and np.array_equal(out, want)
