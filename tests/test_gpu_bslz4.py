"""Device bitshuffle-LZ4 decode (csrc/bslz4_gpu.cuh) through the C ABI: bit-exact against the
original frame and against the host decoder, on chunks produced by the real liblz4 (pyarrow
lz4_raw) and by the pure-python encoder (odd offsets, short periodic matches)."""
import ctypes as C

import numpy as np
import pytest

from bslz4_ref import compress_chunk as py_compress
from ffs_b200 import bslz4_encode, lib, synth
from ffs_b200.spotfinder import FfsError, Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu


def host_decode(chunk, n, dtype):
    out = np.zeros(n, dtype)
    rc = lib.load().ffs_bslz4_decode_chunk(chunk, len(chunk), out.ctypes.data, n, out.dtype.itemsize)
    assert rc == 0
    return out


def frames_for(h, w, dtype, seed):
    rng = np.random.default_rng(seed)
    a = rng.poisson(2.5, (h, w)).astype(dtype)
    a[rng.integers(0, h, 20), rng.integers(0, w, 20)] = np.iinfo(dtype).max      # gap-like pixels
    a[h // 3:h // 3 + 3, :] = np.iinfo(dtype).max                              # a module gap row
    return a


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("shape", [(64, 64), (37, 100), (33, 41), (130, 517), (256, 1028)])
@pytest.mark.parametrize("encoder", ["liblz4", "python"])
def test_decode_parity(shape, dtype, encoder):
    h, w = shape
    if encoder == "python" and h * w > 20000:
        pytest.skip("pure-python encoder is slow")
    a = frames_for(h, w, dtype, seed=h * w)
    chunk = bslz4_encode.compress_chunk(a) if encoder == "liblz4" else py_compress(a)
    with Spotfinder(w, h, dtype, None, n_slots=2) as sf:
        got = sf.decode_chunk(chunk, slot=1)
    assert np.array_equal(got, a)
    assert np.array_equal(host_decode(chunk, a.size, dtype).reshape(h, w), a)


@pytest.mark.parametrize("block", [128, 256, 1024, 4096])
def test_block_sizes_and_patterns(block):
    h, w = 96, 200
    rng = np.random.default_rng(block)
    cases = {
        "zeros": np.zeros((h, w), np.uint16),
        "incompressible": rng.integers(0, 65536, (h, w)).astype(np.uint16),
        "period3": np.tile(np.array([1, 2, 4], np.uint16), h * w // 3 + 1)[:h * w].reshape(h, w),
        "period7": np.tile(np.arange(7, dtype=np.uint16) * 37, h * w // 7 + 1)[:h * w].reshape(h, w),
        "ramp": (np.arange(h * w) // 5 % 251).astype(np.uint16).reshape(h, w),
        "far_repeat": np.tile(rng.integers(0, 4, 997).astype(np.uint16), h * w // 997 + 1)[:h * w].reshape(h, w),
    }
    with Spotfinder(w, h, np.uint16, None, n_slots=1) as sf:
        for name, a in cases.items():
            for chunk in (bslz4_encode.compress_chunk(a, block=block), py_compress(a, block=block)):
                assert np.array_equal(sf.decode_chunk(chunk), a), name


def test_corrupt_chunk_is_rejected():
    a = frames_for(64, 80, np.uint16, 3)
    chunk = bytearray(bslz4_encode.compress_chunk(a))
    with Spotfinder(80, 64, np.uint16, None, n_slots=1) as sf:
        with pytest.raises(FfsError):
            sf.decode_chunk(bytes(chunk[:-9]))              # truncated
        bad = bytearray(chunk)
        bad[12:16] = b"\x00\x00\x00\x05"                    # first block claims 5 bytes
        with pytest.raises(FfsError):
            sf.decode_chunk(bytes(bad))
        bad = bytearray(chunk)
        bad[20] ^= 0xF0                                     # damage a token inside the first LZ4 block
        try:
            got = sf.decode_chunk(bytes(bad))
            assert not np.array_equal(got, a)               # decodable garbage is acceptable ...
        except FfsError:
            pass                                            # ... and so is a format error
        assert np.array_equal(sf.decode_chunk(bytes(chunk)), a)  # the context survives


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_streaming_compressed_equals_raw(dtype):
    g = synth.small_geometry(1040, 520)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 4, dtype, config=1, n_spots=40)
    chunks = [bslz4_encode.compress_chunk(f) for f in frames]
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=4) as sf:
        raw = []
        for i, f in enumerate(frames):
            slot, buf = sf.acquire()
            buf[...] = f
            sf.submit(slot, i)
        while sf.pending:
            raw.append(sf.wait())
        comp = []
        for i, ch in enumerate(chunks):
            slot, buf = sf.acquire_compressed()
            buf[:len(ch)] = np.frombuffer(ch, np.uint8)
            sf.submit_compressed(slot, 100 + i, len(ch))
        while sf.pending:
            comp.append(sf.wait())
    assert len(raw) == len(comp) == 4
    for (r0, s0), (r1, s1) in zip(raw, comp):
        assert r1.status == 0 and r1.frame_id == r0.frame_id + 100
        assert r1.num_strong_pixels == r0.num_strong_pixels > 0
        assert r1.n_spots_total == r0.n_spots_total
        assert np.array_equal(s0, s1)


def test_full_size_frame_roundtrip():
    g = synth.EIGER2_16M
    mask = synth.make_mask(g)
    f = synth.make_frame(g, mask, np.uint16, 2, 0)
    chunk = bslz4_encode.compress_chunk(f)
    assert len(chunk) < f.nbytes / 3
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=1) as sf:
        assert np.array_equal(sf.decode_chunk(chunk), f)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_header_block_size_zero_means_the_default(dtype):
    """The reference calls bshuf_decompress_lz4(buffer + 12, out, n, elem_size, block_size = 0)
    (spotfinder/spotfinder.cc:830-834): it never reads the header's block-size field, it lets
    bitshuffle pick its default (8192 bytes per block).  A chunk whose header says 0 must decode
    like one that names the default."""
    g = synth.small_geometry(500, 300)
    mask = synth.make_mask(g)
    f = synth.make_frame(g, mask, dtype, config=2, frame_index=5, n_spots=20)
    chunk = bytearray(bslz4_encode.compress_chunk(f))
    assert int.from_bytes(chunk[8:12], "big") == 8192
    chunk[8:12] = (0).to_bytes(4, "big")
    assert np.array_equal(host_decode(bytes(chunk), f.size, dtype).reshape(f.shape), f)
    assert np.array_equal(oracle.bslz4_decode(bytes(chunk), dtype, f.size).reshape(f.shape), f)
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=1) as sf:
        assert np.array_equal(sf.decode_chunk(bytes(chunk)), f)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_crafted_sequences(dtype):
    """Every branch of the device decoder's sequence loop on purpose: LZ4 blocks written sequence
    by sequence (literal runs 0..14 from the register window, 15+ through the extension loop and
    the global copy; match lengths around the token limit, one extension byte, an extension byte of
    255, the 128-byte short-copy limit; offsets 1/2/4 runs, other powers of two, odd short periods,
    periods just above a 32-byte round, both sides of 132).  Expected elements come from the
    generator's byte-wise expansion and a numpy bit un-transpose, independent of every decoder."""
    from bslz4_ref import crafted_chunk
    rng = np.random.default_rng(77)
    es = np.dtype(dtype).itemsize
    n_blocks = 24
    w = 256
    h = n_blocks * (8192 // es) // w
    with Spotfinder(w, h, dtype, None, n_slots=1) as sf:
        for _ in range(6):
            chunk, want = crafted_chunk(rng, dtype, n_blocks)
            assert np.array_equal(sf.decode_chunk(chunk).reshape(-1), want)
            assert np.array_equal(host_decode(chunk, want.size, dtype), want)
