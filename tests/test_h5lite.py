"""h5lite (csrc/cli/h5lite.hpp): the in-tree HDF5 reader / writer behind `.nxs` input and
`--save-h5`.  No GPU, no libhdf5.

Pins, in order of strength:
  * the two libhdf5-written files the reference keeps for its own tests
    (dx2/tests/data/empty.h5, cut_strong.refl -> tests/golden/h5/): the reader must return the
    values dx2's tests assert (dx2/tests/test_read_h5_array.cxx:62-84,121-131,
    test_reflection_table.cxx:23-49), and an empty file from the writer must be byte-identical
    to libhdf5's;
  * a second, independent parser (tests/h5dissect.py, Python) that agrees with the C++ reader
    on those files and walks everything the C++ writer produces;
  * for the structures no fixture holds (fixed array / extensible array chunk indexes,
    version 2 / 3 hyperslab encodings), files patched or blobs built here from the format
    specification -- these check the code against the specification as read, not against
    libhdf5 ("parity unpinned" for those).
"""
import json
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

import h5dissect
from bslz4_ref import compress_chunk
from ffs_b200 import synth

ROOT = Path(__file__).resolve().parents[1]
TOOL = ROOT / "fast-feedback-service_b200" / "ffs_b200" / "h5lite_tool"
GOLD = Path(__file__).resolve().parent / "golden" / "h5"

pytestmark = pytest.mark.skipif(not TOOL.exists(), reason="h5lite_tool not built (python __graft_entry__.py)")


def tool(*args, check=True):
    r = subprocess.run([str(TOOL), *map(str, args)], capture_output=True, timeout=120)
    if check:
        assert r.returncode == 0, r.stderr.decode()
    return r


def dump(path):
    return json.loads(tool("dump", path).stdout)


def cat(path, dataset, dtype):
    return np.frombuffer(tool("cat", path, dataset).stdout, dtype)


# ---------------------------------------------------------------------------------------------
# the reference's own fixtures
# ---------------------------------------------------------------------------------------------
def test_reference_fixture_values_cpp_reader():
    refl = GOLD / "cut_strong.refl"
    xyz = cat(refl, "/dials/processing/group_0/xyzobs.px.value", "<f8")
    assert xyz[10] == 528.86470588235295          # test_read_h5_array.cxx:62-72
    flags = cat(refl, "/dials/processing/group_0/flags", "<u8")
    assert flags[5] == 32                         # test_read_h5_array.cxx:74-84
    d = dump(refl)["objects"]
    assert d["/dials/processing/group_0/xyzobs.px.value"]["shape"] == [10, 3]   # :121-131
    assert d["/dials/processing/group_0/xyzobs.px.value"]["dtype"] == "<f8"
    assert d["/dials/processing/empty_dataset"]["shape"] == [0]                   # :99-106
    assert cat(refl, "/dials/processing/empty_dataset", "<f8").size == 0
    g = d["/dials/processing/group_0"]
    assert g["attrs"]["experiment_ids"]["value"] == [0]
    assert g["attrs"]["identifiers"]["value"] == ["86434230-7132-26c2-32dc-438091cbc8e6"]
    assert "xyzobs.px.value" in g["links"] and "id" in g["links"]                # test_reflection_table.cxx:23-36
    # a dataset that does not exist / a file that does not exist: an error, not a crash (:86-97)
    assert tool("cat", refl, "/this/does/not/exist", check=False).returncode == 1
    assert tool("cat", "invalid_file.h5", "/some/dataset", check=False).returncode == 1


def test_cpp_reader_agrees_with_python_parser_on_reference_fixture():
    refl = GOLD / "cut_strong.refl"
    f = h5dissect.File(refl)
    d = dump(refl)["objects"]
    seen = 0
    for path, obj in f.walk():
        assert path in d
        e = d[path]
        assert e["address"] == obj.addr
        if obj.is_dataset:
            assert e["shape"] == obj.shape()
            dt = obj.dtype()
            if dt.cls in (0, 1):
                assert e["dtype"] == dt.np_dtype.str
                assert np.array_equal(cat(refl, path, dt.np_dtype).reshape(obj.shape()), obj.read())
                seen += 1
        else:
            assert sorted(e["links"]) == sorted(obj.links())
        for name, val in obj.attrs().items():
            got = e["attrs"][name]["value"]
            want = [val] if isinstance(val, str) else list(np.atleast_1d(val).tolist())
            assert got == want
    assert seen >= 15


def test_empty_file_is_byte_identical_to_libhdf5(tmp_path):
    out = tmp_path / "e.h5"
    tool("write-empty", out)
    assert out.read_bytes() == (GOLD / "empty.h5").read_bytes()


# ---------------------------------------------------------------------------------------------
# the writer, walked by the independent parser
# ---------------------------------------------------------------------------------------------
def check_structure(path):
    """Invariants every file from the writer must hold."""
    f = h5dissect.File(path)
    assert f.version == 0 and f.eof == len(f.buf) and f.base == 0
    for p, o in f.walk():
        assert o.addr % 8 == 0, p
        if o.is_group and o.first(0x11) is not None:
            btree, heap = struct.unpack_from("<QQ", o.first(0x11), 0)
            assert f.buf[btree:btree + 4] == b"TREE" and f.buf[heap:heap + 4] == b"HEAP"
            names = list(o.links())
            assert names == sorted(names)
    return f


@pytest.mark.parametrize("n", [0, 1, 13])
def test_results_table_rotation(tmp_path, n):
    out = tmp_path / "results_ffs.h5"
    tool("write-table", out, n)
    f = check_structure(out)
    g = f.get("/dials/processing/group_0")
    a = g.attrs()
    assert a["experiment_ids"].dtype == np.uint64 and a["experiment_ids"].tolist() == [0]
    assert len(a["identifiers"]) == 1 and len(a["identifiers"][0]) == 36        # dx2 utils.cxx:44-68
    assert sorted(g.links()) == ["id", "sigma_b_variance", "sigma_m_variance", "spot_extent_z", "xyzobs.px.value"]
    xyz = f.get("/dials/processing/group_0/xyzobs.px.value").read()
    assert xyz.dtype == np.float64 and xyz.shape == (n, 3)
    i = np.arange(n)
    assert np.array_equal(xyz, np.stack([0.5 + i, 100.25 + 2.0 * i, (i % 7) + 0.125], axis=1).reshape(n, 3))
    # N x 1 columns are stored as (N,) (reflection.cxx:236-243); `id` and `spot_extent_z` are C ints
    for name, dt in [("id", np.int32), ("spot_extent_z", np.int32), ("sigma_b_variance", np.float64)]:
        col = f.get(f"/dials/processing/group_0/{name}").read()
        assert col.dtype == dt and col.shape == (n,)
    assert np.array_equal(f.get("/dials/processing/group_0/spot_extent_z").read(), (i % 9) + 1)
    # and the C++ reader reads its own file
    assert np.array_equal(cat(out, "/dials/processing/group_0/xyzobs.px.value", "<f8").reshape(n, 3), xyz)


def test_results_table_stills_has_one_experiment_per_image(tmp_path):
    out = tmp_path / "r.h5"
    tool("write-table", out, 10, "2d")
    f = check_structure(out)
    a = f.get("/dials/processing/group_0").attrs()
    assert a["experiment_ids"].tolist() == [0, 1, 2, 3]
    assert len(set(a["identifiers"])) == 4
    assert f.get("/dials/processing/group_0/id").read().tolist() == [0, 0, 0, 1, 1, 1, 2, 2, 2, 3]


N_STRESS = 300
STRESS_MISSING = [7, 57, 107, 157, 207, 257]


def stress_frames():
    want = np.arange(N_STRESS * 20, dtype=np.uint16).reshape(N_STRESS, 4, 5)
    want[STRESS_MISSING] = 0
    return want


def test_writer_stress_file(tmp_path):
    out = tmp_path / "s.h5"
    tool("write-stress", out)
    f = check_structure(out)
    # 300 children: 38 symbol nodes under a two-level B-tree
    many = f.get("/many")
    assert list(many.links()) == [f"d{i:03d}" for i in range(300)]
    for i in (0, 7, 8, 255, 256, 299):
        assert f.get(f"/many/d{i:03d}").read().tolist() == [i, -i]
        assert cat(out, f"/many/d{i:03d}", "<i4").tolist() == [i, -i]
    # 300-chunk dataset, six chunks never written (they read as zeros)
    fr = f.get("/frames")
    want = stress_frames()
    assert len(fr.chunks()) == N_STRESS - len(STRESS_MISSING)
    assert np.array_equal(fr.read(), want)
    assert np.array_equal(cat(out, "/frames", "<u2").reshape(N_STRESS, 4, 5), want)
    assert tool("chunk", out, "/frames", 57, 0, 0, check=False).returncode == 3
    r = tool("chunk", out, "/frames", N_STRESS - 1, 0, 0)
    assert np.array_equal(np.frombuffer(r.stdout, "<u2"), want[-1].ravel())
    # ragged edge chunks
    rg = np.arange(70, dtype=np.float64).reshape(7, 10) + 0.5
    assert np.array_equal(f.get("/ragged").read(), rg)
    assert np.array_equal(cat(out, "/ragged", "<f8").reshape(7, 10), rg)
    # links and attributes
    assert f.root.links()["alias"] == ("soft", "/many/d042")
    assert cat(out, "/alias", "<i4").tolist() == [42, -42]
    assert cat(out, "/links/relative", "u1").tolist() == [200]
    a = f.get("/links").attrs()
    assert a["fixed"] == b"NXdata" and a["i64"] == -1234567890123 and a["names"] == ["alpha", "", "gamma-delta"]
    assert np.array_equal(a["f32x3"], np.array([1.5, -2.25, 1e-3], np.float32))
    d = dump(out)["objects"]["/links"]["attrs"]
    assert d["fixed"]["value"] == ["NXdata"] and d["i64"]["value"] == [-1234567890123]
    assert d["names"]["value"] == ["alpha", "", "gamma-delta"] and d["names"]["shape"] == [3]
    assert f.get("/links/empty").shape() == [0]


# ---------------------------------------------------------------------------------------------
# NeXus master + data files (what H5Read opens)
# ---------------------------------------------------------------------------------------------
def write_stream(d, frames, mask, bits, omega_increment=None):
    h, w = mask.shape
    hdr = {"nimages": len(frames), "ntrigger": 1, "y_pixels_in_detector": h, "x_pixels_in_detector": w,
           "bit_depth_image": bits, "countrate_correction_count_cutoff": 60000, "wavelength": 0.976,
           "detector_distance": 150.0, "y_pixel_size": 7.5e-5, "x_pixel_size": 7.5e-5,
           "beam_center_x": w / 2, "beam_center_y": h / 2}
    if omega_increment is not None:
        hdr.update(omega_start=10.0, omega_increment=omega_increment)
    (d / "start_1").write_text(json.dumps(hdr))
    (d / "start_4").write_bytes(b"")
    (d / "start_5").write_bytes((mask == 0).astype(np.int32).tobytes())
    for i, f in enumerate(frames):
        (d / f"image_{i:06d}_2").write_bytes(compress_chunk(f))


@pytest.mark.parametrize("dtype,bits", [(np.uint16, 16), (np.uint32, 32)])
def test_stream2nxs_layout(tmp_path, dtype, bits):
    g = synth.small_geometry(96, 64)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 7, dtype, config=8, n_spots=5)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, bits, omega_increment=0.1)
    tool("stream2nxs", sd, tmp_path / "scan", 3)
    maps = json.loads(tool("vds", tmp_path / "scan.nxs", "/entry/data/data").stdout)
    assert [(m["file"], m["dataset"], m["dst_start"][0], m["dst_block"][0]) for m in maps] == [
        ("scan_000001.h5", "/data", 0, 3), ("scan_000002.h5", "/data", 3, 3), ("scan_000003.h5", "/data", 6, 1)]
    assert all(m["src_start"] == [0, 0, 0] and m["dst_block"][1:] == [64, 96] for m in maps)
    m = check_structure(tmp_path / "scan.nxs")
    d = m.get("/entry/data/data")
    assert d.shape() == [7, 64, 96] and d.dtype().np_dtype == np.dtype(dtype)
    pm = m.get("/entry/instrument/detector/pixel_mask").read()
    assert pm.dtype == np.uint32 and np.array_equal(pm == 0, mask != 0)
    assert m.get("/entry/instrument/detector/saturation_value").read() == 60000
    assert abs(float(m.get("/entry/instrument/detector/distance").read()) - 0.15) < 1e-7
    assert np.allclose(m.get("/entry/sample/sample_omega/omega").read(), 10.0 + 0.1 * np.arange(7))
    assert m.get("/entry").attrs()["NX_class"] == "NXentry"
    for k, (first, cnt) in enumerate([(0, 3), (3, 3), (6, 1)]):
        path = tmp_path / f"scan_{k + 1:06d}.h5"
        f = check_structure(path)
        ds = f.get("/data")
        assert ds.shape() == [cnt, 64, 96] and ds.layout()["chunk"] == [1, 64, 96]
        assert ds.filters()[0]["id"] == 32008 and ds.filters()[0]["client_data"][2] == np.dtype(dtype).itemsize
        for i in range(cnt):
            want = compress_chunk(frames[first + i])
            assert bytes(ds.read_chunk_raw((i, 0, 0))[1]) == want
            assert tool("chunk", path, "/data", i, 0, 0).stdout == want


# ---------------------------------------------------------------------------------------------
# robustness: damaged files are refused, never crash the process
# ---------------------------------------------------------------------------------------------
def test_damaged_files_fail_cleanly(tmp_path):
    good = (GOLD / "cut_strong.refl").read_bytes()
    rng = np.random.default_rng(5)
    cases = [good[:n] for n in (0, 7, 8, 50, 96, 500, 2000, 8000)]
    for _ in range(150):
        b = bytearray(good)
        for _ in range(int(rng.integers(1, 6))):
            b[int(rng.integers(8, len(b)))] = int(rng.integers(0, 256))
        cases.append(bytes(b))
    for i, blob in enumerate(cases):
        p = tmp_path / f"bad{i}.h5"
        p.write_bytes(blob)
        for args in (("dump", p), ("cat", p, "/dials/processing/group_0/xyzobs.px.value")):
            r = tool(*args, check=False)
            assert r.returncode in (0, 1), (i, args, r.returncode, r.stderr[-200:])
    # the same on a file with chunk B-trees, a two-level group B-tree and many heaps; corruptions are
    # biased to the metadata region so that they land in structures rather than in pixel data
    tool("write-stress", tmp_path / "s.h5")
    good = (tmp_path / "s.h5").read_bytes()
    for i in range(120):
        b = bytearray(good)
        for _ in range(int(rng.integers(1, 8))):
            hi = len(b) if rng.random() < 0.3 else min(len(b), 95000)
            b[int(rng.integers(8, hi))] = int(rng.integers(0, 256))
        p = tmp_path / "fuzz.h5"
        p.write_bytes(b)
        for args in (("dump", p), ("cat", p, "/frames"), ("cat", p, "/ragged"), ("chunk", p, "/frames", 17, 0, 0)):
            r = tool(*args, check=False)
            assert r.returncode in (0, 1, 3), (i, args, r.returncode, r.stderr[-200:])


# ---------------------------------------------------------------------------------------------
# structures without a libhdf5-written fixture: built here from the format specification
# ---------------------------------------------------------------------------------------------
def test_hyperslab_selection_encodings():
    def sel(blob):
        return json.loads(tool("selection", blob.hex()).stdout)

    # version 1: one block given by its start and end corners (u32)
    v1 = struct.pack("<6I", 2, 1, 0, 8 + 8 * 3, 3, 1) + struct.pack("<3I", 5, 0, 0) + struct.pack("<3I", 14, 63, 95)
    assert sel(v1) == {"used": len(v1), "all": False, "start": [5, 0, 0], "block": [10, 64, 96]}
    # version 2: flags (bit 0 = regular), length, rank, then start / stride / count / block as u64
    body = b"".join(struct.pack("<4Q", s, st, c, b) for s, st, c, b in [(5, 10, 1, 10), (0, 64, 1, 64), (0, 96, 1, 96)])
    v2 = struct.pack("<II", 2, 2) + struct.pack("<BII", 1, len(body), 3) + body
    assert sel(v2) == {"used": len(v2), "all": False, "start": [5, 0, 0], "block": [10, 64, 96]}
    # count blocks laid end to end (stride == block) are one run
    body = b"".join(struct.pack("<4Q", s, st, c, b) for s, st, c, b in [(0, 1, 100, 1), (0, 64, 1, 64), (0, 96, 1, 96)])
    v2 = struct.pack("<II", 2, 2) + struct.pack("<BII", 1, len(body), 3) + body
    assert sel(v2)["block"] == [100, 64, 96]
    # version 3: flags, encoding size, rank, then the four values per dimension in that size
    body = b"".join(struct.pack("<4I", s, st, c, b) for s, st, c, b in [(7, 3, 1, 3), (0, 64, 1, 64), (0, 96, 1, 96)])
    v3 = struct.pack("<II", 2, 3) + struct.pack("<BBI", 1, 4, 3) + body
    assert sel(v3) == {"used": len(v3), "all": False, "start": [7, 0, 0], "block": [3, 64, 96]}
    # H5S_ALL
    assert sel(struct.pack("<4I", 3, 1, 0, 0)) == {"used": 16, "all": True, "start": [], "block": []}
    # irregular / multi-block selections are refused, not misread
    bad = struct.pack("<6I", 2, 1, 0, 8 + 16 * 3, 3, 2) + bytes(48)
    assert tool("selection", bad.hex(), check=False).returncode == 1


CK = h5dissect.with_checksum   # block + its lookup3 checksum


def test_lookup3_reproduces_libhdf5_checksums():
    """The three version 2 object headers of cut_strong.refl carry checksums computed by libhdf5."""
    buf = (GOLD / "cut_strong.refl").read_bytes()
    seen = 0
    for addr in (0x320, 0x65E, 0x563):
        assert buf[addr:addr + 4] == b"OHDR" and buf[addr + 4] == 2
        flags = buf[addr + 5]
        p = addr + 6 + (16 if flags & 0x20 else 0) + (4 if flags & 0x10 else 0)
        n = 1 << (flags & 3)
        end = p + n + int.from_bytes(buf[p:p + n], "little")
        assert h5dissect.lookup3(buf[addr:end]) == struct.unpack_from("<I", buf, end)[0]
        seen += 1
    assert seen == 3


def test_cpp_reader_verifies_checksums(tmp_path):
    """The C++ reader checks them while it walks the file: one flipped payload bit is noticed."""
    bad = bytearray((GOLD / "cut_strong.refl").read_bytes())
    bad[0x320 + 20] ^= 1
    p = tmp_path / "bad.refl"
    p.write_bytes(bad)
    r = tool("dump", p, check=False)
    assert r.returncode == 1 and b"checksum mismatch in an object header" in r.stderr
    assert tool("dump", GOLD / "cut_strong.refl").returncode == 0


def _patch_layout(buf, new_layout):
    """Replace the version 3 chunked layout message of the single dataset in `buf`."""
    # message header (type 8, size 32, flags 0) + version 3, chunked, dimensionality 4 (rank 3 + element size)
    old = struct.pack("<HHI", 8, 32, 0) + bytes([3, 2, 4])
    assert buf.count(old) == 1 and len(new_layout) <= 32
    at = buf.index(old) + 8
    size = struct.unpack_from("<H", buf, at - 6)[0]
    assert size == 32
    buf[at:at + 32] = new_layout + bytes(32 - len(new_layout))
    return at


def _chunks_of(path):
    f = h5dissect.File(path)
    return f, {c[0][0]: (c[3], c[2], c[1]) for c in f.get("/frames").chunks()}


def test_fixed_array_chunk_index(tmp_path):
    """A layout version 4 dataset indexed by a fixed array (what libhdf5 1.10+ writes for a
    chunked dataset with no unlimited dimension under the latest file format), paged and not."""
    src = tmp_path / "s.h5"
    tool("write-stress", src)
    f, chunks = _chunks_of(src)
    for page_bits in (10, 5):  # 300 entries: one block / 10 pages of 32
        buf = bytearray(f.buf)
        nelmts, entry = N_STRESS, 8
        hdr_addr = len(buf)
        dblk_addr = hdr_addr + 28
        hdr = CK(b"FAHD" + bytes([0, 0, entry, page_bits]) + struct.pack("<QQ", nelmts, dblk_addr))
        elems = [struct.pack("<Q", chunks[i][0] if i in chunks else h5dissect.UNDEF) for i in range(nelmts)]
        dblk = b"FADB" + bytes([0, 0]) + struct.pack("<Q", hdr_addr)
        page_n = 1 << page_bits
        if nelmts > page_n:
            npages = -(-nelmts // page_n)
            bitmap = bytearray(-(-npages // 8))
            for p in range(npages):
                bitmap[p // 8] |= 0x80 >> (p % 8)
            dblk = CK(dblk + bytes(bitmap))
            for p in range(npages):
                dblk += CK(b"".join(elems[p * page_n:(p + 1) * page_n]))
        else:
            dblk = CK(dblk + b"".join(elems))
        buf += hdr + dblk
        lay = bytes([4, 2, 0, 4, 4]) + struct.pack("<4I", 1, 4, 5, 2) + bytes([3, page_bits]) + struct.pack("<Q", hdr_addr)
        _patch_layout(buf, lay)
        struct.pack_into("<Q", buf, 40, len(buf))  # end-of-file address
        out = tmp_path / f"fa{page_bits}.h5"
        out.write_bytes(buf)
        assert dump(out)["objects"]["/frames"]["index"] == 3
        assert np.array_equal(cat(out, "/frames", "<u2").reshape(N_STRESS, 4, 5), stress_frames())
        assert tool("chunk", out, "/frames", 107, 0, 0, check=False).returncode == 3


def test_extensible_array_chunk_index(tmp_path):
    """A layout version 4 dataset indexed by an extensible array (one unlimited dimension --
    what a SWMR writer such as Odin produces), with libhdf5's default creation parameters:
    4 elements in the index block, then super block s holds 2^(s/2) data blocks of
    16 * 2^((s+1)/2) elements (1x16, 1x32, 2x32, 2x64, 4x64, ...); the data blocks of the first four
    super blocks are addressed from the index block, later ones through "EASB" super blocks."""
    src = tmp_path / "s.h5"
    tool("write-stress", src)
    f, chunks = _chunks_of(src)
    buf = bytearray(f.buf)
    entry, max_bits, idx_elmts, dmin, smin, page_bits = 8, 32, 4, 16, 4, 10
    off_size = 4
    nelmts = N_STRESS
    addr_of = lambda i: chunks[i][0] if i in chunks else h5dissect.UNDEF

    hdr_addr = len(buf)
    hdr_len = 12 + 48 + 8 + 4
    iblk_addr = hdr_addr + hdr_len
    ndblk_addrs, iblock_nsblks = 2 * (smin - 1), 2 * 2
    nsblks = 1 + (max_bits - 4)
    iblk_len = 4 + 2 + 8 + idx_elmts * entry + 8 * ndblk_addrs + 8 * (nsblks - iblock_nsblks) + 4
    cursor = iblk_addr + iblk_len
    blocks = b""

    def new_dblk(first, n):
        nonlocal cursor, blocks
        addr = cursor
        body = b"EADB" + bytes([0, 0]) + struct.pack("<Q", hdr_addr) + struct.pack("<I", first - idx_elmts)
        body += b"".join(struct.pack("<Q", addr_of(i) if i < nelmts else h5dissect.UNDEF) for i in range(first, first + n))
        body = CK(body)
        blocks += body
        cursor += len(body)
        return addr

    dblk_addrs, sblk_addrs = [], []
    first, s = idx_elmts, 0
    while first < nelmts:
        nd, ne = 1 << (s // 2), (1 << ((s + 1) // 2)) * dmin
        if s < iblock_nsblks:
            for _ in range(nd):
                dblk_addrs.append(new_dblk(first, ne) if first < nelmts else h5dissect.UNDEF)
                first += ne
        else:
            kids = []
            for _ in range(nd):
                kids.append(new_dblk(first, ne) if first < nelmts else h5dissect.UNDEF)
                first += ne
            sb = cursor
            body = b"EASB" + bytes([0, 0]) + struct.pack("<Q", hdr_addr) + struct.pack("<I", 0)
            body = CK(body + b"".join(struct.pack("<Q", k) for k in kids))
            blocks += body
            cursor += len(body)
            sblk_addrs.append(sb)
        s += 1
    dblk_addrs += [h5dissect.UNDEF] * (ndblk_addrs - len(dblk_addrs))
    sblk_addrs += [h5dissect.UNDEF] * (nsblks - iblock_nsblks - len(sblk_addrs))
    assert sblk_addrs[0] != h5dissect.UNDEF  # 300 elements reach the first super block outside the index block

    hdr = b"EAHD" + bytes([0, 0, entry, max_bits, idx_elmts, dmin, smin, page_bits])
    hdr = CK(hdr + struct.pack("<6Q", 0, 0, 0, 0, nelmts, nelmts) + struct.pack("<Q", iblk_addr))
    iblk = b"EAIB" + bytes([0, 0]) + struct.pack("<Q", hdr_addr)
    iblk += b"".join(struct.pack("<Q", addr_of(i)) for i in range(idx_elmts))
    iblk = CK(iblk + b"".join(struct.pack("<Q", a) for a in dblk_addrs) + b"".join(struct.pack("<Q", a) for a in sblk_addrs))
    assert len(hdr) == hdr_len and len(iblk) == iblk_len
    buf += hdr + iblk + blocks
    lay = bytes([4, 2, 0, 4, 2]) + struct.pack("<4H", 1, 4, 5, 2) + bytes([4, max_bits, idx_elmts, smin, dmin, page_bits])
    lay += struct.pack("<Q", hdr_addr)
    _patch_layout(buf, lay)
    struct.pack_into("<Q", buf, 40, len(buf))
    out = tmp_path / "ea.h5"
    out.write_bytes(buf)
    assert dump(out)["objects"]["/frames"]["index"] == 4
    want = stress_frames()
    assert np.array_equal(cat(out, "/frames", "<u2").reshape(N_STRESS, 4, 5), want)
    for i in (0, 3, 4, 19, 20, 51, 52, 83, 84, 115, 116, 179, 180, 243, 244, 299):
        r = tool("chunk", out, "/frames", i, 0, 0)
        assert np.array_equal(np.frombuffer(r.stdout, "<u2"), want[i].ravel()), i


# ---------------------------------------------------------------------------------------------
# H5Read (csrc/cli/reader_h5.hpp): what the frame loop sees of a NeXus file
# ---------------------------------------------------------------------------------------------
def h5read(path):
    r = tool("h5read", path)
    return json.loads(r.stdout.decode().strip().splitlines()[-1]), r.stdout.decode(), r.stderr.decode()


def make_scan(tmp_path, dtype=np.uint16, bits=16, n=7, per_file=3, mode="", omega=0.1):
    g = synth.small_geometry(96, 64)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, n, dtype, config=8, n_spots=5)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, bits, omega_increment=omega)
    args = ["stream2nxs", sd, tmp_path / "scan", per_file] + ([mode] if mode else [])
    tool(*args)
    return frames, mask


@pytest.mark.parametrize("mode", ["", "links", "external"])
def test_h5read_metadata_and_chunks(tmp_path, mode):
    frames, mask = make_scan(tmp_path, mode=mode)
    info, out, err = h5read(tmp_path / "scan.nxs")
    assert info["frames"] == 7 and info["shape"] == [64, 96] and info["element_size"] == 2
    assert info["trusted"] == [0, 60000]                      # underload_value, saturation_value (h5read.c:793-817)
    assert err == ""                                          # a complete NXmx file draws no warnings
    assert info["mask_valid"] == int(mask.sum())              # valid = (pixel_mask == 0) (h5read.c:615-633)
    assert "mask dtype uint32" in out and f"Mask has {mask.size} elements" in out and "Total data size: 7x64x96" in out
    assert info["wavelength"] == pytest.approx(0.976) and info["distance"] == pytest.approx(0.15)
    assert info["pixel_size_x"] == pytest.approx(7.5e-5) and info["beam_center_x"] == pytest.approx(48.0)
    assert info["beam_center_y"] == pytest.approx(32.0)
    assert info["oscillation"] == pytest.approx([10.0, 0.1], rel=1e-5)   # omega[0], omega[1] - omega[0] (:829-860)
    assert not info["raw_pixels"]
    want = [compress_chunk(f) for f in frames]
    assert info["available"] == [len(c) for c in want]
    for i in (0, 2, 3, 6):
        assert tool("h5read", tmp_path / "scan.nxs", i).stdout == want[i]


def test_h5read_stills_have_no_oscillation_and_32_bit(tmp_path):
    frames, mask = make_scan(tmp_path, dtype=np.uint32, bits=32, n=2, per_file=0, omega=None)
    info, _, _ = h5read(tmp_path / "scan.nxs")
    assert info["oscillation"] == [0.0, 0.0] and info["element_size"] == 4 and info["frames"] == 2


def test_h5read_unfiltered_files_hand_out_pixels(tmp_path):
    frames, mask = make_scan(tmp_path, n=3, per_file=2, mode="raw")
    info, _, _ = h5read(tmp_path / "scan.nxs")
    assert info["raw_pixels"] and info["available"] == [64 * 96 * 2] * 3
    for i in range(3):
        got = np.frombuffer(tool("h5read", tmp_path / "scan.nxs", i).stdout, np.uint16).reshape(64, 96)
        assert np.array_equal(got, frames[i])


def test_h5read_data_files_that_are_not_there_yet(tmp_path):
    """The master (with its VDS) exists before the data files do; a missing or half-written data
    file reads as 'image not available' (h5read.c:387-390,396-412), not as an error."""
    frames, mask = make_scan(tmp_path)
    second = tmp_path / "scan_000002.h5"
    blob = second.read_bytes()
    second.unlink()
    info, _, _ = h5read(tmp_path / "scan.nxs")
    sizes = [len(compress_chunk(f)) for f in frames]
    assert info["available"] == sizes[:3] + [0, 0, 0] + sizes[6:]
    assert tool("h5read", tmp_path / "scan.nxs", 4, check=False).returncode == 3
    # the file is there but its tail (chunk index and later chunks) is not on disk yet
    second.write_bytes(blob[:len(blob) // 2])
    info, _, _ = h5read(tmp_path / "scan.nxs")
    assert info["available"][:3] == sizes[:3] and info["available"][6:] == sizes[6:]
    assert all(a in (0, s) for a, s in zip(info["available"][3:6], sizes[3:6])) and 0 in info["available"][3:6]
    second.write_bytes(blob)
    info, _, _ = h5read(tmp_path / "scan.nxs")
    assert info["available"] == sizes


def test_h5read_rejects_what_it_cannot_serve(tmp_path):
    # not a NeXus file: no /entry/data
    tool("write-stress", tmp_path / "s.h5")
    r = tool("h5read", tmp_path / "s.h5", check=False)
    assert r.returncode == 1 and b"Error: Reading" in r.stderr
    # a data file whose frame size contradicts the master (h5read.c:356-375 exits 1)
    frames, mask = make_scan(tmp_path)
    other = tmp_path / "other"
    other.mkdir()
    g = synth.small_geometry(80, 64)
    m2 = synth.make_mask(g)
    f2 = synth.make_frames(g, m2, 3, np.uint16, config=8, n_spots=2)
    sd = other / "stream"
    sd.mkdir()
    write_stream(sd, f2, m2, 16)
    tool("stream2nxs", sd, other / "scan", 3)
    (tmp_path / "scan_000002.h5").write_bytes((other / "scan_000001.h5").read_bytes())
    r = tool("h5read", tmp_path / "scan.nxs", check=False)
    assert r.returncode == 1 and b"Validation Error" in r.stderr and b"fast data has 80 pixels, expected 96" in r.stderr


def test_external_links_both_parsers(tmp_path):
    """The master of `stream2nxs ... external`: /entry/data is a compact (link-message) group holding
    external links; the VDS names its sources as file "." + those links (h5read.c:945-985)."""
    frames, mask = make_scan(tmp_path, mode="external")
    m = h5dissect.File(tmp_path / "scan.nxs")
    links = m.get("/entry/data").links()
    assert links["data_000002"] == ("external", "scan_000002.h5", "/data") and links["data"][0] == "hard"
    d = dump(tmp_path / "scan.nxs")["objects"]["/entry/data"]["links"]
    assert d["data_000003"] == {"external": ["scan_000003.h5", "/data"]} and "hard" in d["data"]
    maps = json.loads(tool("vds", tmp_path / "scan.nxs", "/entry/data/data").stdout)
    assert [(x["file"], x["dataset"]) for x in maps] == [(".", f"/entry/data/data_{k:06d}") for k in (1, 2, 3)]
    # the C++ resolver follows the link into the other file
    want = compress_chunk(frames[4])
    assert tool("chunk", tmp_path / "scan.nxs", "/entry/data/data_000002", 1, 0, 0).stdout == want
    # a data file that is missing: frames behind its link poll as not available
    (tmp_path / "scan_000003.h5").unlink()
    info, _, _ = h5read(tmp_path / "scan.nxs")
    assert info["available"][6] == 0 and all(info["available"][:6])


@pytest.mark.parametrize("layout", ["root-direct", "root-indirect"])
def test_dense_link_storage(tmp_path, layout):
    """A new-format group with more than 8 links: link messages live in a fractal heap and are
    indexed by a version 2 B-tree (what a libver='latest' / SWMR writer produces for e.g.
    /entry/instrument/detector).  Built here from the format specification: 30 hard links,
    heap with a direct root block / with a two-row indirect root block, B-tree of depth 0 / 1."""
    src = tmp_path / "s.h5"
    tool("write-stress", src)
    f = h5dissect.File(src)
    buf = bytearray(f.buf)
    targets = {f"f{i:02d}": f.get(f"/many/d{i * 7:03d}").addr for i in range(30)}
    msgs = {n: bytes([1, 0, len(n)]) + n.encode() + struct.pack("<Q", a) for n, a in targets.items()}

    base = len(buf)
    ohdr_addr = base
    heap_addr = ohdr_addr + 64
    bt_addr = heap_addr + 160
    blocks_addr = bt_addr + 64
    # --- fractal heap: objects packed behind each direct block's 21-byte prefix
    hdr_len = 5 + 8 + 4 + 4
    names = sorted(msgs)
    if layout == "root-direct":
        placement = {0: names}                                   # block index -> names
        cur_rows = 0
    else:
        placement = {0: names[:12], 1: names[12:20], 5: names[20:]}   # row 0: blocks 0-3, row 1: blocks 4-7
        cur_rows = 2
    ids, direct = {}, {}
    for k, ns in placement.items():
        data = b""
        for n in ns:
            off = k * 512 + hdr_len + len(data)
            ids[n] = bytes([0]) + struct.pack("<I", off) + struct.pack("<H", len(msgs[n]))
            data += msgs[n]
        blk = b"FHDB" + bytes([0]) + struct.pack("<Q", heap_addr) + struct.pack("<I", k * 512) + bytes(4) + data
        direct[k] = blk + bytes(512 - len(blk))
    layout_bytes = b""
    block_addr = {}
    for k in sorted(direct):
        block_addr[k] = blocks_addr + len(layout_bytes)
        layout_bytes += direct[k]
    if cur_rows:
        iblk_addr = blocks_addr + len(layout_bytes)
        iblk = b"FHIB" + bytes([0]) + struct.pack("<Q", heap_addr) + struct.pack("<I", 0)
        iblk += b"".join(struct.pack("<Q", block_addr.get(k, h5dissect.UNDEF)) for k in range(8)) + bytes(4)
        layout_bytes += iblk
        root = iblk_addr
    else:
        root = block_addr[0]
    heap = b"FRHP" + bytes([0]) + struct.pack("<HHB", 7, 0, 2) + struct.pack("<I", 4096)
    heap += struct.pack("<12Q", 0, h5dissect.UNDEF, 0, h5dissect.UNDEF, 0, 0, 0, 30, 0, 0, 0, 0)
    heap += struct.pack("<HQQHH", 4, 512, 65536, 32, 1) + struct.pack("<QH", root, cur_rows) + bytes(4)
    assert len(heap) == 146
    # --- version 2 B-tree, record type 5: {hash u32 (not used by the reader), heap id (7 bytes)}
    rec = lambda n: bytes(4) + ids[n]
    nodes_addr = blocks_addr + len(layout_bytes)
    if layout == "root-direct":
        leaf = CK(b"BTLF" + bytes([0, 5]) + b"".join(rec(n) for n in names))
        nodes = leaf
        bt_root, root_nrec, depth = nodes_addr, 30, 0
    else:
        left, mid, right = names[:15], names[15], names[16:]
        leaf1 = CK(b"BTLF" + bytes([0, 5]) + b"".join(rec(n) for n in left))
        leaf2 = CK(b"BTLF" + bytes([0, 5]) + b"".join(rec(n) for n in right))
        a1, a2 = nodes_addr, nodes_addr + len(leaf1)
        internal = CK(b"BTIN" + bytes([0, 5]) + rec(mid) + struct.pack("<QB", a1, len(left)) + struct.pack("<QB", a2, len(right)))
        nodes = leaf1 + leaf2 + internal
        bt_root, root_nrec, depth = a2 + len(leaf2), 1, 1
    bt = CK(b"BTHD" + bytes([0, 5]) + struct.pack("<IHHBB", 512, 11, depth, 100, 40) + struct.pack("<QHQ", bt_root, root_nrec, 30))
    # --- the group's object header (version 2): link info + group info
    linfo = bytes([0, 0]) + struct.pack("<QQ", heap_addr, bt_addr)
    body = bytes([2]) + struct.pack("<H", len(linfo)) + bytes([0]) + linfo + bytes([10]) + struct.pack("<H", 2) + bytes([1, 0, 0])
    ohdr = CK(b"OHDR" + bytes([2, 0, len(body)]) + body)
    assert len(ohdr) <= 64 and len(heap) <= 160 and len(bt) <= 64
    buf += ohdr + bytes(64 - len(ohdr)) + heap + bytes(160 - len(heap)) + bt + bytes(64 - len(bt)) + layout_bytes + nodes
    # --- hang it where /links was
    old = struct.pack("<Q", f.get("/links").addr)
    snod_hits = [i for i in range(len(f.buf) - 8) if f.buf[i:i + 8] == old]
    assert len(snod_hits) == 1
    buf[snod_hits[0]:snod_hits[0] + 8] = struct.pack("<Q", ohdr_addr)
    struct.pack_into("<I", buf, snod_hits[0] + 8, 0)  # cache type 0: nothing cached in the entry
    struct.pack_into("<Q", buf, 40, len(buf))
    out = tmp_path / f"dense_{layout}.h5"
    out.write_bytes(buf)

    d = dump(out)["objects"]
    assert sorted(d["/links"]["links"]) == names
    assert {n: v["hard"] for n, v in d["/links"]["links"].items()} == targets
    for i in (0, 11, 12, 19, 20, 29):
        assert cat(out, f"/links/f{i:02d}", "<i4").tolist() == [i * 7, -i * 7]
