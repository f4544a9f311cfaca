// h5lite_tool -- command-line front end of h5lite.hpp (host only, no CUDA).
//
//   h5lite_tool dump FILE                 JSON description of every object reachable by hard links
//   h5lite_tool cat FILE DATASET          raw element bytes of an unfiltered dataset on stdout
//   h5lite_tool chunk FILE DATASET I [J K ...]   the stored bytes of one chunk (filtered or not)
//   h5lite_tool vds FILE DATASET          the mapping list of a virtual dataset, as JSON
//   h5lite_tool write-empty OUT           what H5Fcreate + H5Fclose leave behind
//   h5lite_tool write-table OUT N [2d]    a results_ffs.h5-shaped reflection table with N rows
//   h5lite_tool h5read FILE.nxs [I]       what H5Read (reader_h5.hpp) makes of a NeXus file: metadata and
//                                         per-frame availability / stored sizes as JSON; with I, frame
//                                         I's stored bytes (chunk, or pixels for an unfiltered file)
//   h5lite_tool write-stress OUT          a file that exercises the writer: 300 datasets in one group
//                                         (two B-tree levels), a 300-chunk dataset, ragged edge chunks,
//                                         soft links, fixed-length string and numeric attributes
//   h5lite_tool selection HEX             decode one serialised dataspace selection
//   h5lite_tool stream2nxs DIR PREFIX [FRAMES_PER_FILE [raw|links|external]]
//                                         an Eiger stream directory (start_1, start_5, image_*_2) as
//                                         PREFIX.nxs (NXmx metadata + VDS) and PREFIX_NNNNNN.h5 data
//                                         files (one bitshuffle-LZ4 chunk per frame, filter 32008;
//                                         `raw`: decoded pixels, no filter; `links`: no VDS, frames
//                                         in the master as /entry/data/data_NNNNNN; `external`:
//                                         VDS sources reached through external links in the master)
//
// tests/test_h5lite.py drives it; `stream2nxs` is also how a recorded stream is turned into the
// file layout the reference's H5Read opens (h5read.c:1088-1127).
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "bslz4.hpp"
#include "h5lite.hpp"
#include "h5_results.hpp"
#include "json_mini.hpp"
#include "reader_h5.hpp"

using namespace h5lite;

static std::string jstr(const std::string& s) { return jsonmini::escape(s); }

static std::string dtype_name(const Datatype& d) {
    char b[64];
    switch (d.cls) {
        case Datatype::Fixed: snprintf(b, sizeof b, "%c%s%u", d.big_endian ? '>' : '<', d.is_signed ? "i" : "u", d.size); break;
        case Datatype::Float: snprintf(b, sizeof b, "%cf%u", d.big_endian ? '>' : '<', d.size); break;
        case Datatype::String: snprintf(b, sizeof b, "S%u", d.size); break;
        case Datatype::Vlen: snprintf(b, sizeof b, d.vlen_string ? "vlen-str" : "vlen"); break;
        case Datatype::Enum: snprintf(b, sizeof b, "enum%u", d.size); break;
        default: snprintf(b, sizeof b, "class%u/%u", d.cls, d.size);
    }
    return b;
}

static std::string dims_json(const std::vector<uint64_t>& v) {
    std::string s = "[";
    for (size_t i = 0; i < v.size(); ++i) s += (i ? "," : "") + std::to_string(v[i]);
    return s + "]";
}

static std::string attr_json(const File& f, const Attribute& a) {
    std::string s = "{\"dtype\":" + jstr(dtype_name(a.dtype)) + ",\"shape\":" + (a.scalar ? "null" : dims_json(a.dims)) +
                    ",\"value\":[";
    if (a.dtype.cls == Datatype::String || (a.dtype.cls == Datatype::Vlen && a.dtype.vlen_string)) {
        const auto v = attribute_strings(f, a);
        for (size_t i = 0; i < v.size(); ++i) s += (i ? "," : "") + jstr(v[i]);
    } else if (a.dtype.cls == Datatype::Float) {
        const auto v = attribute_numbers<double>(a);
        for (size_t i = 0; i < v.size(); ++i) s += (i ? "," : "") + jsonmini::number(v[i]);
    } else if (a.dtype.numeric()) {
        const auto v = attribute_numbers<int64_t>(a);
        for (size_t i = 0; i < v.size(); ++i) s += (i ? "," : "") + std::to_string(v[i]);
    }
    return s + "]}";
}

static void dump_object(const File& f, const std::string& path, const Object& o, std::string& out, bool& first, int depth) {
    if (depth > 64) throw Error("h5lite_tool: hierarchy too deep (link cycle?)");
    out += first ? "" : ",\n";
    first = false;
    out += jstr(path) + ":{\"address\":" + std::to_string(o.address());
    const auto attrs = o.attributes();
    out += ",\"attrs\":{";
    bool fa = true;
    for (const auto& kv : attrs) {
        out += (fa ? "" : ",") + jstr(kv.first) + ":" + attr_json(f, kv.second);
        fa = false;
    }
    out += "}";
    if (o.is_dataset()) {
        const Layout l = o.layout();
        static const char* cls[] = {"compact", "contiguous", "chunked", "virtual"};
        out += ",\"kind\":\"dataset\",\"dtype\":" + jstr(dtype_name(o.dtype())) + ",\"shape\":" + dims_json(o.shape()) +
               ",\"layout\":" + jstr(cls[l.cls]);
        if (l.cls == Layout::Chunked) out += ",\"chunk\":" + dims_json(l.chunk) + ",\"index\":" + std::to_string(l.index);
        out += ",\"filters\":[";
        const auto fl = o.filters();
        for (size_t i = 0; i < fl.size(); ++i) {
            out += (i ? "," : "") + std::string("{\"id\":") + std::to_string(fl[i].id) + ",\"name\":" + jstr(fl[i].name) +
                   ",\"client_data\":[";
            for (size_t k = 0; k < fl[i].client_data.size(); ++k) out += (k ? "," : "") + std::to_string(fl[i].client_data[k]);
            out += "]}";
        }
        out += "]}";
        return;
    }
    out += ",\"kind\":\"group\",\"links\":{";
    const auto links = o.links();
    bool fl = true;
    for (const auto& kv : links) {
        out += (fl ? "" : ",") + jstr(kv.first) + ":";
        fl = false;
        if (kv.second.kind == Link::Hard) out += "{\"hard\":" + std::to_string(kv.second.addr) + "}";
        else if (kv.second.kind == Link::Soft) out += "{\"soft\":" + jstr(kv.second.path) + "}";
        else out += "{\"external\":[" + jstr(kv.second.file) + "," + jstr(kv.second.path) + "]}";
    }
    out += "}}";
    for (const auto& kv : links)
        if (kv.second.kind == Link::Hard)
            dump_object(f, (path == "/" ? "" : path) + "/" + kv.first, f.object_at(kv.second.addr), out, first, depth + 1);
}

static int cmd_dump(const char* file) {
    File f(file);
    std::string out = "{\"superblock\":" + std::to_string(f.superblock_version()) + ",\"eof\":" + std::to_string(f.eof_address()) +
                      ",\"objects\":{\n";
    bool first = true;
    dump_object(f, "/", f.root(), out, first, 0);
    out += "\n}}\n";
    fwrite(out.data(), 1, out.size(), stdout);
    return 0;
}

static int cmd_stream2nxs(const std::string& dir, const std::string& prefix, size_t per_file, const std::string& mode);
static int cmd_write_stress(const char* out);
static int cmd_h5read(const char* file, const char* frame);

int main(int argc, char** argv) {
    try {
        const std::string cmd = argc > 1 ? argv[1] : "";
        if (cmd == "dump" && argc == 3) return cmd_dump(argv[2]);
        if (cmd == "cat" && argc == 4) {
            const auto fp = File::open(argv[2]);
            const Node n = resolve(fp, argv[3]);
            const auto raw = n.obj.read_raw();
            fwrite(raw.data(), 1, raw.size(), stdout);
            return 0;
        }
        if (cmd == "chunk" && argc >= 5) {
            const auto fp = File::open(argv[2]);
            const Node n = resolve(fp, argv[3]);
            std::vector<uint64_t> off;
            for (int i = 4; i < argc; ++i) off.push_back(strtoull(argv[i], nullptr, 10));
            const ChunkInfo c = n.obj.find_chunk(off);
            if (!c.found) {
                fprintf(stderr, "chunk not allocated\n");
                return 3;
            }
            fprintf(stderr, "address %" PRIu64 " size %" PRIu64 " filter_mask %u\n", c.addr, c.size, c.filter_mask);
            fwrite(n.file->at(c.addr, c.size), 1, c.size, stdout);
            return 0;
        }
        if (cmd == "vds" && argc == 4) {
            File f(argv[2]);
            const auto maps = f.get(argv[3]).vds_mappings();
            std::string out = "[";
            for (size_t i = 0; i < maps.size(); ++i) {
                const auto& m = maps[i];
                out += (i ? ",\n" : "") + std::string("{\"file\":") + jstr(m.file) + ",\"dataset\":" + jstr(m.dataset) +
                       ",\"src_all\":" + (m.src_all ? "true" : "false") + ",\"src_start\":" + dims_json(m.src_start) +
                       ",\"src_block\":" + dims_json(m.src_block) + ",\"dst_start\":" + dims_json(m.dst_start) +
                       ",\"dst_block\":" + dims_json(m.dst_block) + "}";
            }
            printf("%s]\n", out.c_str());
            return 0;
        }
        if (cmd == "write-empty" && argc == 3) {
            Writer w;
            w.save(argv[2]);
            return 0;
        }
        if (cmd == "write-table" && argc >= 4) {
            const size_t n = strtoull(argv[3], nullptr, 10);
            const bool two_d = argc > 4 && std::string(argv[4]) == "2d";
            h5results::Table t;
            std::vector<double> xyz(3 * n), vb(n), vm(n);
            std::vector<int> id(n), ext(n);
            for (size_t i = 0; i < n; ++i) {
                xyz[3 * i] = 0.5 + (double)i, xyz[3 * i + 1] = 100.25 + 2.0 * i, xyz[3 * i + 2] = (double)(i % 7) + 0.125;
                vb[i] = 1e-6 * (i + 1), vm[i] = 2e-6 * (i + 1), ext[i] = (int)(i % 9) + 1;
                id[i] = two_d ? (int)(i / 3) : 0;
            }
            if (two_d)
                for (size_t k = 1; k < (n + 2) / 3; ++k) t.generate_new_attributes();
            t.add_column("xyzobs.px.value", n, 3, xyz);
            t.add_column("id", n, 1, id);
            if (!two_d) {
                t.add_column("sigma_b_variance", n, 1, vb);
                t.add_column("sigma_m_variance", n, 1, vm);
                t.add_column("spot_extent_z", n, 1, ext);
            }
            t.write(argv[2], "dials/processing/group_0");
            return 0;
        }
        if (cmd == "h5read" && argc >= 3) return cmd_h5read(argv[2], argc > 3 ? argv[3] : nullptr);
        if (cmd == "write-stress" && argc == 3) return cmd_write_stress(argv[2]);
        if (cmd == "selection" && argc == 3) {  // a serialised dataspace selection given as hex
            std::vector<uint8_t> b;
            for (const char* p = argv[2]; p[0] && p[1]; p += 2) b.push_back((uint8_t)strtoul(std::string(p, 2).c_str(), nullptr, 16));
            std::vector<uint64_t> start, block;
            bool all = false;
            const size_t used = detail::parse_selection(b.data(), b.size(), start, block, all);
            printf("{\"used\":%zu,\"all\":%s,\"start\":%s,\"block\":%s}\n", used, all ? "true" : "false",
                   dims_json(start).c_str(), dims_json(block).c_str());
            return 0;
        }
        if (cmd == "stream2nxs" && argc >= 4)
            return cmd_stream2nxs(argv[2], argv[3], argc > 4 ? strtoull(argv[4], nullptr, 10) : 0,
                                  argc > 5 ? argv[5] : "");
        fprintf(stderr, "usage: h5lite_tool dump|cat|chunk|vds|write-empty|write-table|stream2nxs ... (see the source header)\n");
        return 2;
    } catch (const std::exception& e) {
        fprintf(stderr, "h5lite_tool: %s\n", e.what());
        return 1;
    }
}

// ---- stream2nxs ----------------------------------------------------------------------------
static std::vector<uint8_t> slurp(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw Error("cannot read " + path);
    return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

static int cmd_stream2nxs(const std::string& dir, const std::string& prefix, size_t per_file, const std::string& mode) {
    const bool raw = mode == "raw";
    // "links": no VDS and no separate files -- the frames sit in the master as /entry/data/data_000001,
    // data_000002, ... (the group layout of the Eiger file writer)
    const bool links = mode == "links";
    // "external": the master reaches the data files through external links /entry/data/data_NNNNNN,
    // and the VDS names them as sources in file "." -- the case h5read.c:945-985 dereferences
    const bool external = mode == "external";
    Writer master;
    const auto hdr = slurp(dir + "/start_1");
    const jsonmini::Value j = jsonmini::Parser::parse(std::string(hdr.begin(), hdr.end()));
    const size_t n = (size_t)j.at("nimages").number() * (size_t)j.at("ntrigger").number();
    const uint64_t H = (uint64_t)j.at("y_pixels_in_detector").number(), W = (uint64_t)j.at("x_pixels_in_detector").number();
    const int bits = (int)j.at("bit_depth_image").number();
    const Datatype pix = bits == 32 ? Datatype::u32() : Datatype::u16();
    if (!per_file) per_file = n;
    const std::string base = prefix.substr(prefix.find_last_of('/') == std::string::npos ? 0 : prefix.find_last_of('/') + 1);

    Filter bs;
    bs.id = FILTER_BSLZ4, bs.flags = 0, bs.name = "bitshuffle; see https://github.com/kiyo-masui/bitshuffle";
    bs.client_data = {0, 3, (uint32_t)pix.size, 0, 2};  // major, minor, element size, block size (0 = default), LZ4
    std::vector<VdsMapping> maps;
    size_t nfiles = 0;
    for (size_t f0 = 0; f0 < n; f0 += per_file, ++nfiles) {
        const size_t cnt = std::min(per_file, n - f0);
        char suffix[32];
        snprintf(suffix, sizeof suffix, "_%06zu.h5", nfiles + 1);
        Writer file_writer;
        Writer& w = links ? master : file_writer;
        char dname[32];
        snprintf(dname, sizeof dname, "data_%06zu", nfiles + 1);
        auto* d = w.chunked_dataset(w.group(links ? "/entry/data" : "/"), links ? dname : "data", pix, {cnt, H, W}, {1, H, W},
                                    raw ? std::vector<Filter>{} : std::vector<Filter>{bs});
        std::vector<uint8_t> pixels(H * W * pix.size), scratch;
        for (size_t i = 0; i < cnt; ++i) {
            char name[64];
            snprintf(name, sizeof name, "/image_%06zu_2", f0 + i);
            const auto chunk = slurp(dir + name);
            if (raw) {  // decoded pixels, no filter
                if (bslz4::decompress_chunk(chunk.data(), chunk.size(), pixels.data(), H * W, pix.size, scratch) != 0)
                    throw Error("cannot decode " + dir + name);
                Writer::add_chunk(d, {i, 0, 0}, 0, pixels.data(), pixels.size());
            } else {
                Writer::add_chunk(d, {i, 0, 0}, 0, chunk.data(), chunk.size());
            }
        }
        if (!links) w.save(prefix + suffix);
        VdsMapping m;
        m.file = base + suffix, m.dataset = "/data";
        if (external) {
            Writer::external_link(master.group("/entry/data"), dname, base + suffix, "/data");
            m.file = ".", m.dataset = std::string("/entry/data/") + dname;
        }
        m.src_start = {0, 0, 0}, m.src_block = {cnt, H, W};
        m.dst_start = {f0, 0, 0}, m.dst_block = {cnt, H, W};
        maps.push_back(m);
    }

    Writer& w = master;
    auto* entry = w.group("/entry");
    Writer::attr_string(entry, "NX_class", "NXentry");
    auto* data = w.group("/entry/data");
    Writer::attr_string(data, "NX_class", "NXdata");
    if (!links) w.virtual_dataset(data, "data", pix, {n, H, W}, maps);
    auto* det = w.group("/entry/instrument/detector");
    Writer::attr_string(det, "NX_class", "NXdetector");
    auto scalar_f = [&](Writer::Node* g, const char* name, double v, const char* units) {
        const float f = (float)v;
        auto* ds = w.dataset(g, name, Datatype::f32(), {}, &f);
        if (units) Writer::attr_string(ds, "units", units);
    };
    scalar_f(det, "x_pixel_size", j.at("x_pixel_size").number(), "m");
    scalar_f(det, "y_pixel_size", j.at("y_pixel_size").number(), "m");
    scalar_f(det, "beam_center_x", j.at("beam_center_x").number(), "pixel");
    scalar_f(det, "beam_center_y", j.at("beam_center_y").number(), "pixel");
    scalar_f(det, "distance", j.at("detector_distance").number() / 1000.0, "m");
    {
        const uint64_t cutoff = (uint64_t)j.at("countrate_correction_count_cutoff").number();
        if (bits == 32) {
            const uint32_t v = (uint32_t)cutoff;
            w.dataset(det, "saturation_value", Datatype::u32(), {}, &v);
        } else {
            const uint16_t v = (uint16_t)std::min<uint64_t>(cutoff, 65535);
            w.dataset(det, "saturation_value", Datatype::u16(), {}, &v);
        }
    }
    {   // NXmx underload_value: counts below it are invalid (h5read.c:809-816 reads it, default 0)
        const int32_t under = 0;
        w.dataset(det, "underload_value", Datatype::i32(), {}, &under);
    }
    {   // start_5: int32 mask, non-zero = masked -> NXmx pixel_mask (uint32, non-zero = masked)
        const auto m = slurp(dir + "/start_5");
        if (m.size() != H * W * 4) throw Error("start_5 has the wrong size");
        w.dataset(det, "pixel_mask", Datatype::u32(), {H, W}, m.data());
    }
    if (j.contains("wavelength")) scalar_f(w.group("/entry/instrument/beam"), "incident_wavelength", j.at("wavelength").number(), "angstrom");
    if (j.contains("omega_start") && j.contains("omega_increment") && j.at("omega_increment").number() > 0) {
        std::vector<double> omega(n);
        for (size_t i = 0; i < n; ++i) omega[i] = j.at("omega_start").number() + i * j.at("omega_increment").number();
        w.dataset(w.group("/entry/sample/sample_omega"), "omega", Datatype::f64(), {n}, omega.data());
    }
    w.save(prefix + ".nxs");
    printf("wrote %s.nxs and %zu data file(s)\n", prefix.c_str(), nfiles);
    return 0;
}

// ---- write-stress --------------------------------------------------------------------------
static int cmd_write_stress(const char* out) {
    Writer w;
    auto* many = w.group("/many");
    for (int i = 0; i < 300; ++i) {
        char name[32];
        snprintf(name, sizeof name, "d%03d", i);
        const int32_t v[2] = {i, -i};
        w.dataset(many, name, Datatype::i32(), {2}, v);
    }
    // 300 chunks of one 4 x 5 frame each: five leaf nodes under one internal node (K = 32)
    auto* frames = w.chunked_dataset(w.group("/"), "frames", Datatype::u16(), {300, 4, 5}, {1, 4, 5});
    for (uint64_t f = 0; f < 300; ++f) {
        uint16_t px[20];
        for (int k = 0; k < 20; ++k) px[k] = (uint16_t)(f * 20 + k);
        if (f % 50 != 7) Writer::add_chunk(frames, {f, 0, 0}, 0, px, sizeof px);  // a few chunks never written
    }
    // ragged: 7 x 10 in chunks of 3 x 4 (edge chunks hang over the extent)
    auto* ragged = w.chunked_dataset(w.group("/"), "ragged", Datatype::f64(), {7, 10}, {3, 4});
    for (uint64_t r = 0; r < 7; r += 3)
        for (uint64_t c = 0; c < 10; c += 4) {
            double t[12];
            for (int i = 0; i < 3; ++i)
                for (int k = 0; k < 4; ++k) t[i * 4 + k] = (double)((r + i) * 10 + (c + k)) + 0.5;
            Writer::add_chunk(ragged, {r, c}, 0, t, sizeof t);
        }
    Writer::soft_link(w.group("/"), "alias", "/many/d042");
    Writer::soft_link(w.group("/links"), "relative", "./scalar_u8");  // relative to /links
    auto* g = w.group("/links");
    const char fixed[8] = {'N', 'X', 'd', 'a', 't', 'a', 0, 0};
    Writer::attr_raw(g, "fixed", Datatype::fixed_string(8), {}, fixed);
    const float f3[3] = {1.5f, -2.25f, 1e-3f};
    Writer::attr_raw(g, "f32x3", Datatype::f32(), {3}, f3);
    const int64_t big = -1234567890123LL;
    Writer::attr_raw(g, "i64", Datatype::i64(), {}, &big);
    Writer::attr_strings(g, "names", {"alpha", "", "gamma-delta"});
    const uint8_t scalar8 = 200;
    w.dataset(g, "scalar_u8", Datatype::u8(), {}, &scalar8);
    w.dataset(g, "empty", Datatype::f64(), {0}, nullptr);
    w.save(out);
    return 0;
}

// ---- h5read --------------------------------------------------------------------------------
static int cmd_h5read(const char* file, const char* frame) {
    FILE* real_stdout = stdout;
    if (frame) stdout = stderr;  // the reader's own chatter must not mix with the binary output
    H5Read r(file);
    stdout = real_stdout;
    const auto shape = r.image_shape();
    const size_t px = shape[0] * shape[1];
    if (frame) {
        const size_t i = strtoull(frame, nullptr, 10);
        std::vector<uint8_t> buf(px * r.get_element_size() + 65536);
        if (!r.is_image_available(i)) return 3;
        if (r.chunks_are_raw_pixels()) {
            if (!r.read_pixels(i, buf.data(), px * r.get_element_size())) return 3;
            fwrite(buf.data(), 1, px * r.get_element_size(), stdout);
        } else {
            const size_t n = r.read_raw_chunk(i, buf.data(), buf.size());
            if (n == 0 || n == SIZE_MAX) return 3;
            fwrite(buf.data(), 1, n, stdout);
        }
        return 0;
    }
    size_t valid = 0;
    for (uint8_t m : r.get_mask()) valid += m;
    std::string out = "{\"frames\":" + std::to_string(r.get_number_of_images()) + ",\"shape\":[" + std::to_string(shape[0]) + "," +
                      std::to_string(shape[1]) + "],\"element_size\":" + std::to_string(r.get_element_size()) +
                      ",\"trusted\":[" + std::to_string(r.get_trusted_range()[0]) + "," + std::to_string(r.get_trusted_range()[1]) +
                      "],\"mask_valid\":" + std::to_string(valid) + ",\"raw_pixels\":" + (r.chunks_are_raw_pixels() ? "true" : "false");
    auto opt = [&](const char* k, std::optional<float> v) { out += std::string(",\"") + k + "\":" + (v ? jsonmini::number(*v) : "null"); };
    opt("wavelength", r.get_wavelength());
    opt("distance", r.get_detector_distance());
    const auto ps = r.get_pixel_size(), bc = r.get_beam_center();
    opt("pixel_size_y", ps ? std::optional<float>((*ps)[0]) : std::nullopt);
    opt("pixel_size_x", ps ? std::optional<float>((*ps)[1]) : std::nullopt);
    opt("beam_center_y", bc ? std::optional<float>((*bc)[0]) : std::nullopt);
    opt("beam_center_x", bc ? std::optional<float>((*bc)[1]) : std::nullopt);
    out += ",\"oscillation\":[" + jsonmini::number(r.get_oscillation()[0]) + "," + jsonmini::number(r.get_oscillation()[1]) + "]";
    out += ",\"available\":[";
    std::vector<uint8_t> chunk;
    for (size_t i = 0; i < r.get_number_of_images(); ++i) {
        size_t n = 0;
        if (r.is_image_available(i)) {
            if (r.chunks_are_raw_pixels()) n = px * r.get_element_size();
            else if (r.get_raw_chunk(i, chunk)) n = chunk.size();
        }
        out += (i ? "," : "") + std::to_string(n);
    }
    printf("%s]}\n", out.c_str());
    return 0;
}
