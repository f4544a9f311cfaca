// reader_h5.hpp -- H5Read: NeXus / HDF5 frame ingestion on h5lite (no libhdf5).
//
// Mirrors the reference's H5Read (h5read/include/h5read.h:206-320, h5read/src/h5read.c):
//   * h5read_open (h5read.c:1088-1127): master file, `/entry/data/data` as a virtual dataset whose
//     mapping list names the data files (vds_info, :909-1007; "." + external link is
//     dereferenced as :945-985 does), trusted range from saturation_value / underload_value
//     (:793-817), wavelength (:819-827), detector metadata (:862-907), omega (:829-860),
//     pixel_mask with 4- or 8-byte elements, valid = (value == 0) (:561-650);
//   * is_image_available / get_raw_chunk (h5read.h:229-247, h5read.c:379-456): the stored chunk
//     of one frame, copied without decoding (H5Dget_chunk_storage_size + H5Dread_chunk); a data
//     file or chunk that is not there yet reads as "not available" and is looked for again
//     (H5Drefresh -> File::refresh) -- the SWMR behaviour the frame loop polls on.
// Beyond the reference: `/entry/data/data` may also be a plain 3-D dataset, or absent with
// `/entry/data/data_NNNNNN` links present (the layout the Eiger file writer produces), and
// datasets without the bitshuffle filter are served as raw pixels (read_pixels).
#pragma once
#include <mutex>
#include <shared_mutex>

#include "h5lite.hpp"
#include "reader.hpp"

class H5Read : public Reader {
    struct DataFile {
        std::string filename, dsetname;  // filename empty = the master file
        size_t frames = 0;               // frames this file contributes
        size_t first = 0;                // index of its first frame in the whole scan
        size_t offset = 0;               // index of that frame inside the source dataset
        std::shared_ptr<h5lite::File> file;
        h5lite::Object dataset;
        bool open = false;
        bool filtered = false;
        h5lite::Layout layout;
    };
    std::shared_ptr<h5lite::File> master_;
    std::vector<DataFile> files_;
    mutable std::shared_mutex mutex_;  // shared: chunk copies; exclusive: open / refresh
    size_t frames_ = 0, slow_ = 0, fast_ = 0, elem_size_ = 0;
    bool raw_pixels_ = false;
    std::array<int64_t, 2> trusted_{};
    std::vector<uint8_t> mask_;
    std::optional<float> wavelength_, distance_;
    std::optional<std::array<float, 2>> pixel_size_, beam_center_;
    std::array<float, 2> oscillation_{0, 0};

    std::optional<double> scalar(const std::string& path) const {
        try {
            const auto n = h5lite::resolve(master_, path);
            const auto v = n.obj.read_as<double>();
            if (v.size() != 1) {
                fprintf(stderr, "Error: While reading %s: More than one element.\n", path.c_str());
                return std::nullopt;
            }
            return v[0];
        } catch (const h5lite::Error&) {
            return std::nullopt;
        }
    }

    void describe(DataFile& d, const h5lite::Node& n) {
        d.file = n.file, d.dataset = n.obj;
        const auto dims = d.dataset.shape();
        if (dims.size() != 3) throw std::runtime_error("Error: data in " + d.file->path() + " is not three dimensional");
        if (slow_ == 0) slow_ = dims[1], fast_ = dims[2], elem_size_ = d.dataset.dtype().size;
        if (dims[1] != slow_)  // h5read.c:356-375
            throw std::runtime_error("Validation Error: Data file " + d.file->path() + " slow data has " +
                                     std::to_string(dims[1]) + " pixels, expected " + std::to_string(slow_));
        if (dims[2] != fast_)
            throw std::runtime_error("Validation Error: Data file " + d.file->path() + " fast data has " +
                                     std::to_string(dims[2]) + " pixels, expected " + std::to_string(fast_));
        if (d.dataset.dtype().size != elem_size_) throw std::runtime_error("Error: data files differ in pixel type");
        d.layout = d.dataset.layout();
        d.filtered = false;
        for (const auto& f : d.dataset.filters()) {
            if (f.id != h5lite::FILTER_BSLZ4)
                throw std::runtime_error("Error: " + d.file->path() + ": filter " + std::to_string(f.id) +
                                         " is not handled (bitshuffle-LZ4, id 32008, or none)");
            if (f.client_data.size() > 4 && f.client_data[4] != 2)
                throw std::runtime_error("Error: " + d.file->path() + ": bitshuffle without LZ4 is not handled");
            d.filtered = true;
        }
        if (d.layout.cls == h5lite::Layout::Chunked &&
            (d.layout.chunk.size() != 3 || d.layout.chunk[0] != 1 || d.layout.chunk[1] != slow_ || d.layout.chunk[2] != fast_))
            throw std::runtime_error("Error: " + d.file->path() + ": direct chunk read needs one whole frame per chunk");
        if (d.layout.cls != h5lite::Layout::Chunked && d.layout.cls != h5lite::Layout::Contiguous)
            throw std::runtime_error("Error: " + d.file->path() + ": unsupported data layout");
        d.open = true;
    }

    // Opens a data file if it exists by now.  Caller holds the exclusive lock.
    bool try_open(DataFile& d) {
        if (d.open) return true;
        try {
            if (d.filename.empty()) {
                describe(d, h5lite::resolve(master_, d.dsetname));
            } else {
                struct stat st;
                if (stat(d.filename.c_str(), &st) != 0) return false;  // not written yet
                describe(d, h5lite::resolve(h5lite::File::open(d.filename), d.dsetname));
            }
        } catch (const h5lite::Error&) {
            return false;  // half-written file: look again on the next poll
        } catch (const std::runtime_error& e) {  // a data file that contradicts the master: h5read.c:356-375 exits
            fprintf(stderr, "%s\n", e.what());
            exit(1);
        }
        return true;
    }

    DataFile* file_for(size_t& index) {  // h5read.c:312-330 _find_data_file_for_image
        for (auto& d : files_)
            if (index >= d.first && index < d.first + d.frames) {
                index = index - d.first + d.offset;
                return &d;
            }
        return nullptr;
    }

    // Location of frame `local` inside data file d, refreshing the mapping once if it is missing.
    h5lite::ChunkInfo locate(DataFile& d, size_t local) {
        h5lite::ChunkInfo c;
        for (int attempt = 0; attempt < 2; ++attempt) {
            {
                std::shared_lock<std::shared_mutex> lk(mutex_);
                if (d.open) {
                    try {
                        if (d.layout.cls == h5lite::Layout::Contiguous) {
                            const uint64_t fb = (uint64_t)slow_ * fast_ * elem_size_;
                            c.addr = d.layout.addr + local * fb, c.size = fb;
                            c.found = d.layout.addr != h5lite::UNDEF && d.file->in_file(c.addr, fb);
                        } else {
                            c = d.dataset.find_chunk({local, 0, 0});
                            if (c.found && !d.file->in_file(c.addr, c.size)) c.found = false;  // index ahead of the data
                        }
                    } catch (const h5lite::Error&) {
                        c.found = false;  // index structures caught half-written
                    }
                    if (c.found) return c;
                }
            }
            if (attempt == 0) {
                std::unique_lock<std::shared_mutex> lk(mutex_);
                if (!d.open) {
                    if (!try_open(d)) return c;
                } else {
                    try {
                        d.file->refresh();
                        // the object header may have been rewritten (new index address, new extent)
                        d.dataset = d.file->object_at(d.dataset.address());
                        d.layout = d.dataset.layout();
                    } catch (const h5lite::Error&) {
                        return c;
                    }
                }
            }
        }
        return c;
    }

  public:
    explicit H5Read(const std::string& filename) {
        try {
            master_ = h5lite::File::open(filename);
            const std::string root = h5lite::dirname_of(filename);
            bool have_data = false;
            h5lite::Node data;
            try {
                data = h5lite::resolve(master_, "/entry/data/data");
                have_data = true;
            } catch (const h5lite::Error&) {
            }
            if (have_data && data.obj.is_dataset() && data.obj.layout().cls == h5lite::Layout::Virtual) {
                const auto dims = data.obj.shape();
                if (dims.size() != 3) throw std::runtime_error("VDS data not three dimensional");
                frames_ = dims[0], slow_ = dims[1], fast_ = dims[2];
                elem_size_ = data.obj.dtype().size;
                for (const auto& m : data.obj.vds_mappings()) {
                    if (m.dst_start.size() != 3 || (!m.src_all && m.src_start.size() != 3))
                        throw std::runtime_error("incorrect data dimensionality in the VDS mapping");
                    for (int k = 1; k < 3; ++k)
                        if ((!m.src_all && m.src_start[k] != 0) || m.dst_start[k] != 0)
                            throw std::runtime_error("incorrect chunk start in the VDS mapping");
                    DataFile d;
                    d.dsetname = m.dataset;
                    d.first = m.dst_start[0];
                    d.offset = m.src_all ? 0 : m.src_start[0];
                    d.frames = std::min<uint64_t>(m.dst_block[0], frames_ - std::min<uint64_t>(d.first, frames_));
                    if (m.file != ".") d.filename = (m.file[0] == '/') ? m.file : root + "/" + m.file;
                    files_.push_back(std::move(d));
                }
            } else if (have_data && data.obj.is_dataset()) {
                DataFile d;
                d.dsetname = "/entry/data/data";
                describe(d, data);
                d.frames = frames_ = d.dataset.shape()[0];
                files_.push_back(std::move(d));
            } else {
                // the Eiger file writer's layout: /entry/data/data_000001, data_000002, ... (links)
                const auto links = master_->get("/entry/data").links();
                for (const auto& kv : links) {
                    if (kv.first.rfind("data_", 0) != 0) continue;
                    DataFile d;
                    d.dsetname = "/entry/data/" + kv.first;
                    d.first = frames_;
                    {
                        std::unique_lock<std::shared_mutex> lk(mutex_);
                        if (!try_open(d))
                            throw std::runtime_error("Error: cannot open " + d.dsetname + " of " + filename);
                    }
                    d.frames = d.dataset.shape()[0];
                    frames_ += d.frames;
                    files_.push_back(std::move(d));
                }
                if (files_.empty()) throw std::runtime_error("Error: Reading H5 entry /entry/data/data");
            }
            {   // the first data file decides filtered / raw; later ones must agree when they appear
                std::unique_lock<std::shared_mutex> lk(mutex_);
                if (!files_.empty() && try_open(files_.front())) raw_pixels_ = !files_.front().filtered;
            }

            // trusted range (h5read.c:793-817)
            trusted_ = {0, elem_size_ == 4 ? (int64_t)UINT32_MAX : (int64_t)UINT16_MAX};
            if (const auto v = scalar("/entry/instrument/detector/saturation_value")) trusted_[1] = (int64_t)*v;
            else fprintf(stderr, "Warning: No saturation_value found, using maximum value\n");
            if (const auto v = scalar("/entry/instrument/detector/underload_value")) trusted_[0] = (int64_t)*v;
            else fprintf(stderr, "Warning: No underload_value found, using 0\n");
            if (const auto v = scalar("/entry/instrument/beam/incident_wavelength")) wavelength_ = (float)*v;
            else fprintf(stderr, "No wavelength data found...\n");
            const auto px = scalar("/entry/instrument/detector/x_pixel_size"), py = scalar("/entry/instrument/detector/y_pixel_size");
            if (px && py) pixel_size_ = {{(float)*py, (float)*px}};
            else fprintf(stderr, "Warning: No pixel size found\n");
            const auto bx = scalar("/entry/instrument/detector/beam_center_x"), by = scalar("/entry/instrument/detector/beam_center_y");
            if (bx && by) beam_center_ = {{(float)*by, (float)*bx}};
            else fprintf(stderr, "Warning: No beam center found\n");
            if (const auto v = scalar("/entry/instrument/detector/distance")) distance_ = (float)*v;
            else fprintf(stderr, "Warning: No detector distance found\n");
            if (pixel_size_) printf("Read pixel size: %f\n", (*pixel_size_)[1]);

            // oscillation (h5read.c:829-860): absent = stills
            try {
                const auto omega = h5lite::resolve(master_, "/entry/sample/sample_omega/omega").obj.read_as<double>();
                if (omega.size() < 2) throw std::runtime_error("Error: While reading oscillation, size<2");
                oscillation_ = {(float)omega[0], (float)omega[1] - (float)omega[0]};
            } catch (const h5lite::Error&) {
            }

            // mask (h5read.c:561-650): valid where the NXmx pixel_mask is 0
            const size_t npx = slow_ * fast_;
            try {
                const auto m = h5lite::resolve(master_, "/entry/instrument/detector/pixel_mask");
                const size_t ds = m.obj.dtype().size;
                if (ds == 4) printf("mask dtype uint32\n");
                else if (ds == 8) printf("mask dtype uint64\n");
                else throw std::runtime_error("Error: mask data size (" + std::to_string(ds) + ") != 4,8");
                const auto raw = m.obj.read_as<int64_t>();
                printf("Mask has %zu elements\n", raw.size());
                if (raw.size() != npx) throw std::runtime_error("Error: mask size does not match the image size");
                mask_.resize(npx);
                for (size_t i = 0; i < npx; ++i) mask_[i] = raw[i] == 0;
            } catch (const h5lite::Error&) {
                fprintf(stderr, "Warning: no mask data found at /entry/instrument/detector/pixel_mask\n");
                mask_.assign(npx, 1);
            }
            printf("Total data size: %zux%zux%zu\n", frames_, slow_, fast_);
        } catch (const h5lite::Error& e) {
            throw std::runtime_error(std::string("Error: Reading ") + filename + ": " + e.what());
        }
    }

    static bool is_ready_for_read(const std::string& path) {
        try {
            h5lite::File f(path);
            return f.exists("/entry/data");
        } catch (const std::exception&) {
            return false;
        }
    }

    bool chunks_are_raw_pixels() const override { return raw_pixels_; }

    bool is_image_available(size_t index) override {
        size_t local = index;
        DataFile* d = file_for(local);
        if (!d) {  // beyond every data file: a usage error (--images / --start-index), h5read.c:383-386 exits too
            fprintf(stderr, "Error: Could not find data file for frame %zu\n", index);
            exit(1);
        }
        return locate(*d, local).found;
    }

    size_t read_raw_chunk(size_t index, uint8_t* dst, size_t capacity) override {
        size_t local = index;
        DataFile* d = file_for(local);
        if (!d) return 0;
        const h5lite::ChunkInfo c = locate(*d, local);
        if (!c.found || c.size == 0) return 0;
        if (c.size > capacity) return SIZE_MAX;
        std::shared_lock<std::shared_mutex> lk(mutex_);
        return d->file->read_into(c.addr, c.size, dst) ? c.size : 0;
    }
    bool get_raw_chunk(size_t index, std::vector<uint8_t>& dst) override {
        size_t local = index;
        DataFile* d = file_for(local);
        if (!d) return false;
        const h5lite::ChunkInfo c = locate(*d, local);
        dst.clear();
        if (!c.found) return true;  // "zero length": the caller sleeps and retries
        std::shared_lock<std::shared_mutex> lk(mutex_);
        dst.resize(c.size);
        if (!d->file->read_into(c.addr, c.size, dst.data())) dst.clear();
        return true;
    }
    bool read_pixels(size_t index, void* dst, size_t bytes) override {
        if (!raw_pixels_) return false;
        size_t local = index;
        DataFile* d = file_for(local);
        if (!d) return false;
        const h5lite::ChunkInfo c = locate(*d, local);
        if (!c.found || c.size < bytes) return false;
        std::shared_lock<std::shared_mutex> lk(mutex_);
        return d->file->read_into(c.addr, bytes, dst);
    }

    size_t get_number_of_images() const override { return frames_; }
    std::array<size_t, 2> image_shape() const override { return {slow_, fast_}; }
    size_t get_element_size() const override { return elem_size_; }
    std::array<int64_t, 2> get_trusted_range() const override { return trusted_; }
    const std::vector<uint8_t>& get_mask() const override { return mask_; }
    std::optional<float> get_wavelength() const override { return wavelength_; }
    std::optional<std::array<float, 2>> get_pixel_size() const override { return pixel_size_; }
    std::optional<std::array<float, 2>> get_beam_center() const override { return beam_center_; }
    std::optional<float> get_detector_distance() const override { return distance_; }
    std::array<float, 2> get_oscillation() const override { return oscillation_; }
};
