// reader.hpp -- frame ingestion: the abstract Reader the frame loop sees
// (h5read/include/h5read.h:173-204) and the /dev/shm Eiger-stream reader
// (spotfinder/shmread.cc:13-93): JSON header `start_1`, int32 pixel mask `start_5`
// (valid = !value), one bitshuffle-LZ4 chunk per frame in `image_%06d_2`; the directory is
// ready once `start_1` and `start_4` exist.  NeXus / HDF5 input is reader_h5.hpp (on the
// in-tree h5lite.hpp, no libhdf5); CBF input is out of scope.
#pragma once
#include <array>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <cstdio>
#include <filesystem>
#include <fstream>
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "json_mini.hpp"

class Reader {
  public:
    virtual ~Reader() = default;
    virtual bool is_image_available(size_t index) = 0;
    // Raw compressed chunk (12-byte HDF5 bitshuffle header + body).  Empty = try again.
    virtual bool get_raw_chunk(size_t index, std::vector<uint8_t>& destination) = 0;
    // Reads the chunk straight into caller memory (the pinned upload buffer of a ring slot).
    // Returns the chunk size, 0 when the file is not (completely) there yet, or SIZE_MAX when
    // it does not fit `capacity` (nothing useful is in dst then).
    virtual size_t read_raw_chunk(size_t index, uint8_t* dst, size_t capacity) = 0;
    // Sources whose frames are stored without a compression filter hand out pixels instead of
    // chunks: read_pixels copies frame `index` (bytes = slow * fast * element size) into dst.
    virtual bool chunks_are_raw_pixels() const { return false; }
    virtual bool read_pixels(size_t /*index*/, void* /*dst*/, size_t /*bytes*/) { return false; }
    virtual size_t get_number_of_images() const = 0;
    virtual std::array<size_t, 2> image_shape() const = 0;  // {slow, fast}
    virtual size_t get_element_size() const = 0;
    virtual std::array<int64_t, 2> get_trusted_range() const = 0;
    virtual const std::vector<uint8_t>& get_mask() const = 0;  // 1 = valid
    virtual std::optional<float> get_wavelength() const = 0;
    virtual std::optional<std::array<float, 2>> get_pixel_size() const = 0;   // {y, x} metres
    virtual std::optional<std::array<float, 2>> get_beam_center() const = 0;  // {y, x} pixels
    virtual std::optional<float> get_detector_distance() const = 0;           // metres
    virtual std::array<float, 2> get_oscillation() const = 0;                 // {start, width}
};

class SHMRead : public Reader {
    std::string base_;
    size_t num_images_ = 0;
    std::array<size_t, 2> shape_{};
    size_t elem_size_ = 2;
    std::array<int64_t, 2> trusted_{};
    std::vector<uint8_t> mask_;
    std::optional<float> wavelength_;
    float distance_ = 0;
    std::array<float, 2> pixel_size_{}, beam_center_{}, oscillation_{0, 0};

    std::string image_path(size_t index) const {
        char name[64];
        snprintf(name, sizeof name, "/image_%06zu_2", index);
        return base_ + name;
    }

  public:
    explicit SHMRead(const std::string& path) : base_(path) {
        std::ifstream f(path + "/start_1");
        std::stringstream ss;
        ss << f.rdbuf();
        const jsonmini::Value data = jsonmini::Parser::parse(ss.str());
        num_images_ = (size_t)data.at("nimages").number() * (size_t)data.at("ntrigger").number();
        shape_ = {(size_t)data.at("y_pixels_in_detector").number(), (size_t)data.at("x_pixels_in_detector").number()};
        const int bits = (int)data.at("bit_depth_image").number();
        if (bits == 16) elem_size_ = 2;
        else if (bits == 32) elem_size_ = 4;
        else throw std::runtime_error("Data is unhandled bit-depth: " + std::to_string(bits) + "-bit");
        trusted_ = {0, (int64_t)data.at("countrate_correction_count_cutoff").number()};
        if (data.contains("wavelength")) wavelength_ = (float)data.at("wavelength").number();
        distance_ = (float)data.at("detector_distance").number() / 1000;
        pixel_size_ = {(float)data.at("y_pixel_size").number(), (float)data.at("x_pixel_size").number()};
        beam_center_ = {(float)data.at("beam_center_y").number(), (float)data.at("beam_center_x").number()};
        if (data.contains("omega_start") && data.contains("omega_increment"))
            oscillation_ = {(float)data.at("omega_start").number(), (float)data.at("omega_increment").number()};
        const size_t px = shape_[0] * shape_[1];
        const std::string mask_file = path + "/start_5";
        if (std::filesystem::file_size(mask_file) != px * sizeof(int32_t))
            throw std::runtime_error("Error: Mask file does not match expected size");
        std::vector<int32_t> raw(px);
        std::ifstream fm(mask_file, std::ios::in | std::ios::binary);
        fm.read(reinterpret_cast<char*>(raw.data()), px * sizeof(int32_t));
        mask_.resize(px);
        for (size_t i = 0; i < px; ++i) mask_[i] = !raw[i];
    }

    static bool is_ready_for_read(const std::string& path) {
        return std::filesystem::exists(path + "/start_1") && std::filesystem::exists(path + "/start_4");
    }

    bool is_image_available(size_t index) override { return std::filesystem::exists(image_path(index)); }
    bool get_raw_chunk(size_t index, std::vector<uint8_t>& dst) override {
        std::ifstream f(image_path(index), std::ios::in | std::ios::binary | std::ios::ate);
        if (!f) return false;
        const std::streamsize n = f.tellg();
        if (n <= 0) {  // non-atomic /dev/shm write caught half way (spotfinder.cc:811-819)
            dst.clear();
            return true;
        }
        dst.resize((size_t)n);
        f.seekg(0);
        f.read(reinterpret_cast<char*>(dst.data()), n);
        dst.resize((size_t)f.gcount());
        return true;
    }
    size_t read_raw_chunk(size_t index, uint8_t* dst, size_t capacity) override {
        const int fd = ::open(image_path(index).c_str(), O_RDONLY);
        if (fd < 0) return 0;
        struct stat st {};
        if (fstat(fd, &st) != 0 || st.st_size <= 0) {
            ::close(fd);
            return 0;
        }
        if ((size_t)st.st_size > capacity) {
            ::close(fd);
            return SIZE_MAX;
        }
        size_t got = 0;
        while (got < (size_t)st.st_size) {
            const ssize_t r = ::read(fd, dst + got, (size_t)st.st_size - got);
            if (r <= 0) break;
            got += (size_t)r;
        }
        ::close(fd);
        return got;
    }
    size_t get_number_of_images() const override { return num_images_; }
    std::array<size_t, 2> image_shape() const override { return shape_; }
    size_t get_element_size() const override { return elem_size_; }
    std::array<int64_t, 2> get_trusted_range() const override { return trusted_; }
    const std::vector<uint8_t>& get_mask() const override { return mask_; }
    std::optional<float> get_wavelength() const override { return wavelength_; }
    std::optional<std::array<float, 2>> get_pixel_size() const override { return pixel_size_; }
    std::optional<std::array<float, 2>> get_beam_center() const override { return beam_center_; }
    std::optional<float> get_detector_distance() const override { return distance_; }
    std::array<float, 2> get_oscillation() const override { return oscillation_; }
};
