// reader_sample.hpp -- `--sample`: the six generated Eiger2 XE 16M frames of the reference's
// h5read (h5read/src/h5read.c:203-276 _generate_sample_image, :1130-1189 h5read_generate_samples
// and _generate_e2xe_16m_mask; detector constants h5read/include/eiger2xe.h:6-17).
//
//   0: all zero            1: 1 on every module pixel      2: 100 every 42 pixels in x and y
//   3: I = x               4: I = y                        5: pcg32() % 10 on module pixels
//
// The reference's own executable cannot run on these (its frame loop asks for raw chunks and
// h5read answers "Cannot do direct chunk read with sample data", h5read.c:418-421); here the
// frames are handed out as pixels (Reader::read_pixels), so `spotfinder --sample` is a
// self-contained smoke run with known answers.
#pragma once
#include "reader.hpp"

class SampleRead : public Reader {
    static constexpr size_t MOD_FAST = 1028, MOD_SLOW = 512, GAP_FAST = 12, GAP_SLOW = 38;
    static constexpr size_t SLOW = 4362, FAST = 4148, NSLOW = 8, NFAST = 4, N_IMAGES = 6;
    size_t elem_size_;
    std::vector<uint8_t> mask_;

    struct Pcg32 {  // h5read.c:189-201
        uint64_t state = 0, inc = 0;
        uint32_t next() {
            const uint64_t old = state;
            state = old * 6364136223846793005ULL + (inc | 1);
            const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
            const uint32_t rot = (uint32_t)(old >> 59u);
            return (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31));
        }
    };

    template <typename T>
    void generate(size_t n, T* data) const {
        std::fill(data, data + SLOW * FAST, T(0));
        auto for_module_pixels = [&](auto&& fn) {
            for (size_t mody = 0; mody < NSLOW; ++mody) {
                const size_t row0 = mody * (MOD_SLOW + GAP_SLOW);
                for (size_t modx = 0; modx < NFAST; ++modx) {
                    const size_t col0 = modx * (MOD_FAST + GAP_FAST);
                    for (size_t row = 0; row < MOD_SLOW; ++row)
                        for (size_t x = 0; x < MOD_FAST; ++x) fn(data[FAST * (row0 + row) + col0 + x]);
                }
            }
        };
        switch (n) {
            case 0: break;
            case 1: for_module_pixels([](T& v) { v = 1; }); break;
            case 2:
                for (size_t y = 0; y < SLOW; y += 42)
                    for (size_t x = 0; x < FAST; x += 42) data[y * FAST + x] = 100;
                break;
            case 3:
                for (size_t y = 0; y < SLOW; ++y)
                    for (size_t x = 0; x < FAST; ++x) data[y * FAST + x] = (T)x;
                break;
            case 4:
                for (size_t y = 0; y < SLOW; ++y)
                    for (size_t x = 0; x < FAST; ++x) data[y * FAST + x] = (T)y;
                break;
            default: {
                Pcg32 rng;
                for_module_pixels([&](T& v) { v = (T)(rng.next() % 10); });
            }
        }
    }

  public:
    explicit SampleRead(size_t elem_size) : elem_size_(elem_size), mask_(SLOW * FAST, 1) {
        for (size_t gap = 1; gap < NSLOW; ++gap) {  // horizontal gaps
            const size_t y = gap * MOD_SLOW + (gap - 1) * GAP_SLOW;
            std::fill(mask_.begin() + y * FAST, mask_.begin() + (y + GAP_SLOW) * FAST, 0);
        }
        for (size_t gap = 1; gap < NFAST; ++gap) {  // vertical gaps
            const size_t x = gap * MOD_FAST + (gap - 1) * GAP_FAST;
            for (size_t y = 0; y < SLOW; ++y) std::fill(mask_.begin() + y * FAST + x, mask_.begin() + y * FAST + x + GAP_FAST, 0);
        }
    }
    bool is_image_available(size_t index) override { return index < N_IMAGES; }
    bool get_raw_chunk(size_t, std::vector<uint8_t>&) override { return false; }
    size_t read_raw_chunk(size_t, uint8_t*, size_t) override { return 0; }
    bool chunks_are_raw_pixels() const override { return true; }
    bool read_pixels(size_t index, void* dst, size_t bytes) override {
        if (index >= N_IMAGES || bytes != SLOW * FAST * elem_size_) return false;
        if (elem_size_ == 2) generate(index, static_cast<uint16_t*>(dst));
        else generate(index, static_cast<uint32_t*>(dst));
        return true;
    }
    size_t get_number_of_images() const override { return N_IMAGES; }
    std::array<size_t, 2> image_shape() const override { return {SLOW, FAST}; }
    size_t get_element_size() const override { return elem_size_; }
    // trusted_range_max = (image_t_type)(-1) (h5read.c:1165)
    std::array<int64_t, 2> get_trusted_range() const override { return {0, elem_size_ == 2 ? 65535 : 4294967295LL}; }
    const std::vector<uint8_t>& get_mask() const override { return mask_; }
    std::optional<float> get_wavelength() const override { return 0.0f; }  // calloc'ed, never set
    std::optional<std::array<float, 2>> get_pixel_size() const override { return {{0.75e-6f, 0.75e-6f}}; }
    std::optional<std::array<float, 2>> get_beam_center() const override { return {{SLOW / 2.0f, FAST / 2.0f}}; }
    std::optional<float> get_detector_distance() const override { return 0.5f; }
    std::array<float, 2> get_oscillation() const override { return {0, 0}; }
};
