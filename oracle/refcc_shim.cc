// refcc_shim.cc -- C entry points around the REFERENCE'S OWN host-side connected components,
// reflection filters and Kabsch-space variances, compiled in place from /root/reference
// (oracle/Makefile target `refcc`; stand-in Boost.Graph / Eigen headers: oracle/stub_cc).
// Test infrastructure: it pins the oracle's restatement (ffs_oracle.c) of SURVEY.md 8(a) rows
// a12-a15 and 8(f) rank 3.  Nothing of the product links or loads this.
//
//   ConnectedComponents(strong u8, image, W, H, min_spot_size)        connected_components.cc:16-39
//   ::get_num_strong_pixels / _filtered / get_boxes                   connected_components.hpp:271-292
//   ::find_2d_components(min_spot_size, max_sep)                      connected_components.cc:238-266
//   ::find_3d_components(slices, W, H, min_spot_size, max_sep)        connected_components.cc:270-470
//   Reflection3D::center_of_mass / variances_in_kabsch_space          connected_components.hpp:70-104, .cc:159-203
//   the per-spot variance loop of the executable                      spotfinder/spotfinder.cc:1152-1196
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "connected_components.hpp"

extern "C" {

struct refcc_spot {
    uint32_t x_min, x_max, y_min, y_max;
    int32_t z_min, z_max, num_pixels, depth;
    float com_x, com_y, com_z, pad;
    double sigma_b_variance, sigma_m_variance;
};

struct refcc_box {
    uint32_t l, t, r, b;
    int32_t num_pixels;
};

struct refcc_handle {
    std::vector<std::unique_ptr<ConnectedComponents>> slices;
    std::vector<Reflection3D> last;  // result of the last find_2d / find_3d call
    uint32_t width = 0, height = 0;
};

void* ffs_refcc_new(void) { return new refcc_handle; }
void ffs_refcc_free(void* h) { delete static_cast<refcc_handle*>(h); }
int ffs_refcc_pixel_bytes(void) { return (int)sizeof(pixel_t); }

// One frame -> one ConnectedComponents (what the frame loop builds per image, spotfinder.cc:897-903).
int ffs_refcc_add_slice(void* hv, const uint8_t* strong, const void* image, uint32_t width, uint32_t height,
                        uint32_t min_spot_size, uint32_t* num_strong, uint32_t* num_strong_filtered, uint32_t* n_boxes) {
    auto* h = static_cast<refcc_handle*>(hv);
    try {
        h->slices.push_back(std::make_unique<ConnectedComponents>(strong, static_cast<const pixel_t*>(image), width,
                                                                  height, min_spot_size));
    } catch (const std::exception&) {
        return -1;
    }
    h->width = width, h->height = height;
    const auto& cc = *h->slices.back();
    if (num_strong) *num_strong = cc.get_num_strong_pixels();
    if (num_strong_filtered) *num_strong_filtered = cc.get_num_strong_pixels_filtered();
    if (n_boxes) *n_boxes = (uint32_t)cc.get_boxes().size();
    return (int)h->slices.size() - 1;
}

int ffs_refcc_boxes(void* hv, int slice, refcc_box* out, size_t cap) {
    auto* h = static_cast<refcc_handle*>(hv);
    if (slice < 0 || (size_t)slice >= h->slices.size()) return -1;
    const auto& boxes = h->slices[slice]->get_boxes();
    for (size_t i = 0; i < boxes.size() && i < cap; ++i) out[i] = {boxes[i].l, boxes[i].t, boxes[i].r, boxes[i].b, boxes[i].num_pixels};
    return (int)boxes.size();
}

static void export_spots(const std::vector<Reflection3D>& refl, refcc_spot* out, size_t cap) {
    for (size_t i = 0; i < refl.size() && i < cap; ++i) {
        const auto& r = refl[i];
        refcc_spot s{};
        s.x_min = r.get_x_min(), s.x_max = r.get_x_max(), s.y_min = r.get_y_min(), s.y_max = r.get_y_max();
        s.z_min = r.get_z_min(), s.z_max = r.get_z_max(), s.num_pixels = r.get_num_pixels();
        auto [x, y, z] = r.center_of_mass();
        s.com_x = x, s.com_y = y, s.com_z = z;
        out[i] = s;
    }
}

// Returns the number of reflections after both filters; *n_components = before.
long ffs_refcc_find_2d(void* hv, int slice, uint32_t min_spot_size, float max_sep, refcc_spot* out, size_t cap,
                       uint32_t* n_components) {
    auto* h = static_cast<refcc_handle*>(hv);
    if (slice < 0 || (size_t)slice >= h->slices.size()) return -1;
    try {
        if (n_components) *n_components = (uint32_t)h->slices[slice]->get_boxes().size();
        h->last = h->slices[slice]->find_2d_components(min_spot_size, max_sep);
        export_spots(h->last, out, cap);
    } catch (const std::exception&) {
        return -1;
    }
    return (long)h->last.size();
}

long ffs_refcc_find_3d(void* hv, uint32_t min_spot_size, float max_sep, refcc_spot* out, size_t cap) {
    auto* h = static_cast<refcc_handle*>(hv);
    if (h->slices.empty()) return -1;
    try {
        h->last = ConnectedComponents::find_3d_components(h->slices, h->width, h->height, min_spot_size, max_sep);
        export_spots(h->last, out, cap);
    } catch (const std::exception&) {
        return -1;
    }
    return (long)h->last.size();
}

// The executable's variance loop (spotfinder.cc:1152-1196) over the reflections of the last
// find_3d call; geometry in the units that loop uses (distance / pixel size already in mm).
long ffs_refcc_variances(void* hv, double distance_mm, double beam_center_x_px, double beam_center_y_px,
                         double pixel_size_x_mm, double pixel_size_y_mm, double wavelength, double oscillation_start,
                         double oscillation_width, int num_images, refcc_spot* out, size_t cap) {
    auto* h = static_cast<refcc_handle*>(hv);
    try {
        std::array<int, 2> image_size = {static_cast<int>(h->width), static_cast<int>(h->height)};
        Panel panel(distance_mm, {beam_center_x_px, beam_center_y_px}, {pixel_size_x_mm, pixel_size_y_mm}, image_size);
        Vector3d s0 = {0.0, 0.0, -1.0 / wavelength};
        Scan scan({1, num_images}, {oscillation_start, oscillation_width});
        int image_range_0 = scan.get_image_range()[0];
        Vector3d m2 = {1.0, 0.0, 0.0};
        constexpr double deg_to_rad = M_PI / 180.0;
        for (size_t i = 0; i < h->last.size() && i < cap; ++i) {
            const auto& refl = h->last[i];
            auto [x, y, z] = refl.center_of_mass();
            auto [xmm, ymm] = panel.px_to_mm(x, y);
            Vector3d s1 = panel.get_lab_coord(xmm, ymm);
            double phi = (oscillation_start + (z - image_range_0) * oscillation_width) * deg_to_rad;
            auto [sigma_b_variance, sigma_m_variance, bbox_depth] = refl.variances_in_kabsch_space(s1, s0, m2, panel, scan, phi);
            out[i].sigma_b_variance = sigma_b_variance;
            out[i].sigma_m_variance = sigma_m_variance;
            out[i].depth = bbox_depth;
        }
    } catch (const std::exception&) {
        return -1;
    }
    return (long)h->last.size();
}

}  // extern "C"
