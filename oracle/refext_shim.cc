// refext_shim.cc -- C entry points around the REFERENCE'S OWN CPU threshold classes of
// baseline/spotfinder/baseline.cpp (DIALS's DispersionThreshold and DispersionExtendedThreshold
// as the reference carries them), compiled in place from /root/reference (oracle/Makefile target
// `refext`; stand-in scitbx / DIALS headers: oracle/stub_dials).  Test infrastructure: it pins
// the oracle's restatement (ffs_oracle.c) of SURVEY.md 8(a) rows a3 and a4.  Nothing of the
// product links or loads this.
//
//   DispersionThreshold::threshold / threshold_w_gain                  baseline.cpp:259-310
//   DispersionExtendedThreshold::threshold / threshold_w_gain          baseline.cpp:730-805
//   its stages compute_sat / compute_dispersion_threshold / erode_dispersion_mask /
//   compute_final_threshold (public members)                           baseline.cpp:375-714
//
// Pixels are converted to double first, as the reference's own context does
// (_spotfind_context<image_t_type, double>, baseline.cpp:817-878).
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include <scitbx/array_family/accessors/c_grid.h>
#include <scitbx/array_family/tiny_types.h>
namespace af = scitbx::af;
// what spotfind_test_utils.h defines under HAVE_DIALS (left undefined here: it would pull in local.h)
const af::tiny<int, 2> kernel_size_(3, 3);

#include "baseline.cpp"

namespace {
struct Views {
    std::vector<double> src;
    std::unique_ptr<bool[]> mask, dst;
    af::c_grid<2> grid;
    Views(const void* image, int bpp, const uint8_t* mask_u8, int w, int h) : src((size_t)w * h), grid(h, w) {
        const size_t n = (size_t)w * h;
        mask.reset(new bool[n]);
        dst.reset(new bool[n]);
        for (size_t k = 0; k < n; ++k) {
            src[k] = bpp == 2 ? (double)static_cast<const uint16_t*>(image)[k] : (double)static_cast<const uint32_t*>(image)[k];
            mask[k] = mask_u8[k] != 0;
            dst[k] = false;
        }
    }
    af::const_ref<double, af::c_grid<2>> s() const { return {src.data(), grid}; }
    af::const_ref<bool, af::c_grid<2>> m() const { return {mask.get(), grid}; }
    af::ref<bool, af::c_grid<2>> d() const { return {dst.get(), grid}; }
};
}  // namespace

extern "C" {

// algorithm: 0 dispersion, 1 dispersion_extended.  gain: NULL or w*h doubles.
// stage (extended only): 1 = after compute_dispersion_threshold, 2 = after erode_dispersion_mask,
// 3 (or 0) = the final result.  Returns the number of true pixels in out, or -1 when the
// reference's own argument checks (DIALS_ASSERT) throw.
long ffs_refext_threshold(int algorithm, const void* image, int bytes_per_pixel, const uint8_t* mask, int width, int height,
                          int kernel_half_x, int kernel_half_y, double nsig_b, double nsig_s, double threshold, int min_count,
                          const double* gain, int stage, uint8_t* out) {
    try {
        Views v(image, bytes_per_pixel, mask, width, height);
        const af::tiny<int, 2> size(height, width), kernel(kernel_half_y, kernel_half_x);  // kernel_size_[0] is the slow (y) half-size
        af::const_ref<double, af::c_grid<2>> g(gain, v.grid);
        if (algorithm == 0) {
            baseline::DispersionThreshold algo(size, kernel, nsig_b, nsig_s, threshold, min_count);
            if (gain) algo.threshold_w_gain(v.s(), v.m(), g, v.d());
            else algo.threshold(v.s(), v.m(), v.d());
        } else {
            baseline::DispersionExtendedThreshold algo(size, kernel, nsig_b, nsig_s, threshold, min_count);
            if (stage == 0 || stage == 3) {
                if (gain) algo.threshold_w_gain(v.s(), v.m(), g, v.d());
                else algo.threshold(v.s(), v.m(), v.d());
            } else {  // the body of threshold() (baseline.cpp:742-760 / 785-804), stopped early
                typedef baseline::DispersionExtendedThreshold::Data<double> Cell;
                std::vector<Cell> buffer((size_t)width * height);
                af::ref<Cell> table(buffer.data(), af::trivial_accessor(buffer.size()));
                algo.compute_sat(table, v.s(), v.m());
                if (gain) algo.compute_dispersion_threshold(table, v.s(), v.m(), g, v.d());
                else algo.compute_dispersion_threshold(table, v.s(), v.m(), v.d());
                if (stage >= 2) algo.erode_dispersion_mask(v.m(), v.d());
            }
        }
        long count = 0;
        for (size_t k = 0; k < (size_t)width * height; ++k) {
            out[k] = v.dst[k] ? 1 : 0;
            count += out[k];
        }
        return count;
    } catch (const std::exception&) {
        return -1;
    }
}

// Persistent context for timing runs (bench.py --impl reference): buffers and the algorithm object
// are built once, as the reference's own _spotfind_context does (baseline.cpp:817-842); a call
// converts the pixels to double and runs threshold(), like spotfinder_standard_dispersion (:858-878).
struct refext_ctx {
    int algorithm, width, height;
    std::vector<double> src;
    std::unique_ptr<bool[]> mask, dst;
    std::unique_ptr<baseline::DispersionThreshold> disp;
    std::unique_ptr<baseline::DispersionExtendedThreshold> ext;
};

void* ffs_refext_create(int algorithm, int width, int height, const uint8_t* mask_u8, int kernel_half_x, int kernel_half_y,
                        double nsig_b, double nsig_s, double threshold, int min_count) {
    try {
        auto c = std::make_unique<refext_ctx>();
        const size_t n = (size_t)width * height;
        c->algorithm = algorithm, c->width = width, c->height = height;
        c->src.resize(n);
        c->mask.reset(new bool[n]);
        c->dst.reset(new bool[n]);
        for (size_t k = 0; k < n; ++k) c->mask[k] = mask_u8[k] != 0;
        const af::tiny<int, 2> size(height, width), kernel(kernel_half_y, kernel_half_x);
        if (algorithm == 0) c->disp = std::make_unique<baseline::DispersionThreshold>(size, kernel, nsig_b, nsig_s, threshold, min_count);
        else c->ext = std::make_unique<baseline::DispersionExtendedThreshold>(size, kernel, nsig_b, nsig_s, threshold, min_count);
        return c.release();
    } catch (const std::exception&) {
        return nullptr;
    }
}

void ffs_refext_destroy(void* ctx) { delete static_cast<refext_ctx*>(ctx); }

long ffs_refext_run(void* ctx, const void* image, int bytes_per_pixel, uint8_t* out) {
    auto* c = static_cast<refext_ctx*>(ctx);
    try {
        const size_t n = (size_t)c->width * c->height;
        if (bytes_per_pixel == 2)
            for (size_t k = 0; k < n; ++k) c->src[k] = static_cast<const uint16_t*>(image)[k];
        else
            for (size_t k = 0; k < n; ++k) c->src[k] = static_cast<const uint32_t*>(image)[k];
        const af::c_grid<2> grid(c->height, c->width);
        af::const_ref<double, af::c_grid<2>> s(c->src.data(), grid);
        af::const_ref<bool, af::c_grid<2>> m(c->mask.get(), grid);
        af::ref<bool, af::c_grid<2>> d(c->dst.get(), grid);
        if (c->disp) c->disp->threshold(s, m, d);
        else c->ext->threshold(s, m, d);
        long count = 0;
        for (size_t k = 0; k < n; ++k) {
            out[k] = c->dst[k] ? 1 : 0;
            count += out[k];
        }
        return count;
    } catch (const std::exception&) {
        return -1;
    }
}

}  // extern "C"
