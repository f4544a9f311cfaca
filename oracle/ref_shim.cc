/*
 * ref_shim.cc -- C ABI around the REFERENCE's own CPU spotfinder, compiled in
 * place from /root/reference/baseline/spotfinder/standalone.cc (never copied
 * into this repo).  Output: oracle/_ref/libffs_ref.so (git-ignored).
 *
 * TEST INFRASTRUCTURE ONLY: used to pin oracle/ffs_oracle.c, to generate
 * tests/golden/, and as bench.py's "reference" CPU baseline.
 *
 * Mirrors the reference's C API spotfinder_standard_dispersion
 * (baseline/spotfinder/baseline.cpp:849-881): convert image_t_type -> double,
 * run StandaloneSpotfinder<double>::standard_dispersion, count strong pixels.
 */
#include <standalone.h>

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <span>
#include <thread>
#include <vector>

namespace {
struct RefCtx {
    size_t width, height;
    StandaloneSpotfinder<double> finder;
    std::vector<double> converted;
    RefCtx(size_t w, size_t h) : width(w), height(h), finder(w, h), converted(w * h) {}
};

template <typename T>
int64_t run_one(RefCtx *ctx, const T *image, const uint8_t *mask, uint8_t *dst) {
    const size_t n = ctx->width * ctx->height;
    for (size_t i = 0; i < n; ++i) ctx->converted[i] = image[i];  // baseline.cpp:866-868
    auto result = ctx->finder.standard_dispersion(
      std::span<const double>(ctx->converted.data(), n), std::span<const uint8_t>(mask, n));
    int64_t count = 0;
    for (size_t i = 0; i < n; ++i) count += result[i];  // baseline.cpp:873-876
    if (dst) std::memcpy(dst, result.data(), n);
    return count;
}
}  // namespace

extern "C" {

void *ffs_ref_create(size_t width, size_t height) {
    return new RefCtx(width, height);
}
void ffs_ref_free(void *ctx) {
    delete static_cast<RefCtx *>(ctx);
}

/* Parameters are the reference's compile-time constants (standalone.cc:16-20):
 * kernel 3x3 half-size, min_count 2, threshold 0, nsig_b 6, nsig_s 3. */
int64_t ffs_ref_standard_dispersion(void *ctx, const void *image, int bytes_per_pixel,
                                    const uint8_t *mask, uint8_t *dst) {
    auto *c = static_cast<RefCtx *>(ctx);
    if (bytes_per_pixel == 2) return run_one(c, static_cast<const uint16_t *>(image), mask, dst);
    if (bytes_per_pixel == 4) return run_one(c, static_cast<const uint32_t *>(image), mask, dst);
    return -1;
}

/* Frame-parallel timing driver: n_threads workers, one reference context each
 * (as bm.cpp does per benchmark), pulling frames from a shared atomic counter like
 * the reference's frame loop (spotfinder.cc:686,752).  frames = n_distinct frames
 * stored back to back; frame f uses frames[f % n_distinct].  Returns wall seconds;
 * total_strong (optional) receives the summed strong-pixel count. */
double ffs_ref_run_frames(const void *frames, int bytes_per_pixel, size_t n_distinct,
                          const uint8_t *mask, size_t width, size_t height, size_t n_frames,
                          int n_threads, int64_t *total_strong) {
    const size_t frame_bytes = width * height * (size_t)bytes_per_pixel;
    std::atomic<size_t> next{0};
    std::atomic<int64_t> strong{0};
    std::vector<RefCtx *> ctxs;
    for (int t = 0; t < n_threads; ++t) ctxs.push_back(new RefCtx(width, height));
    auto t0 = std::chrono::steady_clock::now();
    {
        std::vector<std::jthread> workers;
        for (int t = 0; t < n_threads; ++t) {
            workers.emplace_back([&, t] {
                for (;;) {
                    size_t f = next.fetch_add(1);
                    if (f >= n_frames) break;
                    const char *img =
                      static_cast<const char *>(frames) + (f % n_distinct) * frame_bytes;
                    strong += ffs_ref_standard_dispersion(ctxs[t], img, bytes_per_pixel, mask,
                                                          nullptr);
                }
            });
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    for (auto *c : ctxs) delete c;
    if (total_strong) *total_strong = strong.load();
    return std::chrono::duration<double>(t1 - t0).count();
}
}
