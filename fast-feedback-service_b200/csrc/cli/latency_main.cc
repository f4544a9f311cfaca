// latency_main.cc -- `ffs_latency`: per-frame submit -> result-on-host latency of the streaming
// C ABI at a paced input rate (SURVEY.md 8(d) timing protocol (c): "per-frame latency distribution
// p50/p99 for cfg 5", Jungfrau 4M at 2500 Hz).  The reference's counterpart is the per-image
// copy / kernel / post timing block of `--threads 1` (spotfinder/spotfinder.cc:1055-1076).
//
// A producer thread submits frame i at t0 + i / rate from the pinned ring (the slots were filled
// during warm-up, as a reader thread would have filled them before the frame is "available");
// a collector thread takes the results in order.  Latency of frame i = time its result is on the
// host minus its SCHEDULED time, so a producer that falls behind (ring full) shows up as latency,
// not as a lower rate.  One JSON object on stdout.
//
//   ffs_latency --frames F.raw --mask M.raw --width W --height H --bpp 2|4 --n-distinct K
//               [--count 5000] [--rate 2500] [--ring 8] [--algorithm 0|1] [--device 0] [--warmup 200]
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/ffs_b200.h"

using clk = std::chrono::steady_clock;

static std::vector<uint8_t> read_file(const std::string& path, size_t want) {
    std::vector<uint8_t> v(want);
    FILE* f = fopen(path.c_str(), "rb");
    if (!f || fread(v.data(), 1, want, f) != want) {
        fprintf(stderr, "ffs_latency: cannot read %zu bytes from %s\n", want, path.c_str());
        exit(2);
    }
    fclose(f);
    return v;
}

int main(int argc, char** argv) {
    std::string frames_path, mask_path;
    uint32_t width = 0, height = 0, bpp = 2, n_distinct = 1, count = 5000, ring = 8, warmup = 200;
    int algorithm = 0, device = 0;
    double rate = 2500.0;
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string a = argv[i];
        const char* v = argv[i + 1];
        if (a == "--frames") frames_path = v;
        else if (a == "--mask") mask_path = v;
        else if (a == "--width") width = (uint32_t)atoi(v);
        else if (a == "--height") height = (uint32_t)atoi(v);
        else if (a == "--bpp") bpp = (uint32_t)atoi(v);
        else if (a == "--n-distinct") n_distinct = (uint32_t)atoi(v);
        else if (a == "--count") count = (uint32_t)atoi(v);
        else if (a == "--rate") rate = atof(v);
        else if (a == "--ring") ring = (uint32_t)atoi(v);
        else if (a == "--algorithm") algorithm = atoi(v);
        else if (a == "--device") device = atoi(v);
        else if (a == "--warmup") warmup = (uint32_t)atoi(v);
        else {
            fprintf(stderr, "ffs_latency: unknown option %s\n", a.c_str());
            return 2;
        }
    }
    if (!width || !height || frames_path.empty() || mask_path.empty() || rate <= 0 || !count) {
        fprintf(stderr, "usage: ffs_latency --frames F --mask M --width W --height H --bpp B --n-distinct K ...\n");
        return 2;
    }
    const size_t px = (size_t)width * height, frame_bytes = px * bpp;
    const std::vector<uint8_t> frames = read_file(frames_path, frame_bytes * n_distinct);
    const std::vector<uint8_t> mask = read_file(mask_path, px);

    ffs_params p;
    ffs_default_params(&p);
    p.width = width;
    p.height = height;
    p.bytes_per_pixel = bpp;
    p.algorithm = algorithm;
    p.device = device;
    p.n_slots = (int32_t)ring;
    ffs_ctx* ctx = nullptr;
    if (int rc = ffs_create(&p, mask.data(), &ctx)) {
        fprintf(stderr, "ffs_latency: ffs_create failed (%d): %s\n", rc, ffs_last_error(nullptr));
        return 1;
    }
    // warm-up: fills every pinned ring slot with a frame and lets every slot capture its graph
    uint64_t strong_total = 0;
    {
        std::vector<int> seen(ring, 0);
        for (uint32_t i = 0; i < std::max(warmup, 3 * ring); ++i) {
            int slot = -1;
            void* pin = nullptr;
            if (ffs_acquire(ctx, &slot, &pin) != FFS_OK) return 1;
            if (!seen[slot]) {
                memcpy(pin, frames.data() + (size_t)(slot % n_distinct) * frame_bytes, frame_bytes);
                seen[slot] = 1;
            }
            if (ffs_submit(ctx, slot, i) != FFS_OK) return 1;
            if (ffs_pending(ctx) >= (int)ring - 1) {
                ffs_frame_result r;
                if (ffs_wait(ctx, &r, nullptr, 0, nullptr) != FFS_OK) return 1;
            }
        }
        ffs_frame_result r;
        while (ffs_pending(ctx) > 0)
            if (ffs_wait(ctx, &r, nullptr, 0, nullptr) != FFS_OK) return 1;
    }

    std::vector<double> sched(count), done(count, -1.0);
    std::atomic<uint32_t> submitted{0};
    std::atomic<bool> failed{false};
    uint32_t late = 0;
    const auto t0 = clk::now() + std::chrono::milliseconds(2);
    auto now_s = [&] { return std::chrono::duration<double>(clk::now() - t0).count(); };

    std::thread collector([&] {
        uint32_t got = 0;
        ffs_frame_result r;
        while (got < count && !failed) {
            if (got >= submitted.load(std::memory_order_acquire)) {
                std::this_thread::yield();
                continue;
            }
            if (ffs_wait(ctx, &r, nullptr, 0, nullptr) != FFS_OK) {
                failed = true;
                break;
            }
            done[r.frame_id] = now_s();
            strong_total += r.num_strong_pixels;
            if (r.status != 0) failed = true;
            ++got;
        }
    });
    for (uint32_t i = 0; i < count && !failed; ++i) {
        sched[i] = i / rate;
        while (now_s() < sched[i]) {  // coarse sleep, then spin
            const double left = sched[i] - now_s();
            if (left > 200e-6) std::this_thread::sleep_for(std::chrono::microseconds((long)((left - 150e-6) * 1e6)));
        }
        int slot = -1;
        void* pin = nullptr;
        if (ffs_acquire(ctx, &slot, &pin) != FFS_OK || ffs_submit(ctx, slot, i) != FFS_OK) {
            failed = true;
            break;
        }
        if (now_s() - sched[i] > 0.5 / rate) ++late;
        submitted.store(i + 1, std::memory_order_release);
    }
    collector.join();
    const double total = now_s();
    if (failed) {
        fprintf(stderr, "ffs_latency: stream failed: %s\n", ffs_last_error(ctx));
        return 1;
    }
    std::vector<double> lat(count);
    for (uint32_t i = 0; i < count; ++i) lat[i] = (done[i] - sched[i]) * 1e6;
    std::sort(lat.begin(), lat.end());
    auto q = [&](double f) { return lat[std::min<size_t>(count - 1, (size_t)(f * count))]; };
    double mean = 0;
    for (double v : lat) mean += v;
    printf("{\"frames\": %u, \"rate_hz\": %.1f, \"achieved_fps\": %.1f, \"ring\": %u, \"latency_us\": {\"p50\": %.1f, "
           "\"p90\": %.1f, \"p99\": %.1f, \"max\": %.1f, \"mean\": %.1f}, \"late_submits\": %u, "
           "\"mean_strong_pixels\": %.1f, \"h2d_bytes_per_frame\": %zu}\n",
           count, rate, count / total, ring, q(0.50), q(0.90), q(0.99), lat.back(), mean / count, late,
           (double)strong_total / (count + 0.0), frame_bytes);
    ffs_destroy(ctx);
    return 0;
}
