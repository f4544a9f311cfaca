// spotfinder_main.cc -- the `spotfinder` / `spotfinder32` executable (PIXEL_DATA_32BIT), a
// drop-in for the reference's (spotfinder/spotfinder.cc:407-1331) as spawned by the Zocalo
// GPUPerImageAnalysis service (src/ffs/service.py:419-507): same argv, exit codes, stdout
// lines and JSON-lines pipe protocol.  All image processing happens in libffs_b200.so.
#include <atomic>
#include <charconv>
#include <chrono>
#include <csignal>
#include <cstring>
#include <mutex>
#include <thread>
#include <unistd.h>

#include "../../../include/ffs_b200.h"
#include "args.hpp"
#include "bslz4.hpp"
#include "png_writer.hpp"
#include "reader.hpp"
#include "reader_h5.hpp"
#include "reader_sample.hpp"
#include "h5_results.hpp"

#ifndef FFS_VERSION
#define FFS_VERSION "b200-0.1.0"
#endif
#ifdef PIXEL_DATA_32BIT
using pixel_t = uint32_t;
#else
using pixel_t = uint16_t;
#endif

using namespace std::chrono_literals;
using clk = std::chrono::high_resolution_clock;

static std::atomic<bool> g_stop{false};
static std::atomic<int> g_sigints{0};
static void stop_processing(int) {  // spotfinder.cc:43-54: once = graceful, twice = quick_exit
    if (g_sigints.fetch_add(1) >= 1) std::quick_exit(1);
    g_stop = true;
}

static std::string shortest(float v) {
    char b[32];
    auto r = std::to_chars(b, b + sizeof b, v);
    return std::string(b, r.ptr);
}

static double seconds_since(clk::time_point t) { return std::chrono::duration<double>(clk::now() - t).count(); }

// Wait until checker(path) holds, like wait_for_ready_for_read (spotfinder.cc:139-175).
template <class F>
static void wait_for_ready(const std::string& path, F checker, float timeout) {
    if (checker(path)) return;
    const auto start = clk::now();
    while (!checker(path)) {
        printf("\rWaiting for %s to be ready for read [%4.1f s] ", path.c_str(), seconds_since(start));
        fflush(stdout);
        if (seconds_since(start) > timeout) {
            printf("\nError: Waited too long for read availability\n");
            exit(1);
        }
        std::this_thread::sleep_for(80ms);
    }
    printf("\n");
}

struct DetectorGeometry {  // kernels/masking.cuh:14-70
    float pixel_size_x = 0, pixel_size_y = 0;  // metres
    float beam_center_x = 0, beam_center_y = 0;  // pixels
    float distance = 0;  // metres
};

class PipeHandler {  // spotfinder.cc:208-255
    int fd_;
    std::mutex m_;

  public:
    explicit PipeHandler(int fd) : fd_(fd) { printf("PipeHandler initialized with pipe_fd: %d\n", fd); }
    ~PipeHandler() { close(fd_); }
    void send(const std::string& line) {
        std::lock_guard<std::mutex> lk(m_);
        const std::string s = line + "\n";
        if (write(fd_, s.c_str(), s.size()) == -1) fprintf(stderr, "Error writing to pipe: %s\n", strerror(errno));
    }
};

int main(int argc, char** argv) {
    setvbuf(stdout, nullptr, _IOLBF, 0);
    printf("Spotfinder version: %s\n", FFS_VERSION);
    SpotfinderArgs args;
    if (!parse_args(argc, argv, FFS_VERSION, args)) return 1;
    if (args.list_devices) {  // cuda_arg_parser.cc:38-53
        for (int i = 0; i < ffs_device_count(); ++i) {
            char d[256];
            if (ffs_device_description(i, d, sizeof d) == FFS_OK) printf("%d: %s\n", i, d);
        }
        return 0;
    }
    {
        char d[256];
        if (ffs_device_description(args.device, d, sizeof d) != FFS_OK) {
            printf("Error: Could not select CUDA device %d\n", args.device);
            return 1;
        }
        printf("Using %s\n", d);
    }
    std::string algo = args.algorithm;
    for (auto& ch : algo) ch = (char)std::tolower((unsigned char)ch);
    if (algo != "dispersion" && algo != "dispersion_extended") {  // spotfinder.cc:189-202 throws
        printf("Error: Invalid algorithm specified: %s\n", args.algorithm.c_str());
        return 1;
    }
    printf("Algorithm: %s\n", algo == "dispersion" ? "Dispersion" : "Dispersion Extended");
    if (args.threads < 1) {
        printf("Error: Thread count must be >= 1\n");
        return 1;
    }

    // ---- reader ---------------------------------------------------------------------------
    const std::string& file = args.file;
    const bool sample = file.empty();  // --sample, or no file with H5READ_IMPLICIT_SAMPLE set (args.hpp)
    if (!sample && !std::filesystem::exists(file))
        wait_for_ready(file, [](const std::string& s) { return std::filesystem::exists(s); }, args.timeout);
    std::unique_ptr<Reader> reader_ptr;
    if (sample) {
        reader_ptr = std::make_unique<SampleRead>(sizeof(pixel_t));
    } else if (std::filesystem::is_directory(file)) {
        wait_for_ready(file, SHMRead::is_ready_for_read, args.timeout);
        try {
            reader_ptr = std::make_unique<SHMRead>(file);
        } catch (const std::exception& e) {
            printf("Error: %s\n", e.what());
            return 1;
        }
    } else if (file.size() > 4 && file.substr(file.size() - 4) == ".cbf") {
        printf("Error: CBF input is not supported by this build\n");
        return 1;
    } else {
        wait_for_ready(file, H5Read::is_ready_for_read, args.timeout);
        try {
            reader_ptr = std::make_unique<H5Read>(file);
        } catch (const std::exception& e) {
            printf("%s\n", e.what());
            return 1;
        }
    }
    Reader& reader = *reader_ptr;
    const size_t bytes_per_pixel = reader.get_element_size();
    if (bytes_per_pixel != sizeof(pixel_t)) {  // spotfinder.cc:468-476: exit code = bit depth
        printf("Error: Data type mismatch; This executable only accepts %zu bit != %zu\n", sizeof(pixel_t) * 8,
               bytes_per_pixel * 8);
        return (int)(bytes_per_pixel * 8);
    }
    const uint32_t num_images = args.images ? *args.images : (uint32_t)reader.get_number_of_images();
    const uint32_t height = (uint32_t)reader.image_shape()[0], width = (uint32_t)reader.image_shape()[1];

    // ---- detector geometry / wavelength (spotfinder.cc:486-579) -----------------------------
    DetectorGeometry det;
    if (args.detector) {
        try {
            const jsonmini::Value j = jsonmini::Parser::parse(*args.detector);
            const float psx_mm = (float)j.at("pixel_size_x").number(), psy_mm = (float)j.at("pixel_size_y").number();
            det.pixel_size_x = psx_mm / 1000.f;
            det.pixel_size_y = psy_mm / 1000.f;
            det.beam_center_x = (float)j.at("beam_center_x").number() / psx_mm;
            det.beam_center_y = (float)j.at("beam_center_y").number() / psy_mm;
            det.distance = (float)j.at("distance").number() / 1000.f;
        } catch (const std::exception& e) {
            printf("Error: could not parse --detector: %s\n", e.what());
            return 1;
        }
    } else {
        const auto bc = reader.get_beam_center();
        const auto ps = reader.get_pixel_size();
        const auto dist = reader.get_detector_distance();
        if (!bc) { printf("Error: No beam center available from file. Please pass detector metadata with --distance.\n"); return 1; }
        if (!ps) { printf("Error: No pixel size available from file. Please pass detector metadata with --distance.\n"); return 1; }
        if (!dist) { printf("Error: No detector distance available from file. Please pass metadata with --distance.\n"); return 1; }
        det = {(*ps)[1], (*ps)[0], (*bc)[1], (*bc)[0], *dist};
    }
    float wavelength;
    if (args.wavelength) {
        wavelength = *args.wavelength;
    } else {
        const auto w = reader.get_wavelength();
        if (!w) {
            printf("Error: No wavelength provided. Please pass wavelength using: --wavelength\n");
            return 1;
        }
        wavelength = *w;
        printf("Got wavelength from file: %f \xC3\x85\n", wavelength);
    }
    printf("Detector geometry:\n    Distance:    %.1f mm\n    Beam Center: %.1f px %.1f px\nBeam Wavelength: %.2f \xC3\x85\n",
           det.distance * 1000, det.beam_center_x, det.beam_center_y, wavelength);
    const auto [oscillation_start, oscillation_width] = reader.get_oscillation();
    const bool rotation = oscillation_width > 0;
    if (rotation) printf("Oscillation:  Start: %.2f\xC2\xB0  Width: %.2f\xC2\xB0\n", oscillation_start, oscillation_width);
    std::signal(SIGINT, stop_processing);
#include "spotfinder_main_loop.inl"
}
